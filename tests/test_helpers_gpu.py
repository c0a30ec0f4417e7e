"""SURVEY.md section 8(f) N2 - calibration / deployment helpers on the GPU against goldens generated from the UNMODIFIED
reference (oracle/gen_golden_n2.py; models/modules/quantize.py:217, 301-312, 386-396, 547-592): measure mode ->
calibrate -> quantized eval with the running ranges, set_quant_mode, freeze_quant_params, overwrite_params, and
QuantMeasure(momentum=0)."""
import numpy as np
import pytest
import torch

from conftest import golden
from n2_net import N2Net

pytestmark = pytest.mark.gpu
DEV = "cuda"


def rel_l2(a, b):
    a = a.detach().cpu().double().flatten()
    b = torch.from_numpy(np.asarray(b)).double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def _state(g, prefix):
    return {k[len(prefix):]: torch.from_numpy(g[k]) for k in g.files if k.startswith(prefix)}


def _check_buffers(net, g, prefix, exact=()):
    want = _state(g, prefix)
    have = net.state_dict()
    assert sorted(want) == sorted(have)
    for k, v in want.items():
        h = have[k].detach().cpu()
        assert h.shape == v.shape, k
        if k in exact or not v.dtype.is_floating_point:
            assert torch.equal(h, v), (prefix, k)
        else:
            # ranges: the batch means are accumulated in fp64 here and in fp32 by ATen (<= 1 ulp);
            # RangeBN running statistics inherit a few flipped codes of the calibrated activations
            tol = 2e-3 if ("bn1.running" in k) else 1e-6
            np.testing.assert_allclose(h.numpy(), v.numpy(), rtol=tol, atol=tol * float(v.abs().max()) + 1e-12, err_msg=prefix + k)


def test_measure_mode_calibration_and_eval():
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import _abi, quantize as Q
    g = golden("n2_helpers.npz")
    net = N2Net(Q)
    net.load_state_dict(_state(g, "init/"))
    net = net.to(DEV).train()
    xs = torch.from_numpy(g["xs"]).to(DEV)
    xe = torch.from_numpy(g["xe"]).to(DEV)
    # 1. calibration in measure mode: quantizers pass through, QuantMeasure buffers accumulate (cumulative average)
    Q.set_measure_mode(net, True)
    assert not net.conv1.enable_quant and not net.conv1.quantize_input.enable_quant and not net.bn1.quantize_input.enable_quant
    with torch.no_grad():
        for i in range(3):
            out = net(xs[i])
            assert rel_l2(out, g["measure_out%d" % i]) < 1e-4, i
    _check_buffers(net, g, "calibrated/")
    # 2. quantized evaluation with the calibrated (running) ranges; the contractions run on the int8 tensor-core kernels
    Q.set_measure_mode(net, False)
    net.eval()
    with torch.no_grad():
        _abi.profile_begin()
        out = net(xe)
        names = [r[0] for r in _abi.profile_end()]
        assert "qb_conv_i8_fwd" in names and "qb_gemm_i8" in names, names
        assert "qb_measure_stats" not in names, "eval must use the running ranges"
        assert rel_l2(out, g["eval_out"]) < 2e-4
        # 3. set_quant_mode(False): the float network; back on: the same quantized output
        Q.set_quant_mode(net, False)
        assert rel_l2(net(xe), g["float_out"]) < 1e-4
        Q.set_quant_mode(net, True)
        out2 = net(xe)
        assert torch.equal(out, out2)
        assert rel_l2(out2, g["eval_out_again"]) < 2e-4
    # 4. freeze_quant_params: a train-mode forward leaves the activation ranges untouched (and, as in the reference,
    #    sets the misspelt attribute, so weight ranges keep following the weights)
    net.train()
    Q.freeze_quant_params(net)
    assert not net.conv1.quantize_input.training and net.conv1.training
    assert getattr(net.conv1, "freeze_param_dyn_rang") is True and net.conv1.freeze_param_dyn_range is False
    before = {k: v.clone() for k, v in net.state_dict().items() if "quantize_input" in k}
    with torch.no_grad():
        out = net(xs[0])
    assert rel_l2(out, g["frozen_train_out"]) < 2e-4
    for k, v in before.items():
        assert torch.equal(net.state_dict()[k], v), k
    _check_buffers(net, g, "frozen/")
    # 5. overwrite_params bakes the quantized weights (stored ranges) into the parameters: bit-exact given the same ranges
    w_before = net.conv1.weight.detach().clone()
    Q.overwrite_params(net)
    assert not torch.equal(net.conv1.weight.detach(), w_before)
    _check_buffers(net, g, "overwritten/", exact=("conv1.weight", "conv2.weight", "fc.weight", "conv1.bias", "fc.bias"))
    # idempotent: quantizing already-quantized weights with the same ranges changes nothing
    w1 = net.conv1.weight.detach().clone()
    Q.overwrite_params(net)
    assert torch.equal(net.conv1.weight.detach(), w1)


def test_quantmeasure_momentum_zero_is_cumulative():
    """quantize.py:217: `momentum or self.momentum or n/(n+1)` - a momentum of 0 is falsy and selects the cumulative average."""
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import quantize as Q
    g = golden("n2_helpers.npz")
    qm = Q.QuantMeasure(8, momentum=0).to(DEV).train()
    xs = torch.from_numpy(g["xs"]).to(DEV)
    with torch.no_grad():
        for i in range(3):
            out = qm(xs[i])
    for k in ("running_min", "running_max", "running_mean", "running_var", "num_measurements"):
        np.testing.assert_allclose(qm.state_dict()[k].cpu().numpy(), g["mom0/" + k], rtol=3e-6, err_msg=k)
    assert rel_l2(out, g["mom0_out"]) < 1e-4


def test_stochastic_min_anchored_and_group_ranges_are_reachable():
    """quantize(x, num_chunks=<truthy>) reaches `stochastic=True` on the min-anchored branch
    (quantize.py:92-94 with enforce_true_zero=False) - now computed, not raised: codes within one step of the
    deterministic ones, unbiased, and bit-exact against the formula with injected noise."""
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import quantize as Q
    from oracle import qref
    torch.manual_seed(3)
    x = torch.randn(4, 8, 16, 16)
    lo, hi = float(x.min()), float(x.max())
    noise = torch.rand(x.shape) - 0.5
    ref = qref.fq_affine(x, 8, lo, hi, noise=noise)               # pinned to the reference in tests/test_reference_models_cpu.py
    prev = Q.set_noise_provider(lambda shape: noise.to(DEV).reshape(shape))
    try:
        out = Q.quantize(x.to(DEV), 8, lo, hi, num_chunks=4)
    finally:
        Q.set_noise_provider(prev)
    assert torch.equal(out.cpu(), ref)
    det = Q.quantize(x.to(DEV), 8, lo, hi)
    step = (hi - lo) / 255.0
    out_p = Q.quantize(x.to(DEV), 8, lo, hi, num_chunks=4)        # in-kernel Philox noise
    assert float((out_p - det).abs().max()) <= step * 1.0001
    assert abs(float((out_p - x.to(DEV)).mean())) < 0.02 * step
