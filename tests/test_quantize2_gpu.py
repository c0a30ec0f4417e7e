"""quantize2.py flavour (row A9) on the GPU path against the goldens produced by the unmodified reference
(tests/golden/q2_*.npz) and the CPU oracle (oracle/qref2.py).  Elementwise results are bit-exact given the qparams; the
per-sample-mean ranges agree to <= 1 ulp with ATen's CPU mean (a 1-ulp range moves a handful of codes by one step)."""
import numpy as np
import pytest
import torch

from conftest import golden
from oracle import qref2

pytestmark = pytest.mark.gpu
DEV = "cuda"


def T(a):
    return torch.from_numpy(np.asarray(a))


def q2():
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import quantize2
    return quantize2


def bits_equal(a, b, what=""):
    a, b = a.detach().cpu().numpy(), np.asarray(b)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    assert np.array_equal(a.view(np.int32), b.view(np.int32)), "%s: max abs diff %g" % (what, np.abs(a - b).max())


def close_codes(a, b, step, frac=2e-3, what=""):
    """equal up to one quantization step on at most `frac` of the elements (1-ulp range differences)."""
    a, b = a.detach().cpu().double().numpy(), np.asarray(b, dtype=np.float64)
    d = np.abs(a - b)
    assert d.max() <= step * 1.001 + 1e-6 * np.abs(b).max(), (what, d.max(), step)
    assert (d > 1e-6 * (np.abs(b).max() + 1e-30)).mean() <= frac, (what, (d > 0).mean())


class Replay:
    def __init__(self, tensors):
        self.t = [x.to(DEV) for x in tensors]

    def __call__(self, shape):
        return self.t.pop(0).reshape(shape).contiguous()


def test_fq2_given_qparams_bit_exact():
    Q2 = q2()
    g = golden("q2_fq.npz")
    for i in range(int(g["count"])):
        x = T(g["x%d" % i]).to(DEV)
        bits, signed, deq = [int(v) for v in g["meta%d" % i]]
        # per-row qparams as the reference computed them (weights path): exact min/max, exact quantizer
        qp = Q2.calculate_qparams(x, bits, flatten_dims=(1, -1), reduce_dim=None)
        bits_equal(qp.zero_point, g["zp%d" % i])
        bits_equal(qp.range, g["rng%d" % i])
        bits_equal(Q2.quantize(x, qparams=qp, signed=bool(signed), dequantize=bool(deq)), g["yrow%d" % i], "row %d" % i)
        # global range (flatten everything): exact
        bits_equal(Q2.quantize(x, num_bits=bits, flatten_dims=(0, -1), signed=bool(signed), dequantize=bool(deq)),
                   g["yglob%d" % i], "glob %d" % i)
        # per-sample mean range: qparams from the CPU oracle -> bit exact; from the GPU mean -> within one step
        zp, rng = qref2.qparams(T(g["x%d" % i]))
        qp_ref = Q2.QParams(range=rng.to(DEV), zero_point=zp.to(DEV), num_bits=bits)
        bits_equal(Q2.quantize(x, qparams=qp_ref, signed=bool(signed), dequantize=bool(deq)), g["y%d" % i], "mean %d" % i)
        step = float(rng) / (2 ** bits - 1) if deq else 1.0
        close_codes(Q2.quantize(x, num_bits=bits, signed=bool(signed), dequantize=bool(deq)), g["y%d" % i], step, what="m%d" % i)


def test_fq2_stochastic_and_inplace():
    Q2 = q2()
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import quantize as Q
    g = golden("q2_fq.npz")
    x = T(g["sx"])
    for bits, signed, key, nk in ((8, False, "sy", "snoise0"), (6, True, "sy_signed", "snoise1")):
        zp, rng = qref2.qparams(x)
        qp = Q2.QParams(range=rng.to(DEV), zero_point=zp.to(DEV), num_bits=bits)
        prev = Q.set_noise_provider(Replay([T(g[nk])]))
        try:
            xd = x.to(DEV).clone()
            y = Q2.quantize(xd, qparams=qp, signed=signed, stochastic=True, inplace=True)
        finally:
            Q.set_noise_provider(prev)
        assert y.data_ptr() == xd.data_ptr()
        bits_equal(y, g[key], key)
    # Philox noise: unbiased, within one step
    xd = (torch.randn(1 << 20, generator=torch.Generator().manual_seed(1)) * 0.3).to(DEV)
    y = Q2.quantize(xd, num_bits=4, flatten_dims=(0, -1), stochastic=True)
    step = float(xd.max() - xd.min()) / 15
    err = (y - xd).double()
    assert float(err.abs().max()) <= step * 1.0001 and abs(float(err.mean())) < 4 * step / np.sqrt(12 * xd.numel())


def test_grad2():
    Q2 = q2()
    from qtorch_b200 import quantize as Q
    g = golden("q2_grad.npz")
    for i in range(int(g["count"])):
        fd = tuple(int(v) for v in g["fd%d" % i])
        x = torch.zeros(g["go%d" % i].shape, device=DEV, requires_grad=True)
        prev = Q.set_noise_provider(Replay([T(g["noise%d" % i])]))
        try:
            Q2.quantize_grad(x * 1.0, num_bits=8, flatten_dims=fd).backward(T(g["go%d" % i]).to(DEV))
        finally:
            Q.set_noise_provider(prev)
        bits_equal(x.grad, g["dx%d" % i], "grad %d" % i)


def test_measure2():
    Q2 = q2()
    g = golden("q2_measure.npz")
    for tag, kw in (("ema", dict(momentum=0.1)), ("meas", dict(measure=True))):
        m = Q2.QuantMeasure(8, shape_measure=(1, 1, 1, 1), flatten_dims=(1, -1), **kw).to(DEV).train()
        for s in range(3):
            x = T(g["%s_x%d" % (tag, s)]).to(DEV)
            y = m(x)
            assert torch.allclose(m.running_zero_point.cpu(), T(g["%s_zp%d" % (tag, s)]), rtol=3e-7, atol=0)
            assert torch.allclose(m.running_range.cpu(), T(g["%s_rng%d" % (tag, s)]), rtol=3e-7, atol=0)
            bits_equal(m.num_measured, g["%s_n%d" % (tag, s)])
            if tag == "meas":
                assert y is x
            else:
                close_codes(y, g["%s_y%d" % (tag, s)], float(g["%s_x%d" % (tag, s)].max() - g["%s_x%d" % (tag, s)].min()) / 255)
        m.measure = False
        m.eval()
        close_codes(m(T(g[tag + "_xe"]).to(DEV)), g[tag + "_ye"], float(m.running_range) / 255)


def rel_l2(a, b):
    a, b = a.detach().cpu().double(), T(b).double()
    return float((a - b).norm() / (b.norm() + 1e-30))


MODS = {"conv": (lambda Q2: Q2.QConv2d(8, 12, 3, padding=1)), "conv_s2_nobias": (lambda Q2: Q2.QConv2d(16, 8, 3, stride=2, padding=1, bias=False)),
        "conv_nobiprec": (lambda Q2: Q2.QConv2d(8, 8, 1, biprecision=False)), "linear": (lambda Q2: Q2.QLinear(32, 10)),
        "linear_nograd": (lambda Q2: Q2.QLinear(16, 4, num_bits_grad=None))}


@pytest.mark.parametrize("tag", sorted(MODS))
def test_modules2(tag):
    Q2 = q2()
    from qtorch_b200 import quantize as Q
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    g = golden("q2_modules.npz")
    m = MODS[tag](Q2)
    with torch.no_grad():
        m.weight.copy_(T(g[tag + "_w"]))
        if m.bias is not None:
            m.bias.copy_(T(g[tag + "_b"]))
    m = m.to(DEV).train()
    x = T(g[tag + "_x"]).to(DEV).requires_grad_()
    prev = Q.set_noise_provider(Replay([T(g["%s_noise%d" % (tag, k)]) for k in range(int(g[tag + "_nnoise"]))]))
    try:
        out = m(x)
        out.backward(T(g[tag + "_go"]).to(DEV))
    finally:
        Q.set_noise_provider(prev)
    assert rel_l2(out, g[tag + "_out"]) < 1e-4
    assert rel_l2(x.grad, g[tag + "_dx"]) < 2e-3         # a 1-ulp input range moves a few activation codes by one step
    assert rel_l2(m.weight.grad, g[tag + "_dw"]) < 2e-3
    if m.bias is not None:
        assert rel_l2(m.bias.grad, g[tag + "_db"]) < 1e-5
    assert torch.allclose(m.quantize_input.running_zero_point.cpu(), T(g[tag + "_zp"]), rtol=3e-7, atol=1e-9)
    assert torch.allclose(m.quantize_input.running_range.cpu(), T(g[tag + "_rng"]), rtol=3e-7, atol=0)


@pytest.mark.parametrize("tag", ["tc_conv", "tc_conv_s2_nobias", "tc_conv_1x1_nobiprec", "tc_conv_64", "tc_linear", "tc_linear_nograd"])
def test_modules2_on_tensor_cores(tag):
    """quantize2.QConv2d / QLinear (quantize2.py:204-272) on the SAME native contractions as quantize.py's
    modules - int8 tcgen05 forward on the codes of this flavour's quantizer, qb_conv_dgrad / qb_conv_wgrad backward with its
    per-sample 'extreme' gradient quantizer - against goldens from the unmodified reference (oracle/gen_golden2.py)."""
    Q2 = q2()
    from oracle.gen_golden2 import TC_CASES, build_tc
    from qtorch_b200 import _abi, quantize as Q
    g = golden("q2_modules_tc.npz")
    kind, cfg, _ = TC_CASES[tag]
    m = build_tc(Q2, tag)
    with torch.no_grad():
        m.weight.copy_(T(g[tag + "_w"]))
        if m.bias is not None:
            m.bias.copy_(T(g[tag + "_b"]))
    m = m.to(DEV).train()
    x = T(g[tag + "_x"]).to(DEV).requires_grad_()
    prev = Q.set_noise_provider(Replay([T(g["%s_noise%d" % (tag, k)]) for k in range(int(g[tag + "_nnoise"]))]))
    try:
        _abi.profile_begin()
        out = m(x)
        out.backward(T(g[tag + "_go"]).to(DEV))
        names = [r[0] for r in _abi.profile_end()]
    finally:
        Q.set_noise_provider(prev)
    assert ("qb_conv_i8_fwd" if kind == "conv" else "qb_gemm_i8") in names, names
    assert "qb_conv_dgrad" in names and "qb_conv_wgrad" in names and "qb_fakequant2_codes" in names, names
    assert rel_l2(out, g[tag + "_out"]) < 1e-4
    assert rel_l2(x.grad, g[tag + "_dx"]) < 2e-3         # a 1-ulp input range moves a few activation codes by one step
    assert rel_l2(m.weight.grad, g[tag + "_dw"]) < 2e-3
    if m.bias is not None:
        assert rel_l2(m.bias.grad, g[tag + "_db"]) < 1e-5
    assert torch.allclose(m.quantize_input.running_zero_point.cpu(), T(g[tag + "_zp"]), rtol=3e-7, atol=1e-9)
    assert torch.allclose(m.quantize_input.running_range.cpu(), T(g[tag + "_rng"]), rtol=3e-7, atol=0)


@pytest.mark.parametrize("tag,chunks", [("4d", 4), ("2d", 8)])
@pytest.mark.parametrize("nhwc", [False, True])
def test_rangebn2(tag, chunks, nhwc):
    Q2 = q2()
    from qtorch_b200 import quantize as Q
    g = golden("q2_rangebn.npz")
    C = g[tag + "_x"].shape[1]
    m = (Q2.RangeBN if tag == "4d" else Q2.RangeBN1d)(C, num_chunks=chunks)
    with torch.no_grad():
        m.weight.copy_(T(g[tag + "_gamma"]))
        m.bias.copy_(T(g[tag + "_beta"]))
    m = m.to(DEV).train()
    x = T(g[tag + "_x"]).to(DEV).requires_grad_()
    if nhwc and tag != "4d":
        pytest.skip("channels-last needs a 4-D input")
    xin = x * 1.0
    if nhwc:
        xin = xin.contiguous(memory_format=torch.channels_last)
    prev = Q.set_noise_provider(Replay([T(g[tag + "_noise"])]))
    try:
        out = m(xin)
        out.backward(T(g[tag + "_go"]).to(DEV))
    finally:
        Q.set_noise_provider(prev)
    assert rel_l2(out, g[tag + "_out"]) < 2e-3
    assert rel_l2(x.grad, g[tag + "_dx"]) < 5e-3
    assert rel_l2(m.weight.grad, g[tag + "_dgamma"]) < 5e-3
    assert rel_l2(m.bias.grad, g[tag + "_dbeta"]) < 1e-4
    assert torch.allclose(m.running_mean.cpu(), T(g[tag + "_rm"]), rtol=1e-4, atol=1e-6)
    assert torch.allclose(m.running_var.cpu(), T(g[tag + "_rv"]), rtol=1e-4, atol=1e-7)
    m.eval()
    oute = m(T(g[tag + "_xe"]).to(DEV))
    assert rel_l2(oute, g[tag + "_oute"]) < 2e-3


def test_rangebn2_exact_given_same_quantized_input():
    """Statistics + division-normalisation against the oracle on an input that is already on the 8-bit grid (so the
    1-ulp range caveat does not apply): bit-exact forward, tight backward."""
    Q2 = q2()
    gen = torch.Generator().manual_seed(3)
    x = torch.randint(0, 256, (16, 8, 6, 6), generator=gen).float() * 0.0123 - 1.0
    gamma, beta = torch.rand(8, generator=gen) + 0.5, torch.randn(8, generator=gen) * 0.1
    st = dict(running_mean=torch.zeros(8), running_var=torch.zeros(8))
    xr = x.clone().requires_grad_()
    gr, br = gamma.clone().requires_grad_(), beta.clone().requires_grad_()
    ref = qref2.rangebn2_forward(xr, st, gr, br, True, 0.1, 16)
    go = torch.randn(ref.shape, generator=gen)
    ref.backward(go)
    from qtorch_b200 import quantize2 as Q2m
    for nhwc in (False, True):
        xd = x.to(DEV).requires_grad_()
        xin = xd.contiguous(memory_format=torch.channels_last) if nhwc else xd
        gd, bd = gamma.to(DEV).requires_grad_(), beta.to(DEV).requires_grad_()
        rm, rv = torch.zeros(8, device=DEV), torch.zeros(8, device=DEV)
        out = Q2m._RangeBN2Fn.apply(xin, gd, bd, rm, rv, True, 0.1, 16, 1e-5, None)
        out.backward(go.to(DEV))
        assert torch.allclose(out.cpu(), ref, rtol=2e-6, atol=1e-6)
        assert torch.allclose(rm.cpu(), st["running_mean"], rtol=1e-6, atol=1e-7)
        assert torch.allclose(rv.cpu(), st["running_var"], rtol=1e-6, atol=1e-8)
        assert rel_l2(xd.grad, xr.grad) < 1e-5 and rel_l2(gd.grad, gr.grad) < 1e-5 and rel_l2(bd.grad, br.grad) < 1e-5
