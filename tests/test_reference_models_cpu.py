"""north_star: "the same QConv2d/QLinear/QuantMeasure/RangeBN/quantize() Python API stays in place, so resnet_quantized and
mobilenet_quantized run unchanged".

The reference's OWN model files (models/resnet_quantized.py, models/mobilenet_quantized.py - unmodified, read from
/root/reference, which exists in the build container only) are imported with THIS package standing in for
``models/modules/quantize.py`` (the one-line substitution INTEGRATION.md section 1 describes), constructed, and compared
with the repo's builders (quantized.pytorch_b200/models.py, which the GPU box uses because the reference tree does not
travel) and with the reference models on the reference's own modules: same module classes, constructor arguments,
state-dict keys / shapes / dtypes, initialisation statistics and training regime.  Forward / backward of these networks on
the GPU is covered by tests/test_models_gpu.py.
"""
import importlib
import os
import sys
import types

import pytest
import torch

from oracle import ref_import

pytestmark = pytest.mark.skipif(not ref_import.available(), reason="reference tree not present (GPU box)")


def _reference_models_on(qmodule, pkg_name):
    """Import the reference's model files as package ``pkg_name`` whose ``.modules.quantize`` is ``qmodule``."""
    root = os.path.join(ref_import.REF_ROOT, "models")
    sys.dont_write_bytecode = True
    for name in ("matplotlib", "matplotlib.pyplot"):
        sys.modules.setdefault(name, types.ModuleType(name))
    pkg = types.ModuleType(pkg_name)
    pkg.__path__ = [root]
    sub = types.ModuleType(pkg_name + ".modules")
    sub.__path__ = [os.path.join(root, "modules")]
    sys.modules[pkg_name] = pkg
    sys.modules[pkg_name + ".modules"] = sub
    sys.modules[pkg_name + ".modules.quantize"] = qmodule          # <- the substitution
    rq = importlib.import_module(pkg_name + ".resnet_quantized")
    mq = importlib.import_module(pkg_name + ".mobilenet_quantized")
    return rq, mq


@pytest.fixture(scope="module")
def nets():
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import quantize as Q
    ours_r, ours_m = _reference_models_on(Q, "refmodels_on_qb200")
    _, ref_q = ref_import.load()
    ref_r, ref_m = _reference_models_on(ref_q, "refmodels_on_ref")
    return Q, ref_q, ours_r, ours_m, ref_r, ref_m


CONFIGS = [("resnet18", dict(depth=18, dataset="imagenet")), ("resnet20", dict(depth=20, dataset="cifar10")),
           ("resnet50", dict(depth=50, dataset="imagenet")), ("mobilenet", dict())]


def _build(rmod, mmod, name, cfg):
    torch.manual_seed(0)
    if name == "mobilenet":
        return mmod.mobilenet_quantized(**cfg)
    return rmod.resnet_quantized(**cfg)


@pytest.mark.parametrize("name,cfg", CONFIGS)
def test_reference_model_files_run_on_this_package(nets, name, cfg):
    Q, ref_q, ours_r, ours_m, ref_r, ref_m = nets
    from qtorch_b200 import models
    net = _build(ours_r, ours_m, name, cfg)               # the reference's file, our modules
    ref = _build(ref_r, ref_m, name, cfg)                 # the reference's file, the reference's modules
    mine = models.mobilenet_quantized(Q) if name == "mobilenet" else models.resnet_quantized(Q, **cfg)   # the repo's builder
    # every quantized layer of the reference's network is THIS package's class
    kinds = {type(m) for m in net.modules()}
    assert Q.QConv2d in kinds and Q.QLinear in kinds and Q.RangeBN in kinds and Q.QuantMeasure in kinds
    assert not any(type(m).__module__.endswith("refmodels_on_ref.modules.quantize") for m in net.modules())
    sd, sr, sm = net.state_dict(), ref.state_dict(), mine.state_dict()
    assert list(sd) == list(sr), "state-dict keys / order differ from the reference's"
    assert sorted(sd) == sorted(sm), "the repo's builder produces different keys"
    for k in sd:
        assert sd[k].shape == sr[k].shape == sm[k].shape and sd[k].dtype == sr[k].dtype == sm[k].dtype, k
    if name == "resnet18":
        assert len(sd) == 351
    # same seed -> the reference's init code ran identically over our modules (same parameter order and shapes)
    for k in sd:
        if sd[k].dtype.is_floating_point and sd[k].numel() > 1 and "running" not in k:
            assert torch.equal(sd[k], sr[k]), k
    # checkpoints interchange both ways
    net.load_state_dict(sr)
    ref.load_state_dict(net.state_dict())
    mine.load_state_dict(sr)
    # constructor arguments seen by the quantized layers
    for (n1, m1), (n2, m2) in zip(net.named_modules(), ref.named_modules()):
        assert n1 == n2
        for attr in ("num_bits", "num_bits_weight", "num_bits_grad", "biprecision", "stride", "padding", "groups", "kernel_size",
                     "in_channels", "out_channels", "in_features", "out_features", "momentum", "num_chunks", "eps"):
            if hasattr(m2, attr) and not isinstance(getattr(m2, attr), torch.Tensor):
                assert getattr(m1, attr) == getattr(m2, attr), (n1, attr)
    assert net.regime == ref.regime
    if name != "mobilenet":
        assert mine.regime == ref.regime, "models.py regime differs from resnet_quantized.py"


def test_module_api_surface(nets):
    """Every name the reference's consumers import from models/modules/quantize.py exists here with the same signature."""
    import inspect
    Q, ref_q = nets[0], nets[1]
    for name in ("quantize", "quantize_grad", "QConv2d", "QLinear", "RangeBN", "QuantMeasure", "set_measure_mode",
                 "set_global_quantization_method", "UniformQuantize", "UniformQuantizeGrad", "conv2d_biprec", "linear_biprec",
                 "QuantNode", "set_quant_mode", "overwrite_params", "freeze_quant_params", "set_bn_is_train", "is_bn"):
        assert hasattr(Q, name), name
    for name in ("quantize", "quantize_grad", "conv2d_biprec", "linear_biprec"):
        a, b = inspect.signature(getattr(Q, name)), inspect.signature(getattr(ref_q, name))
        assert list(a.parameters) == list(b.parameters), name
        assert [p.default for p in a.parameters.values()] == [p.default for p in b.parameters.values()], name
    for cls in ("QConv2d", "QLinear", "RangeBN", "QuantMeasure"):
        a = inspect.signature(getattr(Q, cls).__init__)
        b = inspect.signature(getattr(ref_q, cls).__init__)
        assert list(a.parameters) == list(b.parameters), cls
        assert [p.default for p in a.parameters.values()] == [p.default for p in b.parameters.values()], cls


def test_oracle_stochastic_min_anchored_branch_matches_reference(nets):
    """quantize(x, num_chunks=<truthy>) of the reference = min-anchored branch WITH noise (quantize.py:92-94, 159-160); the
    oracle restatement (qref.fq_affine(noise=...)) reproduces it bit for bit with the recorded uniform_ noise."""
    from oracle import qref
    ref_q = nets[1]
    torch.manual_seed(11)
    x = torch.randn(4, 8, 16, 16)
    lo, hi = float(x.min()), float(x.max())
    log = []
    with ref_import.record_uniform(log):
        want = ref_q.quantize(x, 8, lo, hi, num_chunks=4)
    assert len(log) == 1
    got = qref.fq_affine(x, 8, lo, hi, noise=log[0])
    assert torch.equal(got, want)
    assert not torch.equal(want, ref_q.quantize(x, 8, lo, hi))
