"""Import the UNMODIFIED reference from /root/reference (build container only).

TEST INFRASTRUCTURE ONLY.  /root/reference does not exist on the GPU box, so
this module is used solely by ``oracle/gen_golden.py`` and by ``-m "not gpu"``
tests that skip when the tree is absent.  No reference file is edited:
  * ``matplotlib`` (absent here; imported at models/modules/quantize.py:6 for a debug plot) is
    stubbed in ``sys.modules``;
  * ``UniformQuantizeGrad.forward`` is wrapped at run time to return ``input.clone()`` - the
    reference returns its input as-is (quantize.py:121), which current torch rejects as soon
    as the model applies an in-place ReLU.  Numerically neutral.
"""
import contextlib
import os
import sys
import types

REF_ROOT = "/root/reference"


def available():
    return os.path.isfile(os.path.join(REF_ROOT, "models", "modules", "quantize.py"))


def load():
    """Returns (models_pkg, quantize_module) of the reference."""
    if not available():
        raise RuntimeError("reference tree not present")
    sys.dont_write_bytecode = True
    for name in ("matplotlib", "matplotlib.pyplot"):
        sys.modules.setdefault(name, types.ModuleType(name))
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    import models as ref_models
    from models.modules import quantize as rq
    if not getattr(rq, "_clone_shim", False):
        def fwd(cls, ctx, input, num_bits=8, min_value=None, max_value=None, stochastic=True, inplace=False):
            ctx.inplace, ctx.num_bits = inplace, num_bits
            ctx.min_value, ctx.max_value, ctx.stochastic = min_value, max_value, stochastic
            return input.clone()
        rq.UniformQuantizeGrad.forward = classmethod(fwd)
        rq._clone_shim = True
    return ref_models, rq


@contextlib.contextmanager
def record_uniform(log):
    """Append every tensor produced by ``Tensor.uniform_`` (the stochastic-rounding noise,
    quantize.py:93) to ``log`` in call order."""
    import torch
    orig = torch.Tensor.uniform_

    def wrapped(self, *a, **k):
        out = orig(self, *a, **k)
        log.append(out.detach().clone())
        return out

    torch.Tensor.uniform_ = wrapped
    try:
        yield log
    finally:
        torch.Tensor.uniform_ = orig
