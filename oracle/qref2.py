"""CPU oracle for the quantize2.py flavour.  TEST INFRASTRUCTURE ONLY: imported by tests/,
never by the product path.  Pinned bit-exactly against tests/golden/q2_*.npz, which oracle/gen_golden2.py produced by
running the UNMODIFIED /root/reference/models/modules/quantize2.py.  Each function cites the reference lines it restates.

Functional restatement: ranges are (lo, hi) tensors, the quantizer is one expression, modules are plain functions over
explicit state dicts - no autograd Functions, no nn.Modules."""
import math

import torch


def qparams(x, flatten_dims=(1, -1), reduce_dim=0, reduce_type="mean", keepdim=False):
    """calculate_qparams, quantize2.py:19-38 -> (zero_point, range) tensors (no gradient)."""
    x = x.detach()
    flat = x.flatten(*flatten_dims)
    if flat.dim() == 1:                                        # :22-24
        lo, hi = flat.min(), flat.max()
    else:                                                      # :25-27
        lo, hi = flat.min(-1)[0], flat.max(-1)[0]
    pad = [1] * (x.dim() - lo.dim())                           # _deflatten_as, :14-16
    lo, hi = lo.reshape(list(lo.shape) + pad), hi.reshape(list(hi.shape) + pad)
    if reduce_dim is not None:
        if reduce_type == "mean":                              # :29-31
            lo, hi = lo.mean(reduce_dim, keepdim=keepdim), hi.mean(reduce_dim, keepdim=keepdim)
        else:                                                  # :32-34 ('extreme')
            lo, hi = lo.min(reduce_dim, keepdim=keepdim)[0], hi.max(reduce_dim, keepdim=keepdim)[0]
    return lo, hi - lo                                         # :36-38


def fq2(x, zero_point, rng, bits, signed=False, dequantize=True, noise=None):
    """UniformQuantize.forward, quantize2.py:58-74.  zero_point / rng broadcast against x; noise = the uniform(-.5,.5)
    tensor the stochastic branch draws (:66-67) or None."""
    qmin = -(2.0 ** (bits - 1)) if signed else 0.0             # :60
    qmax = qmin + 2.0 ** bits - 1.0                            # :61
    scale = rng / (qmax - qmin)                                # :62
    t = (x.detach() + (qmin * scale - zero_point)) / scale     # :64
    if noise is not None:
        t = t + noise                                          # :65-67
    q = t.clamp(qmin, qmax).round()                            # :69
    if not dequantize:
        return q
    return q * scale + (zero_point - qmin * scale)             # :71-73


def quantize(x, bits=None, qp=None, flatten_dims=(1, -1), reduce_dim=0, dequantize=True, signed=False, noise=None):
    """quantize(), quantize2.py:131-132 (+ the qparams-from-data default of :53-56)."""
    if qp is None:
        zp, rng = qparams(x, flatten_dims, reduce_dim)
    else:
        zp, rng = qp
    return fq2(x, zp, rng, bits, signed, dequantize, noise)


def quantize_grad_backward(g, bits, flatten_dims=(0, -1), reduce_dim=0, signed=False, noise=None):
    """UniformQuantizeGrad.backward, quantize2.py:99-112: 'extreme' ranges of the gradient, stochastic quantizer."""
    zp, rng = qparams(g, flatten_dims, reduce_dim, "extreme")
    return fq2(g, zp, rng, bits, signed, True, noise)


def measure_update(state, x, bits, flatten_dims=(1, -1), momentum=0.1, measure=False):
    """QuantMeasure.forward statistics branch, quantize2.py:178-191.  state: dict(running_zero_point, running_range,
    num_measured), updated in place.  Returns the (zero_point, range) this call quantizes with."""
    zp, rng = qparams(x, flatten_dims, 0)
    if measure:                                                # :183-185
        m = state["num_measured"] / (state["num_measured"] + 1)
        state["num_measured"] += 1
    else:
        m = momentum
    state["running_zero_point"].mul_(m).add_(zp * (1 - m))     # :188-189
    state["running_range"].mul_(m).add_(rng * (1 - m))         # :190-191
    return zp, rng


def rangebn2_forward(x, state, gamma, beta, training, momentum=0.1, num_chunks=16, eps=1e-5):
    """quantize2.RangeBN.forward after the input quantizer, :305-342.  x [B,C,H,W] (already fake-quantized); state:
    dict(running_mean, running_var).  Differentiable w.r.t. x, gamma, beta (autograd supplies the reference backward)."""
    B, C, H, W = x.shape
    if training:
        y = x.transpose(0, 1).contiguous().view(C, num_chunks, (B * H * W) // num_chunks)     # :309-310
        mean_max = y.max(-1)[0].mean(-1)                                                       # :311
        mean_min = y.min(-1)[0].mean(-1)                                                       # :312
        mean = y.view(C, -1).mean(-1)                                                          # :313
        fix = (0.5 * 0.35) * (1 + (math.pi * math.log(4)) ** 0.5) / ((2 * math.log(y.size(-1))) ** 0.5)   # :314-315
        scale = (mean_max - mean_min) * fix                                                    # :317
        with torch.no_grad():
            state["running_mean"].mul_(momentum).add_(mean * (1 - momentum))                   # :319-320
            state["running_var"].mul_(momentum).add_(scale * (1 - momentum))                   # :322-323
    else:
        mean, scale = state["running_mean"], state["running_var"]                             # :325-326
    out = (x - mean.view(1, -1, 1, 1)) / (scale.view(1, -1, 1, 1) + eps)                       # :329-330
    if gamma is not None:
        out = out * gamma.view(1, -1, 1, 1)                                                    # :336
    if beta is not None:
        out = out + beta.view(1, -1, 1, 1)                                                     # :341
    return out


# ---------------------------------------------------------------------------------------------------
# autograd glue for module-level checks: straight-through quantizers and the gradient quantizer
# ---------------------------------------------------------------------------------------------------
def ste(x, y):
    """Value of y, gradient of identity w.r.t. x (UniformQuantize.backward, quantize2.py:78-82)."""
    return x + (y - x).detach()


class GradQ(torch.autograd.Function):
    """quantize_grad (quantize2.py:85-112): identity forward; backward quantizes with the next recorded noise tensor."""

    @staticmethod
    def forward(ctx, x, bits, flatten_dims, noise_src):
        ctx.bits, ctx.fd, ctx.noise_src = bits, flatten_dims, noise_src
        return x.clone()

    @staticmethod
    def backward(ctx, g):
        noise = ctx.noise_src.pop(0) if ctx.noise_src is not None else None
        return quantize_grad_backward(g, ctx.bits, ctx.fd, 0, False, noise), None, None, None


def qcontract(x, w, b, state, conv, bits=8, bits_w=8, bits_g=8, biprec=True, training=True, noise_src=None):
    """QConv2d.forward / QLinear.forward, quantize2.py:217-235 / 249-269.  conv = (stride, padding, dilation, groups) or
    None for linear; state = the input QuantMeasure's buffers."""
    import torch.nn.functional as F
    if training:
        zp, rng = measure_update(state, x, bits, (1, -1), 0.1)
    else:
        zp, rng = state["running_zero_point"], state["running_range"]
    qx = ste(x, fq2(x, zp, rng, bits))
    wzp, wrng = qparams(w, (1, -1), None)
    qw = ste(w, fq2(w, wzp, wrng, bits_w))
    qb = None
    if b is not None:
        qb = ste(b, quantize(b, bits_w + bits, flatten_dims=(0, -1)))
    op = (lambda a, ww, bb: F.conv2d(a, ww, bb, *conv)) if conv is not None else F.linear
    fd = (1, -1) if conv is not None else (0, -1)
    if not biprec or bits_g is None:                                         # :225-230
        out = op(qx, qw, qb)
        return GradQ.apply(out, bits_g, fd, noise_src) if bits_g is not None else out
    out1 = op(qx.detach(), qw, qb)                                           # conv2d_biprec, :115-121
    out2 = GradQ.apply(op(qx, qw.detach(), qb.detach() if qb is not None else None), bits_g, fd, noise_src)
    return out1 + out2 - out1.detach()
