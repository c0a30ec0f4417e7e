"""CPU restatement (torch-CPU, fp32) of the reference's fake-quantization path.

TEST INFRASTRUCTURE ONLY - see ``oracle/__init__.py``.  Every function cites the
reference lines (relative to /root/reference) it restates.  The arithmetic of
the reference is ATen-on-CPU, so the oracle is expressed in torch CPU ops too:
true fp32 division, round-half-to-even, separately rounded mul/add, Python
double scalars for ``float()`` ranges.

Pinned by ``tests/test_oracle_golden.py`` against ``tests/golden/*.npz`` that
``oracle/gen_golden.py`` produced by running the unmodified reference.
"""
import math

import torch
import torch.nn.functional as F

# ----------------------------------------------------------------------------
# noise source for the stochastic branch.  ``None`` -> draw from torch's default
# generator exactly like models/modules/quantize.py:93 does.  Tests install a
# callable(shape) -> tensor to replay a recorded sequence.
# ----------------------------------------------------------------------------
_noise_provider = None


def set_noise_provider(fn):
    global _noise_provider
    prev = _noise_provider
    _noise_provider = fn
    return prev


def _draw_noise(like):
    if _noise_provider is not None:
        n = _noise_provider(tuple(like.shape))
        return n.to(like.dtype).reshape(like.shape)
    return torch.empty_like(like).uniform_(-0.5, 0.5)  # quantize.py:93


def levels(bits):
    """qmax - qmin = 2**bits - 1 (quantize.py:68-69)."""
    return 2.0 ** bits - 1.0


# ----------------------------------------------------------------------------
# A1: min-anchored affine fake-quant (what every quantize() call reaches,
# quantize.py:159-160 argument shuffle -> enforce_true_zero=False).
# ----------------------------------------------------------------------------
def data_range(x):
    """quantize.py:45-55 with num_chunks=None -> x.view(1,-1): 1-element tensors
    holding the global min / max."""
    flat = x.reshape(1, -1)
    return flat.min(-1)[0].mean(-1), flat.max(-1)[0].mean(-1)


def affine_scale(lo, hi, bits):
    """quantize.py:71-75.  Python floats -> double arithmetic, max(.,1e-8);
    tensors -> fp32 arithmetic, clamp_(1e-8)."""
    s = (hi - lo) / levels(bits)
    if torch.is_tensor(s):
        return s.clamp(min=1e-8)
    return max(s, 1e-8)


def fq_affine(x, bits=8, lo=None, hi=None, want_codes=False, noise=None):
    """quantize.py:42-75, 90, 95, 100 (deterministic, min-anchored).

    lo/hi: Python floats (activation ranges), broadcastable tensors (per-channel
    weights, 0-dim bias ranges) or None (range taken from the data).
    Returns y (and the integer codes held in fp32 when want_codes)."""
    if lo is None or hi is None:
        dlo, dhi = data_range(x)
        lo = dlo if lo is None else lo
        hi = dhi if hi is None else hi
    qmax = levels(bits)
    s = affine_scale(lo, hi, bits)
    t = x.clone()
    t.add_(-lo).div_(s).add_(0.0)          # :90
    if noise is not None:                  # :92-94 (stochastic := num_chunks through quantize()'s positional shuffle)
        t.add_(noise)
    t.clamp_(0.0, qmax).round_()           # :95
    codes = t.clone() if want_codes else None
    t.add_(-0.0).mul_(s).add_(lo)          # :100
    return (t, codes) if want_codes else t


# ----------------------------------------------------------------------------
# A2: zero-point + stochastic branch (only reached from the gradient path,
# quantize.py:137-138 -> enforce_true_zero=True, stochastic=True).
# ----------------------------------------------------------------------------
def zero_point(lo, scale, bits):
    """quantize.py:78-87: int() truncation of the clamped double."""
    z = 0.0 - lo / scale
    z = min(max(z, 0.0), levels(bits))
    return int(z)


def fq_zero_point(g, bits, lo, hi, noise=None, want_codes=False):
    """quantize.py:66-88, 92-98 with Python-float lo/hi."""
    qmax = levels(bits)
    s = affine_scale(float(lo), float(hi), bits)
    zp = zero_point(float(lo), s, bits)
    t = g.clone()
    t.div_(s).add_(zp)                     # :88
    if noise is not None:
        t.add_(noise)                      # :93-94
    t.clamp_(0.0, qmax).round_()           # :95
    codes = t.clone() if want_codes else None
    t.add_(-zp).mul_(s)                    # :98
    return (t, codes) if want_codes else t


def fq_grad(g, bits=8, noise=None):
    """UniformQuantizeGrad.backward, quantize.py:124-139: global min/max of the
    incoming gradient as Python floats, then A2 with uniform(-.5,.5) noise."""
    lo, hi = float(g.min()), float(g.max())
    if noise is None:
        noise = _draw_noise(g)
    return fq_zero_point(g, bits, lo, hi, noise)


class _STE(torch.autograd.Function):
    """fq_affine forward, identity backward (quantize.py:105-109)."""

    @staticmethod
    def forward(ctx, x, bits, lo, hi):
        return fq_affine(x, bits, lo, hi)

    @staticmethod
    def backward(ctx, g):
        return g, None, None, None


class _GradQ(torch.autograd.Function):
    """quantize_grad: identity forward (quantize.py:115-121; a clone so that
    in-place ReLU / += stay legal on current torch), A2 backward."""

    @staticmethod
    def forward(ctx, x, bits):
        ctx.bits = bits
        return x.clone()

    @staticmethod
    def backward(ctx, g):
        return fq_grad(g, ctx.bits), None


def quantize(x, bits=8, lo=None, hi=None):
    """Effective behaviour of reference quantize() (quantize.py:159-160)."""
    return _STE.apply(x, bits, lo, hi)


def quantize_grad(x, bits=8):
    """reference quantize_grad (quantize.py:163-164)."""
    return _GradQ.apply(x, bits)


# ----------------------------------------------------------------------------
# A5: QuantMeasure
# ----------------------------------------------------------------------------
ACIQ_ALPHA = {2: 2.83, 3: 3.89, 4: 5.03, 5: 6.2, 6: 7.41, 7: 8.64, 8: 9.89}  # quantize.py:214


class RangeTracker:
    """State + update rule of QuantMeasure (quantize.py:198-268)."""

    def __init__(self, bits=8, momentum=None, method="avg"):
        self.running_min = torch.zeros(1)
        self.running_max = torch.zeros(1)
        self.num_measurements = torch.zeros(1)
        self.running_var = torch.ones(1)
        self.running_mean = torch.zeros(1)
        self.momentum = momentum
        self.bits = bits
        self.method = method
        self.enable_quant = True
        self.training = True

    def _blend(self, new, stat):
        # quantize.py:216-219
        m = self.momentum or self.num_measurements / (self.num_measurements + 1)
        stat.mul_(m).add_(new * (1 - m))

    def aciq_range(self, std, mean, tmin, tmax):
        # quantize.py:256-268 (std is modified in place there too)
        std += 1e-8
        clip = std * ACIQ_ALPHA[self.bits]
        clip = min((tmax - tmin) / 2, clip)
        return max(tmin, mean - clip), min(tmax, mean + clip)

    def observe(self, x):
        """Returns (lo, hi) as used for this call and updates the running
        buffers when training (quantize.py:221-246)."""
        x = x.detach()
        if self.training:
            rows = x.reshape(x.size(0), -1)
            lo = rows.min(-1)[0].mean()
            self._blend(lo, self.running_min)
            hi = rows.max(-1)[0].mean()
            self._blend(hi, self.running_max)
            mean = x.mean()
            std = x.std(unbiased=True)
            self._blend(mean, self.running_mean)
            self._blend(std, self.running_var)
            self.num_measurements += 1
            if self.method == "aciq":
                lo, hi = self.aciq_range(std, mean, lo, hi)
        elif self.method == "aciq":
            lo, hi = self.aciq_range(self.running_var, self.running_mean,
                                     self.running_min, self.running_max)
        else:
            lo, hi = self.running_min, self.running_max
        return float(lo), float(hi)

    def __call__(self, x):
        lo, hi = self.observe(x)
        if self.enable_quant:
            return quantize(x, self.bits, lo, hi)   # quantize.py:249
        return x


# ----------------------------------------------------------------------------
# A6/A7: QConv2d / QLinear forward with the bi-precision backward
# ----------------------------------------------------------------------------
def weight_row_range(w):
    """quantize.py:318-323: per-output-channel min/max, shaped [C,1,...]."""
    shape = (w.size(0),) + (1,) * (w.dim() - 1)
    flat = w.flatten(1)
    return flat.min(-1)[0].view(shape), flat.max(-1)[0].view(shape)


def biprec(op, x, w, b, bits_grad):
    """conv2d_biprec / linear_biprec, quantize.py:142-156: weight/bias gradients
    from the full-precision output gradient, input gradient from the quantized
    one; values equal op(x,w,b)."""
    out1 = op(x.detach(), w, b)
    out2 = op(x, w.detach(), b.detach() if b is not None else None)
    out2 = quantize_grad(out2, bits_grad)
    return out1 + out2 - out1.detach()


def qcontract(op, xq, weight, bias, bits_w, bits_grad, biprecision, bias_quant):
    """Shared body of QConv2d.forward / QLinear.forward, quantize.py:316-349 and
    :400-428, after the input has been fake-quantized."""
    wlo, whi = weight_row_range(weight)
    qw = quantize(weight, bits_w, wlo, whi)
    if bias is not None and bias_quant:
        qb = quantize(bias, bits_w, bias.min(), bias.max())
    else:
        qb = bias
    if not biprecision or bits_grad is None:
        out = op(xq, qw, qb)
        if bits_grad is not None:
            out = quantize_grad(out, bits_grad)
        return out
    return biprec(op, xq, qw, qb, bits_grad)


# ----------------------------------------------------------------------------
# A8: RangeBN
# ----------------------------------------------------------------------------
def rangebn_scale_fix(chunk_len):
    """quantize.py:473-474 (Python double)."""
    return (0.5 * 0.35) * (1 + (math.pi * math.log(4)) ** 0.5) / ((2 * math.log(chunk_len)) ** 0.5)


def rangebn_stats(xq, num_chunks=16):
    """quantize.py:467-476 minus eps: per-channel mean, mean-of-chunk-max,
    mean-of-chunk-min, and chunk length."""
    B, C, H, W = xq.shape
    y = xq.transpose(0, 1).contiguous().view(C, num_chunks, B * H * W // num_chunks)
    return y.view(C, -1).mean(-1), y.max(-1)[0].mean(-1), y.min(-1)[0].mean(-1), y.size(-1)


def rangebn_forward(xq, gamma, beta, running_mean, running_var, training=True, momentum=0.1,
                    num_chunks=16, eps=1e-5, bits=8, bits_grad=8):
    """quantize.py:463-505 given the already fake-quantized input xq (4-D or 2-D).
    Differentiable through torch autograd exactly like the reference."""
    squeeze = xq.dim() == 2
    if squeeze:
        xq = xq.unsqueeze(-1).unsqueeze(-1)
    if training:
        mean, cmax, cmin, clen = rangebn_stats(xq, num_chunks)
        scale = 1 / ((cmax - cmin) * rangebn_scale_fix(clen) + eps)
        running_mean.detach().mul_(momentum).add_(mean * (1 - momentum))   # :478-479
        running_var.detach().mul_(momentum).add_(scale * (1 - momentum))   # :481-482
    else:
        mean, scale = running_mean, running_var
    scale = quantize(scale, bits, float(scale.min()), float(scale.max()))  # :486-487
    out = (xq - mean.view(1, -1, 1, 1)) * scale.view(1, -1, 1, 1)
    if gamma is not None:
        qg = quantize(gamma, bits, float(gamma.min()), float(gamma.max()))
        out = out * qg.view(1, -1, 1, 1)
    if beta is not None:
        qb = quantize(beta, bits)                                          # :498 (range from data)
        out = out + qb.view(1, -1, 1, 1)
    if bits_grad is not None:
        out = quantize_grad(out, bits_grad)
    if out.size(3) == 1 and out.size(2) == 1:
        out = out.squeeze(-1).squeeze(-1)
    return out
