"""Drop-in mirror of the reference's ``models/modules/quantize.py`` hot path on B200 kernels.

Same names, constructor arguments, buffers / state-dict keys and (effective) semantics as the reference
(file:line citations are relative to /root/reference/models/modules/quantize.py), but every tensor
pass runs in a hand-written sm_100a kernel behind the C ABI (include/qb200.h) and NO value is synced to
the host on the step path: ranges live in device memory (the reference calls ``float(tensor)`` 242 times
per ResNet-18 step), so a whole training step can be captured in a CUDA graph.

Differences that are deliberate and documented in DESIGN.md:
  * the batch-mean of per-sample minima / maxima and the RangeBN chunk means are accumulated in fp64 and
    rounded once (ATen sums in fp32 in an unspecified order): ranges agree to <= 1 ulp, and given the same
    range the integer codes are bit-identical;
  * stochastic-rounding noise comes from Philox4x32-7 inside the kernel (torch's CPU mt19937 stream cannot
    be reproduced on a GPU); ``set_noise_provider`` injects a recorded noise tensor for parity tests.
"""
import os
import itertools
import math

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.autograd.function import Function, InplaceFunction

from . import _abi

NATIVE_GLUE = True      # models.py: use glue.MaxPool2d (libqb200) between these modules
from ._abi import QB200Error, QB_SCALE_DOUBLE, QB_SCALE_FLOAT

__all__ = ["UniformQuantize", "UniformQuantizeGrad", "quantize", "quantize_grad", "conv2d_biprec", "linear_biprec",
           "QuantNode", "QuantMeasure", "QConv2d", "QLinear", "RangeBN", "set_measure_mode", "set_quant_mode",
           "overwrite_params", "freeze_quant_params", "set_bn_is_train", "set_global_quantization_method",
           "distill_set_train", "is_bn", "is_quant", "recursive_apply", "set_noise_provider", "manual_seed",
           "noise_step", "set_stat_reducer", "set_noise_rank"]

# ---------------------------------------------------------------------------------------------------
# noise source of the stochastic branch + cross-rank statistic hooks
# ---------------------------------------------------------------------------------------------------
# Activations flow channels-last (NHWC) through the network: QConv2d converts a 4-D NCHW-contiguous input once and
# every kernel preserves the layout.  Logical shapes, indexing and results are unchanged (torch.channels_last).
CHANNELS_LAST = True

_noise_provider = None
_seed_base = 0x5DEECE66D
_seed_counter = itertools.count(1)
_step_tensors = {}
_reducer = None


def set_noise_provider(fn):
    """fn(shape) -> CUDA float32 tensor of uniform(-0.5,0.5) noise (or None for Philox).  Parity tests replay
    the noise the reference drew (quantize.py:93).  Returns the previous provider."""
    global _noise_provider
    prev, _noise_provider = _noise_provider, fn
    return prev


def manual_seed(seed):
    """Re-key the in-kernel Philox stream (the call counter restarts too)."""
    global _seed_base, _seed_counter
    _seed_base = int(seed) & (2 ** 63 - 1)
    _seed_counter = itertools.count(1)


def noise_step(device=None):
    """Device-resident step counter mixed into every Philox key.  ``noise_step().add_(1)`` inside a captured
    CUDA graph makes each replay draw fresh noise (host-side seeds are frozen at capture)."""
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    t = _step_tensors.get(device)
    if t is None:
        t = _step_tensors[device] = torch.zeros(1, dtype=torch.int64, device=device)
    return t


_seed_rank = 0


def set_noise_rank(rank):
    """Batch-sharded training (dist.py): mix the rank into every Philox key so that the shards draw INDEPENDENT noise
    fields (the same local element index on two ranks is a different element of the global batch)."""
    global _seed_rank
    _seed_rank = int(rank)


def _next_seed():
    return (_seed_base + next(_seed_counter) * 0x9E3779B97F4A7C15 + _seed_rank * 0xD1B54A32D192ED03) & (2 ** 64 - 1)


def set_stat_reducer(reducer):
    """Install the cross-rank statistic reducer used by batch-sharded training (see ``dist.py``): an object with
    ``measure(stats)``, ``rangebn(red)``, ``rangebn_bwd(sums)`` and ``grad_range(range)`` that all-reduce IN PLACE.
    ``None`` restores single-process behaviour."""
    global _reducer
    prev, _reducer = _reducer, reducer
    return prev


# Producer-supplied gradient ranges.  RangeBN's backward kernel computes min/max of the dx it writes; the QConv2d in
# front of it needs exactly that range for its own gradient quantizer (quantize.py:126,132).  The producer stashes
# (dx, range) here; the consumer takes it if - and only if - the gradient it receives is that very tensor (same
# storage, shape, strides and version; the strong reference keeps the storage from being recycled meanwhile).
_grad_ranges = {}


def _stash_grad_range(t, rng):
    if len(_grad_ranges) >= 4:
        _grad_ranges.clear()
    _grad_ranges[t.data_ptr()] = (t, t._version, rng)


def _take_grad_range(g):
    ent = _grad_ranges.pop(g.data_ptr(), None)
    if ent is None:
        return None
    t, ver, rng = ent
    if t.shape == g.shape and t.stride() == g.stride() and t.dtype == g.dtype and g._version == ver:
        return rng
    return None


# Producer-supplied forward statistics, same protocol: the int8 convolution epilogue emits the QuantMeasure statistics
# block of its output; the quantize_input of the RangeBN that follows takes it instead of reading the tensor again.
_fwd_stats = {}


def _stash_fwd_stats(t, stats):
    if len(_fwd_stats) >= 4:
        _fwd_stats.clear()
    _fwd_stats[t.data_ptr()] = (t, t._version, stats)


def _take_fwd_stats(x):
    ent = _fwd_stats.pop(x.data_ptr(), None)
    if ent is None:
        return None
    t, ver, stats = ent
    if t.shape == x.shape and t.stride() == x.stride() and t.dtype == x.dtype and x._version == ver:
        return stats
    return None


def _grad_noise(g, stochastic):
    noise = _noise_provider(tuple(g.shape)) if (stochastic and _noise_provider is not None) else None
    if noise is not None and _abi.is_nhwc(g):
        noise = noise.reshape(g.shape).contiguous(memory_format=torch.channels_last)   # same logical element <-> same noise
    return noise


def _quantize_gradient(g, bits, stochastic=True):
    """UniformQuantizeGrad.backward (quantize.py:124-139): global min/max, zero-point stochastic quantizer."""
    g = _abi._f32c(g, "grad")
    noise = _grad_noise(g, stochastic)
    r = _take_grad_range(g)
    if _reducer is None and r is None:
        out, _ = _abi.fakequant_grad_fused(g, bits, noise=noise, stochastic=stochastic, seed=_next_seed(),
                                           step_dev=noise_step(g.device))
        return out
    if r is None:
        r = _abi.minmax(g)
    if _reducer is not None:
        _reducer.grad_range(r)
    return _abi.fakequant_grad(g, bits, range_dev=r, noise=noise, stochastic=stochastic, seed=_next_seed(),
                               step_dev=noise_step(g.device))


def _pack_gradient(g, bits, want_codes, want_split, pad_cols=0):
    """UniformQuantizeGrad.backward's quantizer (quantize.py:124-139) feeding the native backward contractions: one pass
    over g -> (range, g (padded), bf16 codes (q - zp) | None, bf16 hi | None, bf16 lo | None).  ``pad_cols`` zero columns
    are appended to a 2-D g first (linear layers whose width is not a multiple of 8); the range is that of the unpadded
    tensor and a padded zero keeps code zp, i.e. contributes nothing."""
    g = _abi._f32c(g, "grad")
    r = _take_grad_range(g)
    if r is None:
        r = _abi.minmax(g)
    if _reducer is not None:
        _reducer.grad_range(r)
    noise = _grad_noise(g, True) if want_codes else None
    if pad_cols:
        g = F.pad(g, (0, pad_cols))
        if noise is not None:
            noise = F.pad(noise.reshape(g.shape[0], -1), (0, pad_cols))
    gq, hi, lo = _abi.bwd_pack(g, r, bits, noise=noise, stochastic=True, seed=_next_seed() if want_codes else 0,
                               step_dev=noise_step(g.device), want_codes=want_codes, want_split=want_split)
    return r, g, gq, hi, lo


# ---------------------------------------------------------------------------------------------------
# UniformQuantize / UniformQuantizeGrad / quantize / quantize_grad         (quantize.py:39-164)
# ---------------------------------------------------------------------------------------------------
def _is_scalar_like(v):
    return (not torch.is_tensor(v)) or v.numel() == 1


class UniformQuantize(InplaceFunction):
    """quantize.py:39-109.  Same positional signature as the reference's forward."""

    @classmethod
    def forward(cls, ctx, input, num_bits=8, min_value=None, max_value=None, stochastic=False, inplace=False,
                enforce_true_zero=True, num_chunks=None, out_half=False):
        num_chunks = input.shape[0] if num_chunks is None else num_chunks
        if (min_value is None or max_value is None) and max(input.shape[0] // num_chunks, 1) != 1:
            raise QB200Error("UniformQuantize: per-group data ranges (B // num_chunks > 1) do not broadcast in the "
                             "reference either; unsupported")
        ctx.inplace, ctx.num_bits, ctx.stochastic = inplace, num_bits, stochastic
        ctx.min_value, ctx.max_value = min_value, max_value
        x = input.contiguous()
        if inplace:
            ctx.mark_dirty(input)
        out = input if (inplace and x is input) else None
        dev_range, lo, hi, mode = None, 0.0, 0.0, QB_SCALE_DOUBLE
        per_row = False
        if min_value is None or max_value is None:
            dev_range = _abi.minmax(x)                       # :45-55 -> 1-element tensors => tensor-scale branch
            mode = QB_SCALE_FLOAT
            if min_value is not None or max_value is not None:
                given_lo = dev_range[0] if min_value is None else torch.as_tensor(min_value, device=x.device).reshape(())
                given_hi = dev_range[1] if max_value is None else torch.as_tensor(max_value, device=x.device).reshape(())
                dev_range = torch.stack([given_lo, given_hi]).float()
        elif torch.is_tensor(min_value) or torch.is_tensor(max_value):
            mode = QB_SCALE_FLOAT
            if _is_scalar_like(min_value) and _is_scalar_like(max_value):
                dev_range = torch.stack([torch.as_tensor(min_value, device=x.device).reshape(()).float(),
                                         torch.as_tensor(max_value, device=x.device).reshape(()).float()])
            else:
                rows = x.shape[0]
                if not (torch.is_tensor(min_value) and torch.is_tensor(max_value) and min_value.numel() == rows
                        and max_value.numel() == rows and min_value.shape[0] == rows):
                    raise QB200Error("UniformQuantize: tensor ranges must be scalar or one value per row of dim 0")
                per_row = True
        else:
            lo, hi = float(min_value), float(max_value)
        if enforce_true_zero:
            if per_row:
                raise QB200Error("UniformQuantize: enforce_true_zero needs a scalar range (as in the reference)")
            noise = _noise_provider(tuple(x.shape)) if (stochastic and _noise_provider is not None) else None
            y = _abi.fakequant_grad(x, num_bits, range_dev=dev_range, lo=lo, hi=hi, noise=noise,
                                    stochastic=bool(stochastic), seed=_next_seed(), step_dev=noise_step(x.device), out=out)
        else:
            if stochastic:
                if per_row:
                    raise QB200Error("UniformQuantize: stochastic rounding with per-row ranges is not implemented")
                noise = _noise_provider(tuple(x.shape)) if _noise_provider is not None else None
                if noise is not None and _abi.is_nhwc(x):
                    noise = noise.reshape(x.shape).contiguous(memory_format=torch.channels_last)
                y = _abi.fakequant_affine_stochastic(x, num_bits, range_dev=dev_range, lo=lo, hi=hi, scale_mode=mode, noise=noise,
                                                     seed=_next_seed(), step_dev=noise_step(x.device), out=out)
            elif per_row:
                y, _, _ = _abi.fakequant_rows(x, num_bits, row_min=min_value.detach().reshape(-1).contiguous().float(),
                                              row_max=max_value.detach().reshape(-1).contiguous().float(),
                                              compute_range=False, out=out)
            else:
                y = _abi.fakequant_affine(x, num_bits, range_dev=dev_range, lo=lo, hi=hi, scale_mode=mode, out=out)
        if inplace and y is not input:
            input.copy_(y)
            y = input
        if out_half and num_bits <= 16:
            y = y.half()
        return y

    @staticmethod
    def backward(ctx, grad_output):
        # straight-through estimator; one None per remaining forward argument actually passed
        return (grad_output,) + (None,) * 8


class UniformQuantizeGrad(InplaceFunction):
    """quantize.py:112-139: identity forward, quantized gradient backward."""

    @classmethod
    def forward(cls, ctx, input, num_bits=8, min_value=None, max_value=None, stochastic=True, inplace=False):
        ctx.inplace, ctx.num_bits, ctx.stochastic = inplace, num_bits, stochastic
        ctx.min_value, ctx.max_value = min_value, max_value
        # the reference returns `input` itself (:121), which current torch rejects once a consumer modifies it
        # in place; a fresh tensor is the numerically neutral equivalent
        return input.clone()

    @staticmethod
    def backward(ctx, grad_output):
        if ctx.min_value is None and ctx.max_value is None:
            grad_input = _quantize_gradient(grad_output, ctx.num_bits, ctx.stochastic)
        else:
            g = grad_output.contiguous()
            r = _abi.minmax(g) if (ctx.min_value is None or ctx.max_value is None) else None
            lo = r[0] if ctx.min_value is None else torch.as_tensor(float(ctx.min_value), device=g.device)
            hi = r[1] if ctx.max_value is None else torch.as_tensor(float(ctx.max_value), device=g.device)
            rng = torch.stack([lo.reshape(()), hi.reshape(())]).float()
            noise = _noise_provider(tuple(g.shape)) if (ctx.stochastic and _noise_provider is not None) else None
            grad_input = _abi.fakequant_grad(g, ctx.num_bits, range_dev=rng, noise=noise, stochastic=bool(ctx.stochastic),
                                             seed=_next_seed(), step_dev=noise_step(g.device))
        return grad_input, None, None, None, None, None


def quantize(x, num_bits=8, min_value=None, max_value=None, num_chunks=None, stochastic=False, inplace=False):
    """quantize.py:159-160 INCLUDING its positional shuffle: the reference passes
    (x, num_bits, min, max, num_chunks, stochastic, inplace) into forward's
    (input, num_bits, min, max, stochastic, inplace, enforce_true_zero, ...) slots, i.e. effectively
    stochastic := num_chunks, inplace := stochastic, enforce_true_zero := inplace (False by default)."""
    return UniformQuantize().apply(x, num_bits, min_value, max_value, num_chunks, stochastic, inplace)


def quantize_grad(x, num_bits=8, min_value=None, max_value=None, stochastic=True, inplace=False):
    """quantize.py:163-164."""
    return UniformQuantizeGrad().apply(x, num_bits, min_value, max_value, stochastic, inplace)


# ---------------------------------------------------------------------------------------------------
# contraction with the bi-precision backward                                  (quantize.py:142-156)
# ---------------------------------------------------------------------------------------------------
def _pair(v):
    return tuple(v) if isinstance(v, (tuple, list)) else (v, v)


class _QContraction(Function):
    """out = conv2d / linear of already fake-quantized operands.

    backward, per the reference:
      grad_mode 'full'   (num_bits_grad None)                : dX, dW, db from g
      grad_mode 'biprec' (conv2d_biprec / linear_biprec)     : dW, db from g;  dX from Q(g)    (:143-148)
      grad_mode 'quant'  (quantize_grad after the op, :342-346): dX, dW, db all from Q(g)
    The reference evaluates the convolution twice in biprec mode (out1 + out2 - out1.detach(), identical
    values); here it is evaluated once and only the backward differs.
    """

    @staticmethod
    def forward(ctx, x, w, b, conv, bits_grad, grad_mode, grad_quantizer=None):
        ctx.conv, ctx.bits_grad, ctx.grad_mode = conv, bits_grad, grad_mode
        ctx.grad_quantizer = grad_quantizer          # quantize2 flavour: callable(g) -> Q(g); None = quantize.py's
        ctx.has_bias = b is not None
        ctx.save_for_backward(x, w)
        if ALLOW_LIBRARY_FALLBACK or not x.is_cuda:
            if conv is None:
                return F.linear(x, w, b)
            stride, padding, dilation, groups = conv
            return F.conv2d(x, w, b, stride, padding, dilation, groups)
        # no tensor-core kernel for this geometry: the native direct kernels (conv_generic.cu), never cuDNN / cuBLAS
        if conv is None and x.dim() != 2:
            return _abi.conv_generic_fwd(x.reshape(-1, x.shape[-1]), w, b, None).reshape(x.shape[:-1] + (w.shape[0],))
        return _abi.conv_generic_fwd(x, w, b, conv)

    @staticmethod
    def backward(ctx, g):
        x, w = ctx.saved_tensors
        return _contraction_backward(ctx, g, x, w) + (None, None, None, None)


def _contraction_backward(ctx, g, x, w):
    """Gradients of out = op(x, w, b) w.r.t. (x, w, b) under ctx.grad_mode (see _QContraction)."""
    need_x, need_w, need_b = ctx.needs_input_grad[0], ctx.needs_input_grad[1], ctx.has_bias and ctx.needs_input_grad[2]
    g = _abi._f32c(g, "grad")
    gq = None
    if ctx.grad_mode != "full" and (need_x or ctx.grad_mode == "quant"):
        custom = getattr(ctx, "grad_quantizer", None)
        gq = custom(g) if custom is not None else _quantize_gradient(g, ctx.bits_grad)
    g_param = gq if ctx.grad_mode == "quant" else g          # weight / bias branch
    g_input = g if ctx.grad_mode == "full" else gq           # input branch
    dx = dw = db = None
    if g.is_cuda and not ALLOW_LIBRARY_FALLBACK and not (
            ctx.conv is not None and USE_DW_KERNEL and ctx.conv[3] > 1 and ctx.conv[3] == x.shape[1] and ctx.conv[2] == (1, 1)
            and ctx.conv[0][0] == ctx.conv[0][1] and ctx.conv[1][0] == ctx.conv[1][1]
            and _abi.conv_dw_bwd_supported(g, x, w, ctx.conv[0][0], ctx.conv[1][0])):
        # native direct kernels for the geometries without a tensor-core / depthwise kernel (conv_generic.cu)
        if ctx.conv is None and x.dim() != 2:
            x2, gi2, gp2 = x.reshape(-1, x.shape[-1]), g_input.reshape(-1, w.shape[0]), g_param.reshape(-1, w.shape[0])
        else:
            x2, gi2, gp2 = x, g_input, g_param
        if need_x:
            dx = _abi.conv_generic_dgrad(gi2, w, tuple(x2.shape), ctx.conv, like=x2).reshape(x.shape)
        if need_w or need_b:
            dw, db = _abi.conv_generic_wgrad(gp2, x2, tuple(w.shape), ctx.conv, want_w=bool(need_w), want_bias=bool(need_b))
        return dx, dw, db
    if ctx.conv is None:
        if need_x:
            dx = g_input.matmul(w)
        if need_w:
            dw = g_param.reshape(-1, g_param.shape[-1]).t().matmul(x.reshape(-1, x.shape[-1]))
        if need_b:
            db = g_param.reshape(-1, g_param.shape[-1]).sum(0)
    else:
        stride, padding, dilation, groups = ctx.conv
        bias_sizes = [w.shape[0]] if ctx.has_bias else None
        if (USE_DW_KERNEL and groups > 1 and groups == x.shape[1] and dilation == (1, 1) and stride[0] == stride[1]
                and padding[0] == padding[1] and g.is_cuda and _abi.conv_dw_bwd_supported(g, x, w, stride[0], padding[0])):
            # depthwise 3x3: native streaming kernels (conv_dw.cu) instead of cuDNN's grouped backward + a bias reduction
            if need_x:
                dx = _abi.conv_dw_dgrad(g_input, w, x.shape, stride[0], padding[0])
            if need_w or need_b:
                dw, db = _abi.conv_dw_wgrad(g_param, x, stride[0], padding[0], bool(need_b))
            return dx, (dw if need_w else None), db
        if need_x:
            dx = torch.ops.aten.convolution_backward(g_input, x, w, bias_sizes, stride, padding, dilation, False,
                                                     [0, 0], groups, [True, False, False])[0]
        if need_w or need_b:
            _, dw, db = torch.ops.aten.convolution_backward(g_param, x, w, bias_sizes, stride, padding, dilation,
                                                            False, [0, 0], groups, [False, need_w, bool(need_b)])
    return dx, dw, db


def conv2d_biprec(input, weight, bias=None, stride=1, padding=0, dilation=1, groups=1, num_bits_grad=None):
    """quantize.py:142-148."""
    conv = (_pair(stride), _pair(padding), _pair(dilation), groups)
    return _QContraction.apply(input, weight, bias, conv, num_bits_grad, "biprec" if num_bits_grad is not None else "full")


def linear_biprec(input, weight, bias=None, num_bits_grad=None):
    """quantize.py:151-156."""
    return _QContraction.apply(input, weight, bias, None, num_bits_grad, "biprec" if num_bits_grad is not None else "full")


# ---------------------------------------------------------------------------------------------------
# QuantNode / QuantMeasure                                                   (quantize.py:177-268)
# ---------------------------------------------------------------------------------------------------
class QuantNode():
    def __init__(self):
        self.enable_quant = True
        self.freeze_param_dyn_range = False

    def set_measure_mode(self, measure, momentum=None):
        self.enable_quant = not measure
        if momentum and isinstance(self, QuantMeasure):
            self.momentum = momentum
        elif isinstance(self, nn.Module):
            for q in self._modules.values():
                if isinstance(q, QuantNode):
                    q.set_measure_mode(measure, momentum=momentum)

    def overwrite_params(self, logging=None):
        if isinstance(self, nn.Module):
            for q in self._modules.values():
                if isinstance(q, QuantNode):
                    q.overwrite_params(logging)


class _AffineSTE(Function):
    """Min-anchored fake-quant with a device-resident range; straight-through backward (:105-109)."""

    @staticmethod
    def forward(ctx, x, range_dev, bits):
        return _abi.fakequant_affine(x, bits, range_dev=range_dev, scale_mode=QB_SCALE_DOUBLE)

    @staticmethod
    def backward(ctx, g):
        return g, None, None


class QuantMeasure(nn.Module, QuantNode):
    """Running activation-range tracker + activation fake-quant (quantize.py:198-268)."""
    _QMEASURE_SUPPORTED_METHODS = ['avg', 'aciq']

    def __init__(self, num_bits=8, momentum=None, method='avg'):
        super(QuantMeasure, self).__init__()
        QuantNode.__init__(self)
        assert method in QuantMeasure._QMEASURE_SUPPORTED_METHODS
        self.register_buffer('running_min', torch.zeros(1))
        self.register_buffer('running_max', torch.zeros(1))
        self.register_buffer('num_measurements', torch.zeros(1))
        self.register_buffer('running_var', torch.ones(1))
        self.register_buffer('running_mean', torch.zeros(1))
        self.momentum = momentum
        self.num_bits = num_bits
        self.method = method
        self.laplace_alpha = {2: 2.83, 3: 3.89, 4: 5.03, 5: 6.2, 6: 7.41, 7: 8.64, 8: 9.89}.get(num_bits)

    def _buffers5(self):
        return (self.running_min, self.running_max, self.num_measurements, self.running_var, self.running_mean)

    def measure(self, input):
        """Update the running buffers (training) and return the device range float[2] this call uses.
        quantize.py:223-246 without the float() syncs."""
        method = 1 if self.method == 'aciq' else 0
        if method:
            assert self.laplace_alpha, 'aciq not supported for module num bits'
        x = input.detach()
        if self.training:
            stats = _take_fwd_stats(x)
            if stats is None:
                stats = _abi.measure_stats(x)
            if _reducer is not None:
                _reducer.measure(stats)
            return _abi.measure_finalize(stats, self._buffers5(), self.momentum, True, method, self.num_bits)
        return _abi.measure_finalize(None, self._buffers5(), self.momentum, False, method, self.num_bits)

    def forward(self, input):
        rng = self.measure(input)
        if self.enable_quant:
            return _AffineSTE.apply(input, rng, self.num_bits)
        return input

    def _get_measured_range(self):
        return float(self.running_min), float(self.running_max)


# ---------------------------------------------------------------------------------------------------
# QConv2d / QLinear                                                          (quantize.py:271-432)
# ---------------------------------------------------------------------------------------------------
class _RowsSTE(Function):
    """Per-output-channel weight fake-quant (:318-323, 332-334), straight-through backward.  The row ranges are
    written straight into the module's weight_min / weight_max buffers (the reference re-assigns them each
    forward); ``compute_range`` False reuses the stored ranges (frozen dynamic range, overwrite_params)."""

    @staticmethod
    def forward(ctx, w, row_min, row_max, bits, compute_range):
        out, _, _ = _abi.fakequant_rows(w, bits, row_min=row_min, row_max=row_max, compute_range=compute_range)
        return out

    @staticmethod
    def backward(ctx, g):
        return g, None, None, None, None


class _ScalarRangeSTE(Function):
    """Per-tensor fake-quant with the range taken from the tensor itself and a tensor-typed scale: the bias
    path (:328-330, 336-338).  Returns (quantized, range float[2])."""

    @staticmethod
    def forward(ctx, v, bits):
        rng = _abi.minmax(v)
        out = _abi.fakequant_affine(v, bits, range_dev=rng, scale_mode=QB_SCALE_FLOAT)
        ctx.mark_non_differentiable(rng)
        return out, rng

    @staticmethod
    def backward(ctx, g, _g_rng):
        return g, None


class _QBase(QuantNode):
    """State and forward logic shared by QConv2d / QLinear."""

    def _init_quant(self, num_bits, num_bits_weight, num_bits_grad, biprecision, bias_quant, per_channel, bias):
        QuantNode.__init__(self)
        self.num_bits = num_bits
        self.num_bits_weight = num_bits_weight or num_bits
        self.num_bits_grad = num_bits_grad
        self.quantize_input = QuantMeasure(self.num_bits)
        self.biprecision = biprecision
        self.bias_quant = bias_quant and bias
        self.per_channel = per_channel
        if not per_channel:
            # the reference assigns a Python float to a registered buffer in forward and raises
            raise QB200Error("per_channel=False is not usable in the reference (quantize.py:325-326 raises); "
                             "only per-output-channel weight ranges are implemented")
        n_channels = self.weight.size(0)
        self.scale_shape = (n_channels,) + (1,) * (self.weight.dim() - 1)
        with torch.no_grad():
            flat = self.weight.flatten(1)
            self.register_buffer('weight_min', flat.min(-1)[0].view(self.scale_shape).clone())
            self.register_buffer('weight_max', flat.max(-1)[0].view(self.scale_shape).clone())
            if self.bias_quant:
                self.register_buffer('bias_min', self.bias.min().clone())
                self.register_buffer('bias_max', self.bias.max().clone())

    def _quantized_params(self):
        """qweight, qbias as the forward uses them (:317-340)."""
        update = not self.freeze_param_dyn_range
        qweight = _RowsSTE.apply(self.weight, self.weight_min, self.weight_max, self.num_bits_weight, update)
        return qweight, self._quantized_bias()

    def _quantized_bias(self):
        update = not self.freeze_param_dyn_range
        qbias = self.bias
        if self.bias is not None and (self.bias_quant or update):
            if update:
                qb, rng = _ScalarRangeSTE.apply(self.bias, self.num_bits_weight)
                if hasattr(self, 'bias_min'):
                    with torch.no_grad():
                        self.bias_min.copy_(rng[0])
                        self.bias_max.copy_(rng[1])
                if self.bias_quant:
                    qbias = qb
            else:
                qbias = UniformQuantize().apply(self.bias, self.num_bits_weight, self.bias_min, self.bias_max, False,
                                                False, False)
        return qbias

    def _grad_mode(self):
        if self.num_bits_grad is None:
            return "full"
        return "biprec" if self.biprecision else "quant"

    def overwrite_params(self, logging=None):
        """quantize.py:301-312 / 386-396: bake the quantized weights (ranges as stored) into the parameters."""
        with torch.no_grad():
            w, _, _ = _abi.fakequant_rows(self.weight, self.num_bits_weight, row_min=self.weight_min,
                                          row_max=self.weight_max, compute_range=False)
            self.weight.copy_(w)
            if self.bias_quant:
                self.bias.copy_(UniformQuantize().apply(self.bias, self.num_bits_weight, self.bias_min, self.bias_max,
                                                        False, False, False))


# Route QConv2d / QLinear through the int8 tcgen05 kernels whenever the geometry is supported (conv_i8.cu,
# gemm_i8.cu); otherwise the contraction of the already fake-quantized tensors goes to cuDNN / cuBLAS.
USE_INT8_TENSOR_CORES = True
USE_NATIVE_BACKWARD = True    # dX / dW / db of the dense contractions on conv_bwd.cu (tcgen05 bf16) instead of cuDNN / cuBLAS
# The reference's F.conv2d / F.linear on the fake-quantized tensors (cuDNN / cuBLAS) is NOT a product path: a geometry the
# tensor-core kernels do not take runs on the native direct kernels (conv_generic.cu).  Tests set this flag to compare against
# the library.
ALLOW_LIBRARY_FALLBACK = False
USE_STEM_KERNEL = True
FUSE_BN_POOL = os.environ.get("QB_FUSE_BN_POOL", "1") != "0"     # stem: max-pool straight from RangeBN's codes (RangeBN.forward_relu_pool)
FUSE_QUANT_ROWSUM = os.environ.get("QB_FUSE_ROWSUM", "1") != "0"      # conv input quantizer also emits the per-pixel code sums (qb_fakequant_affine_rowsum)
# one process: RangeBN statistics + per-channel finalize in one call (qb_rangebn_stats_finalize_nhwc).  Off by default: measured
# 0.3 % SLOWER per training step than the two launches (profiles/ab_fusions_r02r.txt) - the single CTA serialises the fold of
# the chunk accumulators and the finalize, two small kernels overlap their tails inside the CUDA graph
FUSE_RANGEBN_FINALIZE = os.environ.get("QB_FUSE_RBN_FINALIZE", "0") == "1"
STEM_NCHW_DIRECT = os.environ.get("QB_STEM_NCHW", "1") != "0"       # first layer: quantize the NCHW image batch straight into channels-last codes (no fp32 layout copy)
USE_DW_KERNEL = True       # depthwise 3x3 convs on the codes (conv_dw.cu); False: F.conv2d on the fake-quantized tensors


def _conv_dw_supported(mod, x):
    """Depthwise (groups == C_in == C_out) 3x3 geometry: streaming integer kernel on the codes (conv_dw.cu)."""
    st, pd, dl = _pair(mod.stride), _pair(mod.padding), _pair(mod.dilation)
    kh, kw = mod.kernel_size
    return (USE_DW_KERNEL and mod.groups == mod.in_channels == mod.out_channels and mod.groups > 1 and dl == (1, 1)
            and st[0] == st[1] and pd[0] == pd[1] and not isinstance(pd[0], str) and mod.num_bits <= 8
            and mod.num_bits_weight <= 8 and (x.shape[2] + 2 * pd[0] - kh) // st[0] + 1 >= 1
            and (x.shape[3] + 2 * pd[1] - kw) // st[1] + 1 >= 1 and x.numel() < 2 ** 31
            and _abi.conv_dw_supported(mod.in_channels, kh, kw, st[0], pd[0]))


def _conv_i8_supported(mod, x):
    if not (USE_INT8_TENSOR_CORES and x.dim() == 4 and mod.padding_mode == 'zeros'):
        return False
    if mod.groups != 1:
        return _conv_dw_supported(mod, x)
    kh, kw = mod.kernel_size
    st, pd, dl = _pair(mod.stride), _pair(mod.padding), _pair(mod.dilation)
    if not (dl == (1, 1) and st[0] == st[1] and st[0] in (1, 2) and pd[0] == pd[1] and not isinstance(pd[0], str)
            and mod.num_bits <= 8 and mod.num_bits_weight <= 8
            and (x.shape[2] + 2 * pd[0] - kh) // st[0] + 1 >= 1 and (x.shape[3] + 2 * pd[1] - kw) // st[1] + 1 >= 1):
        return False
    if kh * kw <= 9 and mod.in_channels % 16 == 0 and mod.out_channels % 16 == 0:
        return True
    return _conv_stem_supported(mod, x)


def _conv_stem_supported(mod, x):
    """First-layer geometry (3 input channels): tcgen05 over a thread-built im2col operand (conv_stem.cu)."""
    kh, kw = mod.kernel_size
    return (USE_STEM_KERNEL and mod.in_channels % 16 != 0
            and _abi.conv_stem_supported(x.shape[2], x.shape[3], mod.in_channels, mod.out_channels, kh, kw,
                                         _pair(mod.stride)[0], _pair(mod.padding)[0]))


def _stem_takes_nchw(mod, x):
    """The image batch of the first layer stays NCHW: its quantizer writes the channels-last codes itself
    (qb_fakequant_affine_nchw3), so the fp32 batch is read once by the statistics pass and once by the quantizer and never
    copied.  Only where nothing downstream wants the fp32 x_hat (no input gradient; weight gradient from the codes)."""
    return (STEM_NCHW_DIRECT and mod.enable_quant and x.is_cuda and x.dtype == torch.float32 and x.is_contiguous()
            and mod.in_channels == 3 and mod.groups == 1 and (x.shape[2] * x.shape[3]) % 4 == 0 and x.data_ptr() % 16 == 0
            and not (x.requires_grad and torch.is_grad_enabled()) and getattr(mod, "_q2_params", None) is None
            and mod.out_channels in (32, 64) and mod.kernel_size[0] in (3, 7) and mod.kernel_size[0] == mod.kernel_size[1]
            and _conv_i8_supported(mod, x))


def _linear_i8_supported(mod, x):
    return (USE_INT8_TENSOR_CORES and x.dim() == 2 and mod.in_features % 16 == 0 and mod.num_bits <= 8
            and mod.num_bits_weight <= 8)


def _native_bwd_supported(mod, x):
    """Geometry of the dense backward kernels (conv_bwd.cu): groups 1, C_in % 16 == 0, at most 9 taps, stride 1 or 2."""
    if not USE_NATIVE_BACKWARD or mod.num_bits > 8:
        return False
    if isinstance(mod, nn.Conv2d):
        kh, kw = mod.kernel_size
        return (mod.groups == 1 and mod.in_channels % 16 == 0 and mod.out_channels % 8 == 0
                and _abi.conv_bwd_supported(mod.in_channels, mod.out_channels, kh, kw, _pair(mod.stride)[0], _pair(mod.padding)[0]))
    return mod.in_features % 16 == 0


def _pack_wt(wq):
    """bf16 hi / lo planes of w_hat for the input-gradient kernel; linear layers: width padded to 16-byte rows."""
    pad = (-wq.shape[0]) % 8
    return _abi.wt_pack(F.pad(wq, (0, 0) * (wq.dim() - 1) + (0, pad)) if pad else wq)


# Weight preparation aside.  The per-layer weight kernels of a step (per-channel fake-quant + codes + tap sums, hi / lo planes for
# the input gradient: ~2 single-wave launches per layer) depend on nothing but the weights, which are final when the step
# starts.  prepare_weights(model) issues them all on a side stream at the top of the model's forward; each layer then waits
# for its own event instead of running them in line.  Inside a captured CUDA graph this is a parallel branch.
# Off by default: measured 1 % (training) / 3.6 % (evaluation) SLOWER than in-line preparation inside the captured graph
# (profiles/ab_fusions_r02s.txt) - the cross-branch dependencies cost more than the ~40 tiny launches they take off the main path.
PREP_WEIGHTS_ASIDE = os.environ.get("QB_PREP_ASIDE", "0") == "1"
_side_streams = {}
_prep_tail = {}


def prepare_weights(model):
    if not PREP_WEIGHTS_ASIDE or not torch.cuda.is_available():
        return
    mods = model.__dict__.get("_qb_prep_list")
    if mods is None:
        mods = [m for m in model.modules() if isinstance(m, (QConv2d, QLinear))]
        model.__dict__["_qb_prep_list"] = mods
    todo = [m for m in mods if m.__dict__.get("_tc_last") is not None and m.weight.is_cuda and m.enable_quant]
    if not todo:
        return                                   # first forward: the layers prepare in line and record what they needed
    dev = todo[0].weight.device
    main = torch.cuda.current_stream(dev)
    side = _side_streams.get(dev)
    if side is None:
        side = _side_streams[dev] = torch.cuda.Stream(device=dev)
    side.wait_stream(main)                       # the weights (and last step's users of the recycled buffers) are done
    grad = torch.is_grad_enabled()
    with torch.cuda.stream(side):
        for m in todo:
            want_y, want_wt = m.__dict__["_tc_last"]
            want_y = want_y and grad and m.weight.requires_grad
            w = m.weight.detach()
            w4 = w if isinstance(m, nn.Conv2d) else w.view(w.shape[0], w.shape[1], 1, 1)
            update = not getattr(m, "freeze_param_dyn_range", False)
            wq, wcodes, tapsum = _abi.fakequant_rows_conv(w4, m.num_bits_weight, m.weight_min, m.weight_max,
                                                          compute_range=update, want_y=want_y)
            dense = not isinstance(m, nn.Conv2d) or m.groups == 1
            wt = _pack_wt(wq) if (want_y and want_wt and dense) else None
            ev = torch.cuda.Event()
            ev.record(side)
            m.__dict__["_prepped"] = ((w.data_ptr(), w._version), (wq, wcodes, tapsum, wt), ev)
        tail = torch.cuda.Event()
        tail.record(side)
    _prep_tail[dev] = tail


def join_prepared(device=None):
    """The main stream waits for whatever prepare_weights issued (a layer that did not run leaves its branch unjoined, which a
    graph capture does not accept)."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    tail = _prep_tail.pop(dev, None)
    if tail is not None:
        torch.cuda.current_stream(dev).wait_event(tail)


class _QTensorCoreFn(Function):
    """QConv2d / QLinear on the integer codes.  Forward: activation fake-quant (uint8 codes for the int8 tensor cores, the
    same codes as bf16 for the weight gradient), per-channel weight fake-quant (centred int8 codes; bf16 hi / lo planes
    of w_hat for the input gradient), int8 tcgen05 implicit GEMM with the dequantising epilogue.  Backward
    (quantize.py:142-156, 342-346; straight-through for both quantizers, :105-109): one pass packs the output gradient
    (quantizer -> bf16 codes, raw g -> bf16 hi / lo), then qb_conv_dgrad / qb_conv_wgrad (conv_bwd.cu)."""

    @staticmethod
    def forward(ctx, input, weight, qbias, mod, rng):
        conv = isinstance(mod, nn.Conv2d)
        need_bwd = any(ctx.needs_input_grad[:3])
        x = _abi._f32c(input, "input")
        bits_a, bits_w = mod.num_bits, mod.num_bits_weight
        update = not getattr(mod, "freeze_param_dyn_range", False)
        nchw_stem = conv and not ctx.needs_input_grad[0] and _stem_takes_nchw(mod, x)    # left NCHW by QConv2d.forward
        codes = None if nchw_stem else torch.empty_like(x, dtype=torch.uint8)           # same memory format as x
        # first layer whose input needs no gradient: the weight gradient is taken from the codes (conv_stem.cu)
        stem_bwd = (conv and USE_STEM_KERNEL and x.shape[1] % 32 != 0 and not ctx.needs_input_grad[0]
                    and mod.out_channels in (32, 64) and mod.kernel_size[0] in (3, 7)
                    and mod.kernel_size[0] == mod.kernel_size[1])
        dense = not conv or (mod.groups == 1 and x.shape[1] % 16 == 0)
        native_bwd = need_bwd and dense and _native_bwd_supported(mod, x)
        want_x16 = native_bwd and ctx.needs_input_grad[1]
        x16 = torch.empty_like(x, dtype=torch.bfloat16) if want_x16 else None
        q2 = getattr(mod, "_q2_params", None)        # quantize2 flavour: (zero_point, range) device tensors of its quantizer
        want_y = need_bwd and not stem_bwd and not native_bwd
        pixsum = None
        if (conv and FUSE_QUANT_ROWSUM and q2 is None and not want_y and not nchw_stem and mod.groups == 1 and x.shape[1] % 16 == 0
                and x.is_contiguous(memory_format=torch.channels_last) and x.data_ptr() % 32 == 0
                and _abi.fakequant_affine_rowsum_supported(x.shape[1], bits_a)):
            # the per-pixel code sums of the zero-point correction come out of the quantizer's registers
            xq = None
            pixsum = _abi.fakequant_affine_rowsum(x, x.numel() // x.shape[1], x.shape[1], bits_a, rng, codes, x16)
        elif nchw_stem:
            if q2 is not None or (need_bwd and not stem_bwd):
                raise _abi.QB200Error("first-layer NCHW path entered with a backward that needs x_hat")
            xq, codes = None, _abi.fakequant_affine_nchw3(x, bits_a, rng)
        elif q2 is not None:
            xq = _abi.fakequant2_codes(x, q2[0], q2[1], bits_a, codes=codes, codes16=x16,
                                       want_y=need_bwd and not stem_bwd and not native_bwd)
        else:
            xq = _abi.fakequant_affine(x, bits_a, range_dev=rng, codes=codes, codes16=x16,
                                       want_y=need_bwd and not stem_bwd and not native_bwd)
        w4 = weight if conv else weight.view(weight.shape[0], weight.shape[1], 1, 1)
        wt_pre, pre = None, mod.__dict__.pop("_prepped", None)
        if pre is not None and pre[0] == (weight.data_ptr(), weight._version) and (pre[1][0] is not None or not need_bwd):
            torch.cuda.current_stream().wait_event(pre[2])                 # prepared aside by prepare_weights(model)
            wq, wcodes, tapsum, wt_pre = pre[1]
        else:
            wq, wcodes, tapsum = _abi.fakequant_rows_conv(w4, bits_w, mod.weight_min, mod.weight_max, compute_range=update,
                                                          want_y=need_bwd)
        if conv:
            B, Cin, H, W = x.shape
            kh, kw = mod.kernel_size
            if mod.groups != 1:                                             # depthwise: streaming kernel on the codes
                stats = torch.empty(8, dtype=torch.float64, device=x.device) if mod.training else None
                out = _abi.conv_dw_fwd(codes, wcodes, B, H, W, Cin, kh, kw, _pair(mod.stride)[0], _pair(mod.padding)[0],
                                       rng, bits_a, mod.weight_min, mod.weight_max, bits_w, bias=qbias, out_stats=stats)
                if stats is not None:
                    _stash_fwd_stats(out, stats)
            elif Cin % 16 != 0:                                             # first layer: thread-built im2col operand
                out = _abi.conv_stem_fwd(codes, wcodes, B, H, W, Cin, mod.out_channels, kh, kw, _pair(mod.stride)[0],
                                         _pair(mod.padding)[0], rng, bits_a, mod.weight_min, mod.weight_max, bits_w, tapsum,
                                         bias=qbias)
            else:
                if pixsum is None:
                    pixsum = _abi.rowsum_u8(codes, B * H * W, Cin)
                stats = torch.empty(8, dtype=torch.float64, device=x.device) if mod.training else None
                out = _abi.conv_i8_fwd(codes, wcodes, B, H, W, Cin, mod.out_channels, kh, kw, _pair(mod.stride)[0],
                                       _pair(mod.padding)[0], rng, bits_a, mod.weight_min, mod.weight_max, bits_w, pixsum,
                                       tapsum, bias=qbias, out_stats=stats)
                if stats is not None:
                    _stash_fwd_stats(out, stats)
            ctx.conv = (_pair(mod.stride), _pair(mod.padding), _pair(mod.dilation), mod.groups)
        else:
            M, K = x.shape
            N = mod.out_features
            rowsum = _abi.rowsum_u8(codes, M, K)
            out = _abi.gemm_i8(codes, wcodes, M, N, K, rng, bits_a, mod.weight_min, mod.weight_max, bits_w, rowsum,
                               tapsum, bias=qbias, b_offset=1 << (bits_w - 1))
            wq = wq.view(N, K) if wq is not None else None
            ctx.conv = None
        ctx.bits_grad, ctx.grad_mode, ctx.has_bias = mod.num_bits_grad, mod._grad_mode(), qbias is not None
        ctx.grad_quantizer = getattr(mod, "_grad_quantizer", None)
        ctx.stem = ctx.native = None
        # residual block: the skip gradient of this layer's input is handed over by the fused RangeBN+add+ReLU node of the
        # block (RangeBN.forward_relu(skip=...)) and the input-gradient kernel accumulates into it
        holder, ctx.skip = getattr(mod, "_skip_holder", None), None
        if holder is not None:
            mod._skip_holder = None
            if native_bwd and conv and ctx.needs_input_grad[0]:
                ctx.skip = holder
                holder["armed"] = True
        # ... and the downsample convolution of a block (same input) leaves ITS input gradient in the holder instead of returning
        # it to autograd, so that the first convolution accumulates into it
        holder, ctx.skip_out = getattr(mod, "_skip_out_holder", None), None
        if holder is not None:
            mod._skip_out_holder = None
            if holder.get("armed") and native_bwd and conv and ctx.needs_input_grad[0]:
                ctx.skip_out = holder
        if need_bwd and stem_bwd:
            ctx.stem = (bits_a, mod.out_channels, kh, kw, _pair(mod.stride)[0], _pair(mod.padding)[0])
            ctx.save_for_backward(codes, rng)
        elif native_bwd:
            n_out = wq.shape[0]
            pad = (-n_out) % 8                                              # linear layers: width padded to 16-byte rows
            wt = None
            if ctx.needs_input_grad[0]:
                wt = wt_pre if wt_pre is not None else _pack_wt(wq)
            ctx.native = (bits_a, tuple(x.shape), tuple(w4.shape), pad)
            ctx.save_for_backward(x16, rng, wt)
        elif need_bwd:
            ctx.save_for_backward(xq, wq)
        mod.__dict__["_tc_last"] = (need_bwd, bool(native_bwd and ctx.needs_input_grad[0]))    # what prepare_weights readies next
        return out

    @staticmethod
    def backward(ctx, g):
        if ctx.stem is not None:
            codes, rng = ctx.saved_tensors
            g = _abi._f32c(g, "grad")
            if ctx.grad_mode == "quant":
                g = _quantize_gradient(g, ctx.bits_grad)
            g = g.contiguous(memory_format=torch.channels_last)
            dw = None
            if ctx.needs_input_grad[1]:
                bits_a, cout, kh, kw, st, pd = ctx.stem
                B, _, H, W = codes.shape
                if USE_NATIVE_BACKWARD and kh == kw and _abi.conv_stem_wgrad16_supported(B, H, W, cout, kh, st, pd):
                    _, hi, lo = _abi.bwd_pack(g, want_codes=False)
                    dw = _abi.conv_stem_wgrad16(hi, lo, codes, rng, bits_a, cout, kh, st, pd)
                else:
                    dw = _abi.conv_stem_wgrad(g, codes, rng, *ctx.stem)
            db = g.sum((0, 2, 3)) if ctx.has_bias and ctx.needs_input_grad[2] else None
            return None, dw, db, None, None
        if ctx.native is not None:
            return _native_backward(ctx, g) + (None, None)
        x, w = ctx.saved_tensors
        return _contraction_backward(ctx, g, x, w) + (None, None)


def _native_backward(ctx, g):
    """dX, dW, db of the dense contraction on conv_bwd.cu, for the three gradient modes of _QContraction."""
    x16, a_rng, wt = ctx.saved_tensors
    bits_a, x_shape, w_shape, pad = ctx.native
    need_x, need_w, need_b = ctx.needs_input_grad[0], ctx.needs_input_grad[1], ctx.has_bias and ctx.needs_input_grad[2]
    conv = ctx.conv is not None
    n_out, c_in, kh, kw = w_shape
    stride, padding = (ctx.conv[0][0], ctx.conv[1][0]) if conv else (1, 0)
    g = _abi._f32c(g, "grad")
    if conv:
        g = g.contiguous(memory_format=torch.channels_last)
    elif not g.is_contiguous():
        g = g.contiguous()
    mode, bits = ctx.grad_mode, ctx.bits_grad
    scale_range = None
    a0 = a1 = None                      # planes of the input-gradient branch
    if mode == "biprec" and ctx.grad_quantizer is None and bits <= 8:
        scale_range, g, a0, hi, lo = _pack_gradient(g, bits, need_x, need_w or need_b, pad_cols=pad)
        g_param = g
    else:
        if mode == "full":
            g_in = g_param = g
        else:
            gq = ctx.grad_quantizer(g) if ctx.grad_quantizer is not None else _quantize_gradient(g, bits)
            g_in, g_param = gq, (gq if mode == "quant" else g)
        same = g_in is g_param
        if pad:
            g_in = F.pad(g_in, (0, pad))
            g_param = g_in if same else F.pad(g_param, (0, pad))
        hi = lo = None
        if need_w or need_b:
            _, hi, lo = _abi.bwd_pack(g_param, want_codes=False)
        if need_x:
            if g_in is g_param and hi is not None:
                a0, a1 = hi, lo
            else:
                _, a0, a1 = _abi.bwd_pack(g_in, want_codes=False)
    shape4 = x_shape if conv else (x_shape[0], x_shape[1], 1, 1)
    dx = dw = db = None
    if need_x:
        a0v = a0 if conv else a0.view(a0.shape[0], a0.shape[1], 1, 1)
        a1v = a1 if (a1 is None or conv) else a1.view(a1.shape[0], a1.shape[1], 1, 1)
        skip = getattr(ctx, "skip", None)
        acc = skip.pop("g", None) if skip is not None else None
        dx = _abi.conv_dgrad(a0v, a1v, wt, shape4, n_out + pad, kh, kw, stride, padding, scale_range=scale_range,
                             scale_bits=bits if scale_range is not None else 8, accumulate_into=acc)
        if not conv:
            dx = dx.permute(0, 2, 3, 1).reshape(x_shape)
        if getattr(ctx, "skip_out", None) is not None:       # handed to the block's first convolution (runs later in backward)
            ctx.skip_out["g"] = dx
            dx = None
    if need_b and not need_w:             # frozen weight, trainable bias: a plain column sum
        db = g_param.sum((0, 2, 3) if conv else 0)[:n_out]
    elif need_w:
        v4 = (lambda t: t) if conv else (lambda t: t.view(t.shape[0], t.shape[1], 1, 1))
        dw, db = _abi.conv_wgrad(v4(hi), v4(lo), g_param, v4(x16), n_out + pad, kh, kw, stride, padding, a_rng, bits_a,
                                 want_bias=bool(need_b))
        if pad:
            dw = dw[:n_out]
            db = db[:n_out] if db is not None else None
        if not conv:
            dw = dw.reshape(n_out, c_in)
    return dx, dw, db


class QConv2d(nn.Conv2d, _QBase):
    """quantize.py:271-354."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1, bias=True,
                 num_bits=8, num_bits_weight=None, num_bits_grad=None, biprecision=False, bias_quant=True,
                 per_channel=True, **kwargs):
        super(QConv2d, self).__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, bias,
                                      **kwargs)
        self._init_quant(num_bits, num_bits_weight, num_bits_grad, biprecision, bias_quant, per_channel, bias)

    def forward(self, input):
        if (CHANNELS_LAST and input.dim() == 4 and not input.is_contiguous(memory_format=torch.channels_last)
                and not _stem_takes_nchw(self, input)):
            input = input.contiguous(memory_format=torch.channels_last)      # library-native activation layout
        if self.enable_quant and _conv_i8_supported(self, input):
            rng = self.quantize_input.measure(input)
            return _QTensorCoreFn.apply(input, self.weight, self._quantized_bias(), self, rng)
        input_ = self.quantize_input(input)
        conv = (_pair(self.stride), _pair(self.padding), _pair(self.dilation), self.groups)
        if not self.enable_quant:
            # measure mode / set_quant_mode(False): the float layer (:351-352), in true fp32 on the native direct kernels
            if input.is_cuda and input.dtype == torch.float32 and self.padding_mode == 'zeros' and not ALLOW_LIBRARY_FALLBACK:
                return _QContraction.apply(input, self.weight, self.bias, conv, None, "full")
            return F.conv2d(input, self.weight, self.bias, self.stride, self.padding, self.dilation, self.groups)
        qweight, qbias = self._quantized_params()
        return _QContraction.apply(input_, qweight, qbias, conv, self.num_bits_grad, self._grad_mode())


class QLinear(nn.Linear, _QBase):
    """quantize.py:357-432."""

    def __init__(self, in_features, out_features, bias=True, num_bits=8, num_bits_weight=None, num_bits_grad=None,
                 biprecision=False, bias_quant=True, per_channel=True, **kwargs):
        super(QLinear, self).__init__(in_features, out_features, bias, **kwargs)
        self._init_quant(num_bits, num_bits_weight, num_bits_grad, biprecision, bias_quant, per_channel, bias)

    def forward(self, input):
        if self.enable_quant and _linear_i8_supported(self, input):
            rng = self.quantize_input.measure(input)
            return _QTensorCoreFn.apply(input, self.weight, self._quantized_bias(), self, rng)
        input_ = self.quantize_input(input)
        if not self.enable_quant:
            if input.is_cuda and input.dtype == torch.float32 and not ALLOW_LIBRARY_FALLBACK:
                return _QContraction.apply(input, self.weight, self.bias, None, None, "full")
            return F.linear(input, self.weight, self.bias)
        qweight, qbias = self._quantized_params()
        return _QContraction.apply(input_, qweight, qbias, None, self.num_bits_grad, self._grad_mode())


# ---------------------------------------------------------------------------------------------------
# RangeBN                                                                    (quantize.py:435-505)
# ---------------------------------------------------------------------------------------------------
class _RangeBNFn(Function):
    """Statistics + normalisation of the fake-quantized input and, in backward, the output-gradient quantizer
    followed by the closed-form gradient of :467-499 (rangebn.cu / rangebn_nhwc.cu).

    Channels-last fusion (``q_range`` given): ``x`` is the RAW input and QuantMeasure's fake-quant (:463) runs inside
    the statistics / apply / backward-reduce kernels, so x_hat never exists in HBM; with C % 8 == 0 the gradient
    quantizer runs inside the two backward passes as well and the dx range is handed to the preceding QConv2d."""

    @staticmethod
    def forward(ctx, xq, gamma, beta, running_mean, running_var, training, momentum, num_chunks, eps, bits, bits_grad,
                q_range, q_bits, deferred=None):
        xq = _abi._f32c(xq, "x")
        nhwc = _abi.is_nhwc(xq)
        B, Cn = xq.shape[0], xq.shape[1]
        HW = xq.numel() // (B * Cn)
        count, chunk_len, chunks_global = float(B * HW), 0.0, num_chunks
        red = amax = amin = xc = vec = None
        ctx.local_chunks = num_chunks
        if training:
            local_chunks = num_chunks
            if _reducer is not None:
                local_chunks, count = _reducer.rangebn_layout(B, HW, num_chunks)
            want_arg = any(ctx.needs_input_grad[:3])
            if (nhwc and q_range is not None and q_bits <= 8 and Cn % 8 == 0 and xq.numel() < 2 ** 32
                    and xq.data_ptr() % 32 == 0):
                # x_hat leaves the statistics pass as uint8 codes: apply and backward read 1 B/elem, the raw conv output
                # is not kept for backward
                xc = torch.empty(xq.numel(), dtype=torch.uint8, device=xq.device)
            chunk_len = count / chunks_global
            ctx.local_chunks = local_chunks
            if nhwc and _reducer is None and FUSE_RANGEBN_FINALIZE and chunk_len > 1:
                # one process: nothing to exchange between the statistics and the per-channel finalize - one call, and the
                # finalize rides at the tail of the statistics pass' last kernel
                vec, amax, amin = _abi.rangebn_stats_finalize_nhwc(xq, B, Cn, HW, num_chunks, eps, momentum, bits, gamma, beta,
                                                                   running_mean, running_var, want_arg=want_arg,
                                                                   q_range=q_range, q_bits=q_bits, x_codes_out=xc)
            else:
                red, amax, amin = _abi.rangebn_stats(xq, B, Cn, HW, local_chunks, want_arg=want_arg, nhwc=nhwc,
                                                     q_range=q_range, q_bits=q_bits, x_codes_out=xc)
                if _reducer is not None:
                    _reducer.rangebn(red)
        if vec is None:
            vec = _abi.rangebn_finalize(red, Cn, chunks_global, count, chunk_len, eps, momentum, training, bits, gamma, beta,
                                        running_mean, running_var)
        ctx.dims = (B, Cn, HW, chunks_global, count, chunk_len, training, bits_grad, gamma is not None, beta is not None)
        ctx.nhwc, ctx.q_bits = nhwc, q_bits
        ctx.save_for_backward(None if xc is not None else xq, vec, amax, amin, q_range, xc)
        ctx.deferred = deferred
        if deferred is not None:
            # RangeBN.forward_relu: the normalisation runs later, fused with the ReLU (_ApplyReLUFn); this node only owns the
            # statistics and the RangeBN backward.  Its output is a storage-less placeholder of the right shape.
            if xc is None:
                raise _abi.QB200Error("RangeBN.forward_relu: the fused pass needs the uint8 codes of the statistics pass")
            deferred.update(xc=xc, vec=vec, like=xq, dims=(B, Cn, HW), has_gamma=gamma is not None, has_beta=beta is not None,
                            q_range=q_range, q_bits=q_bits, bits_grad=bits_grad)
            return xq.new_empty(1).expand(xq.shape)
        return _abi.rangebn_apply(xq, B, Cn, HW, vec, gamma is not None, beta is not None, nhwc=nhwc, q_range=q_range,
                                  q_bits=q_bits, x_codes=xc)

    @staticmethod
    def backward(ctx, g):
        xq, vec, amax, amin, q_range, xc = ctx.saved_tensors
        B, Cn, HW, K, count, chunk_len, training, bits_grad, has_gamma, has_beta = ctx.dims
        g = g.contiguous(memory_format=torch.channels_last) if ctx.nhwc else g.contiguous()
        hook = _reducer.rangebn_bwd if (_reducer is not None and training) else None
        common = dict(reduce_hook=hook, chunks_global=K, nhwc=ctx.nhwc, q_range=q_range, q_bits=ctx.q_bits, x_codes=xc)
        lc = ctx.local_chunks if training else 0
        # _ApplyReLUFn left the backward of the fused ReLU to this node: g is still the gradient w.r.t. the ReLU OUTPUT
        relu_bits = ctx.deferred.pop("relu_bits", None) if ctx.deferred is not None else None
        fused = (ctx.nhwc and Cn % 8 == 0 and g.numel() < 2 ** 32 and g.data_ptr() % 32 == 0
                 and (xc is not None or xq.data_ptr() % 32 == 0))
        if relu_bits is not None and not (fused and bits_grad is not None and bits_grad <= 8):
            g, rr = _abi.relu_bwd_minmax_bits(_abi._f32c(g, "grad"), relu_bits)
            _stash_grad_range(g, rr)
            relu_bits = None
        if fused:
            gquant = None
            if bits_grad is not None:
                g = _abi._f32c(g, "grad")
                if relu_bits is not None:                    # range of the MASKED gradient; the reduce pass masks on the fly
                    r = _abi.relu_bwd_minmax_bits(g, relu_bits, want_dx=False)[1]
                    common["relu_bits"] = relu_bits
                else:
                    r = _take_grad_range(g)                  # left by the ReLU backward that produced g (glue.py)
                if r is None:
                    r = _abi.minmax(g)
                if _reducer is not None:
                    _reducer.grad_range(r)
                gquant = dict(range=r, bits=bits_grad, noise=_grad_noise(g, True), stochastic=True, seed=_next_seed(),
                              step_dev=noise_step(g.device))
            dx, dgamma, dbeta, dx_range = _abi.rangebn_backward(g, xq, B, Cn, HW, lc, count, chunk_len, training, vec,
                                                                has_gamma, amax, amin, grad_quant=gquant,
                                                                want_dx_range=True, **common)
            _stash_grad_range(dx, dx_range)
        else:
            gq = _quantize_gradient(g, bits_grad) if bits_grad is not None else g
            dx, dgamma, dbeta = _abi.rangebn_backward(gq, xq, B, Cn, HW, lc, count, chunk_len, training, vec, has_gamma,
                                                      amax, amin, **common)
        return (dx if ctx.needs_input_grad[0] else None, dgamma if has_gamma else None, dbeta if has_beta else None,
                None, None, None, None, None, None, None, None, None, None, None)


class _ApplyReLUFn(Function):
    """Second half of RangeBN.forward_relu: normalisation from the codes [+ residual] + ReLU + the statistics of the result
    in one pass (qb_rangebn_apply_relu_measure).  ``h`` is the placeholder output of the _RangeBNFn node that ran the
    statistics; the masked gradient goes back through it (and to the residual).  Keeping the RangeBN backward in its own
    node - created where the reference calls ``bn(x)`` - keeps autograd's execution order, hence the order of the
    stochastic-rounding noise draws, that of the unfused graph."""

    @staticmethod
    def forward(ctx, h, residual, d):
        from . import glue
        B, Cn, HW = d["dims"]
        out, glue._last["stats"], bits = _abi.rangebn_apply_relu_measure(d["xc"], d["like"], B, Cn, HW, d["vec"], d["has_gamma"],
                                                                         d["has_beta"], d["q_range"], d["q_bits"],
                                                                         residual=residual, want_bits=True)
        # the ReLU backward needs one bit per element, not the fp32 output
        ctx.has_bits = bits is not None
        ctx.save_for_backward(bits if ctx.has_bits else out)
        ctx.d = d
        ctx.has_residual = residual is not None
        return out

    @staticmethod
    def backward(ctx, g):
        from . import glue
        (saved,) = ctx.saved_tensors
        g = _abi._f32c(g, "grad").contiguous(memory_format=torch.channels_last)
        if not ctx.has_bits:
            dg = glue._relu_backward(g, saved)
        elif not ctx.has_residual and g.data_ptr() % 32 == 0:
            ctx.d["relu_bits"] = saved                      # the RangeBN node that follows masks on the fly (no dx tensor)
            return g, None, None
        else:
            dg, rng = _abi.relu_bwd_minmax_bits(g, saved)
            _stash_grad_range(dg, rng)
        skip = ctx.d.get("skip")
        if skip is not None:                                # the block's first convolution adds its dX into this tensor
            skip["g"] = dg
        return dg, (dg if ctx.needs_input_grad[1] else None), None


class _ApplyPoolFn(Function):
    """Second half of RangeBN.forward_relu_pool: max-pooling straight from the codes of the statistics pass (the normalised
    tensor is never written), then the ReLU and the next QuantMeasure's statistics on the 4x smaller pooled tensor - the
    relu(maxpool(.)) evaluation of glue.MaxPool2d.relu_pool.  Backward: the one-pass gather of qb_maxpool2d_relu_nhwc_bwd
    (gradient w.r.t. the RangeBN output with its range), handed to the _RangeBNFn node through the placeholder ``h``."""

    @staticmethod
    def forward(ctx, h, d, k, stride, pad):
        from . import glue
        B, Cn, HW = d["dims"]
        shape = tuple(d["like"].shape)
        out, idx = _abi.rangebn_apply_maxpool(d["xc"], shape, d["vec"], d["has_gamma"], d["has_beta"], d["q_range"], d["q_bits"],
                                              k, stride, pad)
        out, stats = _abi.relu_measure(out)
        glue._last["stats"] = stats
        ctx.save_for_backward(idx, out)
        ctx.geom = (shape, k, stride, pad)
        return out

    @staticmethod
    def backward(ctx, g):
        idx, out = ctx.saved_tensors
        shape, k, stride, pad = ctx.geom
        g = _abi._f32c(g, "grad").contiguous(memory_format=torch.channels_last)
        dx, rng = _abi.maxpool2d_relu_nhwc_bwd(g, idx, out, shape, k, stride, pad)
        _stash_grad_range(dx, rng)
        return dx, None, None, None, None


class RangeBN(nn.Module):
    """quantize.py:435-505 ("normalized RangeBN")."""

    def __init__(self, num_features, dim=1, momentum=0.1, affine=True, num_chunks=16, eps=1e-5, num_bits=8,
                 num_bits_grad=8):
        super(RangeBN, self).__init__()
        self.register_buffer('running_mean', torch.zeros(num_features))
        self.register_buffer('running_var', torch.zeros(num_features))
        self.momentum = momentum
        self.dim = dim
        if affine:
            self.bias = nn.Parameter(torch.Tensor(num_features))
            self.weight = nn.Parameter(torch.Tensor(num_features))
        else:
            self.register_parameter('bias', None)
            self.register_parameter('weight', None)
        self.num_bits = num_bits
        self.num_bits_grad = num_bits_grad
        self.quantize_input = QuantMeasure(self.num_bits)
        self.eps = eps
        self.num_chunks = num_chunks
        self.reset_params()

    def reset_params(self):
        if self.weight is not None:
            self.weight.data.uniform_()
        if self.bias is not None:
            self.bias.data.zero_()

    def forward(self, x):
        two_d = x.dim() == 2
        if two_d:
            x = x.unsqueeze(-1).unsqueeze(-1)
        qi = self.quantize_input
        q_range, q_bits = None, 0
        if x.dim() == 4 and x.is_cuda and x.dtype == torch.float32 and _abi.is_nhwc(x) and qi.enable_quant:
            q_range, q_bits = qi.measure(x), qi.num_bits     # the fake-quant itself runs inside the RangeBN kernels
        else:
            x = qi(x)
        out = _RangeBNFn.apply(x, self.weight, self.bias, self.running_mean, self.running_var, self.training,
                               self.momentum, self.num_chunks, self.eps, self.num_bits, self.num_bits_grad, q_range, q_bits)
        if two_d:
            out = out.squeeze(-1).squeeze(-1)
        return out

    def forward_relu(self, x, relu, residual=None, skip=None):
        """``relu(self(x))`` - or, with ``residual``, ``out = self(x); out += residual; relu(out)``, the tail of a residual
        block (models/resnet_quantized.py:62-68, 105-112) - with the normalisation, the add, the clamp and the statistics of
        the next QuantMeasure in ONE pass over the uint8 codes (qb_rangebn_apply_relu_measure): the normalised tensor is
        never written.  ``residual`` may be a callable (the block's downsample branch): it is evaluated AFTER the
        statistics of this module, where the reference evaluates it.  Same values and gradients as the composition, which is
        what runs whenever the fused pass does not apply."""
        from . import glue
        qi = self.quantize_input
        ok = (isinstance(relu, glue.ReLU) and relu.training and self.training and x.dim() == 4 and x.is_cuda
              and x.dtype == torch.float32 and _abi.is_nhwc(x) and qi.enable_quant and qi.num_bits <= 8 and x.shape[1] % 8 == 0
              and x.numel() < 2 ** 32 and x.data_ptr() % 32 == 0 and torch.is_grad_enabled())
        if not ok:
            out = self(x)
            if residual is None:
                return relu(out)
            res = residual() if callable(residual) else residual
            if hasattr(relu, "add_relu"):
                return relu.add_relu(out, res)
            out += res
            return relu(out)
        d = {}
        h = _RangeBNFn.apply(x, self.weight, self.bias, self.running_mean, self.running_var, self.training, self.momentum,
                             self.num_chunks, self.eps, self.num_bits, self.num_bits_grad, qi.measure(x), qi.num_bits, d)
        res = residual() if callable(residual) else residual
        if res is not None:
            res = res.contiguous(memory_format=torch.channels_last)
            # ``skip``: the holder the block handed to its FIRST convolution, whose input IS this residual; if that layer
            # armed it, the skip gradient does not travel through autograd (no separate accumulation pass): the masked
            # gradient is left in the holder and the convolution's input-gradient kernel accumulates into it
            if skip is not None and skip.get("armed") and res.requires_grad:
                d["skip"] = skip
                res = res.detach()
        out = _ApplyReLUFn.apply(h, res, d)
        stats, glue._last["stats"] = glue._last["stats"], None
        if stats is not None:
            _stash_fwd_stats(out, stats)
        if out.requires_grad:
            glue._relu_out["t"] = (out, out._version)
        return out


    def forward_relu_pool(self, x, relu, pool):
        """``pool(relu(self(x)))`` - the stem tail conv1 -> bn1 -> relu -> maxpool of models/resnet_quantized.py:147-150 - with
        the pooling reading the uint8 codes of this module's statistics pass and normalising the nine window elements on the
        fly (qb_rangebn_apply_maxpool_nhwc): the full-size normalised tensor (822 MB for ResNet-18 at batch 256) is neither
        written nor read.  Same values and gradients as ``pool.relu_pool(self(x), relu)``, which runs whenever this does not apply."""
        from . import glue
        qi = self.quantize_input
        k, s, p = pool.kernel_size, pool.stride, pool.padding
        # (single process only for now: the multi-rank equivalence tests last ran before this route existed)
        ok = (FUSE_BN_POOL and _reducer is None and isinstance(relu, glue.ReLU) and isinstance(pool, glue.MaxPool2d) and relu.training
              and self.training
              and x.dim() == 4 and x.is_cuda and x.dtype == torch.float32 and _abi.is_nhwc(x) and qi.enable_quant
              and qi.num_bits <= 8 and x.shape[1] % 8 == 0 and x.numel() < 2 ** 32 and x.data_ptr() % 32 == 0
              and torch.is_grad_enabled() and x.requires_grad and pool.dilation == 1 and not pool.ceil_mode
              and not pool.return_indices and glue._native_ok(x, k, s, p))
        if not ok:
            return pool.relu_pool(self(x), relu)
        d = {}
        h = _RangeBNFn.apply(x, self.weight, self.bias, self.running_mean, self.running_var, self.training, self.momentum,
                             self.num_chunks, self.eps, self.num_bits, self.num_bits_grad, qi.measure(x), qi.num_bits, d)
        out = _ApplyPoolFn.apply(h, d, k, s, p)
        stats, glue._last["stats"] = glue._last["stats"], None
        if stats is not None:
            _stash_fwd_stats(out, stats)
        return out


# ---------------------------------------------------------------------------------------------------
# model-walk helpers                                                          (quantize.py:508-610)
# ---------------------------------------------------------------------------------------------------
def is_bn(m):
    return isinstance(m, nn.BatchNorm2d) or isinstance(m, nn.BatchNorm1d)


def is_quant(m):
    return isinstance(m, QuantNode) or isinstance(m, QLinear) or isinstance(m, QConv2d)


def recursive_apply(model, func, *args):
    for m in model.children():
        func(m, *args)
        recursive_apply(m, func, *args)


def set_bn_is_train(model, train, logger=None, reload_running_estimators=False, reset_running_estimators=False,
                    freeze_affine=False):
    """quantize.py:523-544."""
    def func(m, *args):
        if not is_bn(m):
            return
        if reload_running_estimators or reset_running_estimators:
            if (reload_running_estimators and not hasattr(m, 'locked_running_mean')) or reset_running_estimators:
                m.locked_running_mean = m.running_mean.data.clone()
                m.locked_running_var = m.running_var.data.clone()
            else:
                m.running_mean.data = m.locked_running_mean.clone()
                m.running_var.data = m.locked_running_var.clone()
        if freeze_affine:
            for p in m.parameters():
                p.requires_grad = False
        m.train(train)

    recursive_apply(model, func)


def set_measure_mode(model, measure, momentum=None, logger=None):
    """quantize.py:547-558: calibration switch - quantizers pass through while ranges keep accumulating."""
    def func(m, *args):
        if is_bn(m):
            m.train(not measure)
        elif is_quant(m):
            m.set_measure_mode(measure, momentum=momentum)

    recursive_apply(model, func)


def set_quant_mode(model, quant, logger=None):
    def func(m, *args):
        if is_quant(m):
            m.enable_quant = quant

    recursive_apply(model, func)


def overwrite_params(model, logger=None):
    def func(m, *args):
        if is_quant(m):
            m.overwrite_params(logger)

    recursive_apply(model, func)


def freeze_quant_params(model, freeze=True, include_param_dyn_range=True, momentum='same', logger=None):
    """quantize.py:579-592, INCLUDING its misspelt attribute: the reference sets ``freeze_param_dyn_rang`` (sic), so
    weight ranges are in fact never frozen by this call.  Kept for parity."""
    def func(m, *args):
        if isinstance(m, QuantMeasure):
            m.train(not freeze)
            if momentum != 'same':
                m.momentum = momentum
        if include_param_dyn_range and isinstance(m, QuantNode):
            m.freeze_param_dyn_rang = freeze

    recursive_apply(model, func)


def distill_set_train(model, train):
    model.train(train)
    if train:
        freeze_quant_params(model)
        set_bn_is_train(model, False)


def set_global_quantization_method(model, method='aciq', logger=None):
    assert method in QuantMeasure._QMEASURE_SUPPORTED_METHODS

    def func(m, *args):
        if isinstance(m, QuantMeasure):
            m.method = method

    recursive_apply(model, func)
