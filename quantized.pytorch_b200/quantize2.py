"""Drop-in mirror of the reference's ``models/modules/quantize2.py`` on the B200 kernels.

Same names and signatures (file:line citations are relative to /root/reference/models/modules/quantize2.py).  QParams
are device tensors (the reference never calls ``float()`` in this flavour either).  Every O(n) pass - row / global
min-max, the quantizer itself, RangeBN statistics / normalisation / backward - is a kernel behind the C ABI
(fakequant2.cu, fakequant_grad.cu, rangebn*.cu); what stays in torch is O(batch)- or O(C)-sized glue on the device
(the mean of B per-sample minima, ``max - min``, the running-buffer blends), with no host synchronisation.
There is no CPU fallback: CPU tensors raise QB200Error."""
from collections import namedtuple

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.autograd.function import Function, InplaceFunction

from . import _abi
from . import quantize as _q1
from ._abi import QB200Error

QParams = namedtuple('QParams', ['range', 'zero_point', 'num_bits'])

_DEFAULT_FLATTEN = (1, -1)
_DEFAULT_FLATTEN_GRAD = (0, -1)

__all__ = ["QParams", "calculate_qparams", "UniformQuantize", "UniformQuantizeGrad", "conv2d_biprec", "linear_biprec",
           "quantize", "quantize_grad", "QuantNode", "QuantMeasure", "QConv2d", "QLinear", "RangeBN", "RangeBN1d"]


def _deflatten_as(x, x_full):
    return x.view(*(list(x.shape) + [1] * (x_full.dim() - x.dim())))


def _as_tuple(fd):
    return tuple(fd) if isinstance(fd, (tuple, list)) else (fd,)


def calculate_qparams(x, num_bits, flatten_dims=_DEFAULT_FLATTEN, reduce_dim=0, reduce_type='mean', keepdim=False,
                      true_zero=False):
    """quantize2.py:19-38.  min / max over the flattened trailing block run in qb_minmax (one group) or qb_row_minmax
    (one CTA per group); the reduction over ``reduce_dim`` acts on the resulting small tensor."""
    with torch.no_grad():
        xd = _abi._f32c(x.detach(), "x")
        x_flat = xd.flatten(*_as_tuple(flatten_dims))
        if x_flat.dim() == 1:
            r = _abi.minmax(x_flat.contiguous())
            min_values, max_values = _deflatten_as(r[0], xd), _deflatten_as(r[1], xd)
        else:
            lead = x_flat.shape[:-1]
            rows = x_flat.reshape(-1, x_flat.shape[-1])
            mn, mx = _abi.row_minmax(rows.contiguous())
            min_values, max_values = _deflatten_as(mn.view(lead), xd), _deflatten_as(mx.view(lead), xd)
        if reduce_dim is not None:
            if reduce_type == 'mean':
                min_values = min_values.mean(reduce_dim, keepdim=keepdim)
                max_values = max_values.mean(reduce_dim, keepdim=keepdim)
            else:
                min_values = min_values.min(reduce_dim, keepdim=keepdim)[0]
                max_values = max_values.max(reduce_dim, keepdim=keepdim)[0]
        return QParams(range=max_values - min_values, zero_point=min_values, num_bits=num_bits)


def _noise_for(x, stochastic):
    if stochastic and _q1._noise_provider is not None:
        n = _q1._noise_provider(tuple(x.shape))
        if n is not None and _abi.is_nhwc(x):
            n = n.reshape(x.shape).contiguous(memory_format=torch.channels_last)
        return n
    return None


def _apply_quantizer(x, qparams, dequantize, signed, stochastic, out=None):
    """One qb_fakequant2 launch.  qparams hold one value per tensor or one per row of dim 0."""
    zp, rng = qparams.zero_point, qparams.range
    n_q = zp.numel()
    if n_q == 1 and rng.numel() == 1:
        per_row = False
    elif n_q == x.shape[0] and rng.numel() == n_q and zp.shape[0] == n_q:
        per_row = True
        if _abi.is_nhwc(x):
            x = x.contiguous()                       # rows must be contiguous blocks
            out = None
    else:
        raise QB200Error("quantize2: qparams of shape %s do not map to one value per tensor or per row of dim 0 of %s"
                         % (tuple(zp.shape), tuple(x.shape)))
    return _abi.fakequant2(x, zp, rng, qparams.num_bits, per_row=per_row, signed=signed, dequantize=dequantize,
                           noise=_noise_for(x, stochastic), stochastic=bool(stochastic), seed=_q1._next_seed(),
                           step_dev=_q1.noise_step(x.device), out=out)


class UniformQuantize(InplaceFunction):
    """quantize2.py:41-82."""

    @staticmethod
    def forward(ctx, input, num_bits=None, qparams=None, flatten_dims=_DEFAULT_FLATTEN, reduce_dim=0, dequantize=True,
                signed=False, stochastic=False, inplace=False):
        ctx.inplace = inplace
        x = _abi._f32c(input, "input")
        out = None
        if inplace:
            ctx.mark_dirty(input)
            out = input if x is input else None
        if qparams is None:
            assert num_bits is not None, "either provide qparams of num_bits to quantize"
            qparams = calculate_qparams(input, num_bits=num_bits, flatten_dims=flatten_dims, reduce_dim=reduce_dim)
        y = _apply_quantizer(x, qparams, dequantize, signed, stochastic, out=out)
        if inplace and y is not input:
            input.copy_(y)
            y = input
        return y

    @staticmethod
    def backward(ctx, grad_output):
        return (grad_output,) + (None,) * 8          # straight-through estimator (:78-82)


def _quantize_grad_tensor(g, num_bits, qparams, flatten_dims, reduce_dim, signed, stochastic):
    """UniformQuantizeGrad.backward body (:99-112): 'extreme' ranges of the gradient, then the quantizer."""
    g = _abi._f32c(g, "grad")
    if qparams is None:
        assert num_bits is not None, "either provide qparams of num_bits to quantize"
        qparams = calculate_qparams(g, num_bits=num_bits, flatten_dims=flatten_dims, reduce_dim=reduce_dim,
                                    reduce_type='extreme')
    return _apply_quantizer(g, qparams, True, signed, stochastic)


class UniformQuantizeGrad(InplaceFunction):
    """quantize2.py:85-112: identity forward, quantized gradient backward."""

    @staticmethod
    def forward(ctx, input, num_bits=None, qparams=None, flatten_dims=_DEFAULT_FLATTEN_GRAD, reduce_dim=0,
                dequantize=True, signed=False, stochastic=True):
        ctx.num_bits, ctx.qparams, ctx.flatten_dims = num_bits, qparams, flatten_dims
        ctx.stochastic, ctx.signed, ctx.dequantize, ctx.reduce_dim = stochastic, signed, dequantize, reduce_dim
        ctx.inplace = False
        return input.clone()        # the reference returns `input` itself (:96); see quantize.UniformQuantizeGrad

    @staticmethod
    def backward(ctx, grad_output):
        with torch.no_grad():
            grad_input = _quantize_grad_tensor(grad_output, ctx.num_bits, ctx.qparams, ctx.flatten_dims, ctx.reduce_dim,
                                               ctx.signed, ctx.stochastic)
        return grad_input, None, None, None, None, None, None, None


def quantize(x, num_bits=None, qparams=None, flatten_dims=_DEFAULT_FLATTEN, reduce_dim=0, dequantize=True, signed=False,
             stochastic=False, inplace=False):
    return UniformQuantize.apply(x, num_bits, qparams, flatten_dims, reduce_dim, dequantize, signed, stochastic, inplace)


def quantize_grad(x, num_bits=None, qparams=None, flatten_dims=_DEFAULT_FLATTEN_GRAD, reduce_dim=0, dequantize=True,
                  signed=False, stochastic=True):
    return UniformQuantizeGrad.apply(x, num_bits, qparams, flatten_dims, reduce_dim, dequantize, signed, stochastic)


def _grad_quantizer(num_bits, flatten_dims):
    return lambda g: _quantize_grad_tensor(g, num_bits, None, flatten_dims, 0, False, True)


def conv2d_biprec(input, weight, bias=None, stride=1, padding=0, dilation=1, groups=1, num_bits_grad=None):
    """quantize2.py:115-121: dW, db from the unquantized gradient, dX from Q(g) (per-sample flatten, extreme ranges).
    The reference evaluates the convolution twice with identical values; once is enough (see quantize._QContraction)."""
    conv = (_q1._pair(stride), _q1._pair(padding), _q1._pair(dilation), groups)
    return _q1._QContraction.apply(input, weight, bias, conv, num_bits_grad, "biprec", _grad_quantizer(num_bits_grad, (1, -1)))


def linear_biprec(input, weight, bias=None, num_bits_grad=None):
    """quantize2.py:124-129."""
    return _q1._QContraction.apply(input, weight, bias, None, num_bits_grad, "biprec",
                                   _grad_quantizer(num_bits_grad, _DEFAULT_FLATTEN_GRAD))


class QuantNode():
    """quantize2.py:139-154."""

    def __init__(self):
        self.enable_quant = True
        self.measure = False

    def set_measure_mode(self, measure, momentum=None):
        self.enable_quant = not measure
        self.measure = measure
        if momentum and isinstance(self, QuantMeasure):
            self.momentum = momentum
        elif isinstance(self, nn.Module):
            for m in self.children():
                if isinstance(m, QuantNode):
                    m.set_measure_mode(measure, momentum=momentum)


class QuantMeasure(nn.Module, QuantNode):
    """quantize2.py:157-199: running (zero_point, range) tracker + activation fake-quant."""

    def __init__(self, num_bits=8, shape_measure=(1,), flatten_dims=_DEFAULT_FLATTEN, inplace=False, dequantize=True,
                 stochastic=False, momentum=0.1, measure=False):
        super(QuantMeasure, self).__init__()
        QuantNode.__init__(self)
        self.register_buffer('running_zero_point', torch.zeros(*shape_measure))
        self.register_buffer('running_range', torch.zeros(*shape_measure))
        self.register_buffer('num_measured', torch.zeros(1))
        self.measure = measure
        self.flatten_dims = flatten_dims
        self.momentum = momentum
        self.dequantize = dequantize
        self.stochastic = stochastic
        self.inplace = inplace
        self.num_bits = num_bits

    def forward(self, input, qparams=None):
        if self.training or self.measure:
            if qparams is None:
                qparams = calculate_qparams(input, num_bits=self.num_bits, flatten_dims=self.flatten_dims, reduce_dim=0)
            with torch.no_grad():
                if self.measure:
                    momentum = self.num_measured / (self.num_measured + 1)
                    self.num_measured += 1
                else:
                    momentum = self.momentum
                self.running_zero_point.mul_(momentum).add_(qparams.zero_point * (1 - momentum))
                self.running_range.mul_(momentum).add_(qparams.range * (1 - momentum))
        else:
            qparams = QParams(range=self.running_range, zero_point=self.running_zero_point, num_bits=self.num_bits)
        if self.measure:
            return input
        return quantize(input, qparams=qparams, dequantize=self.dequantize, stochastic=self.stochastic, inplace=self.inplace)


def _tc_route(mod, input):
    """quantize2.QConv2d / QLinear on the native tensor-core contractions (row A9): possible when the input quantizer is the
    plain per-tensor unsigned one (its codes are then the int8 operand) and the geometry is one quantize.py's kernels take."""
    qi = mod.quantize_input
    if not (_q1.USE_INT8_TENSOR_CORES and input.is_cuda and input.dtype == torch.float32 and qi.dequantize and not qi.stochastic
            and not qi.measure and qi.num_bits <= 8 and mod.num_bits_weight <= 8 and tuple(qi.running_range.shape) in ((1, 1, 1, 1), (1,))
            and not getattr(qi, "inplace", False)):
        return False
    if isinstance(mod, nn.Conv2d):
        return input.dim() == 4 and mod.groups == 1 and _q1._conv_i8_supported(mod, input) and mod.in_channels % 16 == 0
    return input.dim() == 2 and mod.in_features % 16 == 0


def _tc_forward(mod, input, conv):
    """The input quantizer's bookkeeping (QuantMeasure.forward, quantize2.py:176-196), then quantize._QTensorCoreFn on its
    (zero_point, range): activation codes by qb_fakequant2_codes, weights per row (quantize2.py:225-226 is the same affine
    quantizer as quantize.py:318-334), int8 tcgen05 forward, native backward with this flavour's gradient quantizer."""
    qi = mod.quantize_input
    x = input
    if conv is not None and _q1.CHANNELS_LAST and not x.is_contiguous(memory_format=torch.channels_last):
        x = x.contiguous(memory_format=torch.channels_last)
    if qi.training:
        qp = calculate_qparams(x, num_bits=qi.num_bits, flatten_dims=qi.flatten_dims, reduce_dim=0)
        with torch.no_grad():
            qi.running_zero_point.mul_(qi.momentum).add_(qp.zero_point * (1 - qi.momentum))
            qi.running_range.mul_(qi.momentum).add_(qp.range * (1 - qi.momentum))
        zp, rg = qp.zero_point, qp.range
    else:
        zp, rg = qi.running_zero_point, qi.running_range
    zp, rg = zp.detach().reshape(1).float(), rg.detach().reshape(1).float()
    qbias = None
    if mod.bias is not None:
        qbias = quantize(mod.bias, num_bits=mod.num_bits_weight + mod.num_bits, flatten_dims=(0, -1))
    n = mod.weight.shape[0]
    mod._q2_params = (zp, rg)
    mod.weight_min = torch.empty(n, dtype=torch.float32, device=x.device)      # scratch of the row kernel (not buffers here)
    mod.weight_max = torch.empty(n, dtype=torch.float32, device=x.device)
    fd = (1, -1) if conv is not None else _DEFAULT_FLATTEN_GRAD
    mod._grad_quantizer = _grad_quantizer(mod.num_bits_grad, fd) if mod.num_bits_grad is not None else None
    try:
        return _q1._QTensorCoreFn.apply(x, mod.weight, qbias, mod, torch.cat([zp, zp + rg]))
    finally:
        del mod._q2_params, mod.weight_min, mod.weight_max


def _q2_grad_mode(mod):
    if mod.num_bits_grad is None:
        return "full"
    return "biprec" if mod.biprecision else "quant"


def _contract(mod, qinput, qweight, qbias, conv):
    """Shared tail of QConv2d / QLinear forward (:225-234, :261-268)."""
    fd = (1, -1) if conv is not None else _DEFAULT_FLATTEN_GRAD
    if mod.num_bits_grad is None:
        mode, gq = "full", None
    else:
        mode, gq = ("biprec" if mod.biprecision else "quant"), _grad_quantizer(mod.num_bits_grad, fd)
    return _q1._QContraction.apply(qinput, qweight, qbias, conv, mod.num_bits_grad, mode, gq)


class QConv2d(nn.Conv2d):
    """quantize2.py:202-235."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1, bias=True,
                 num_bits=8, num_bits_weight=8, num_bits_grad=8, biprecision=True):
        super(QConv2d, self).__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, bias)
        self.num_bits = num_bits
        self.num_bits_weight = num_bits_weight or num_bits
        self.num_bits_grad = num_bits_grad
        self.quantize_input = QuantMeasure(self.num_bits, shape_measure=(1, 1, 1, 1), flatten_dims=(1, -1))
        self.biprecision = biprecision

    def _grad_mode(self):
        return _q2_grad_mode(self)

    def forward(self, input):
        conv = (_q1._pair(self.stride), _q1._pair(self.padding), _q1._pair(self.dilation), self.groups)
        if _tc_route(self, input):
            return _tc_forward(self, input, conv)
        qinput = self.quantize_input(input)
        weight_qparams = calculate_qparams(self.weight, num_bits=self.num_bits_weight, flatten_dims=(1, -1), reduce_dim=None)
        qweight = quantize(self.weight, qparams=weight_qparams)
        qbias = None
        if self.bias is not None:
            qbias = quantize(self.bias, num_bits=self.num_bits_weight + self.num_bits, flatten_dims=(0, -1))
        return _contract(self, qinput, qweight, qbias, conv)


class QLinear(nn.Linear):
    """quantize2.py:238-269."""

    def __init__(self, in_features, out_features, bias=True, num_bits=8, num_bits_weight=8, num_bits_grad=8,
                 biprecision=True):
        super(QLinear, self).__init__(in_features, out_features, bias)
        self.num_bits = num_bits
        self.num_bits_weight = num_bits_weight or num_bits
        self.num_bits_grad = num_bits_grad
        self.biprecision = biprecision
        self.quantize_input = QuantMeasure(self.num_bits)

    def _grad_mode(self):
        return _q2_grad_mode(self)

    def forward(self, input):
        if _tc_route(self, input):
            return _tc_forward(self, input, None)
        qinput = self.quantize_input(input)
        weight_qparams = calculate_qparams(self.weight, num_bits=self.num_bits_weight, flatten_dims=(1, -1), reduce_dim=None)
        qweight = quantize(self.weight, qparams=weight_qparams)
        qbias = None
        if self.bias is not None:
            qbias = quantize(self.bias, num_bits=self.num_bits_weight + self.num_bits, flatten_dims=(0, -1))
        return _contract(self, qinput, qweight, qbias, None)


class _RangeBN2Fn(Function):
    """quantize2.RangeBN after its input quantizer (:305-345): chunked range statistics (the quantize.RangeBN kernels),
    un-quantized scale / gamma / beta, normalisation by true division, and in backward the output-gradient quantizer
    (:343-345, per-sample flatten, extreme ranges) followed by the closed-form gradient (rangebn.cu) with
    qscale := 1 / (scale + eps), qgamma := gamma."""

    @staticmethod
    def forward(ctx, xq, gamma, beta, running_mean, running_var, training, momentum, num_chunks, eps, bits_grad):
        xq = _abi._f32c(xq, "x")
        nhwc = _abi.is_nhwc(xq)
        B, Cn = xq.shape[0], xq.shape[1]
        HW = xq.numel() // (B * Cn)
        count, chunk_len = float(B * HW), 0.0
        red = amax = amin = None
        if training:
            red, amax, amin = _abi.rangebn_stats(xq, B, Cn, HW, num_chunks, want_arg=any(ctx.needs_input_grad[:3]), nhwc=nhwc)
            chunk_len = count / num_chunks
        vec = _abi.rangebn2_finalize(red, Cn, num_chunks, count, chunk_len, eps, momentum, training, running_mean, running_var)
        out = _abi.rangebn2_apply(xq, B, Cn, HW, vec, gamma, beta, nhwc=nhwc)
        ctx.dims = (B, Cn, HW, num_chunks, count, chunk_len, training, bits_grad, gamma is not None, beta is not None)
        ctx.nhwc = nhwc
        g_row = gamma.detach() if gamma is not None else torch.ones_like(vec[0])
        # rows as the backward kernels index them: mean, scale_raw, qscale, qgamma
        ctx.save_for_backward(xq, torch.stack([vec[0], vec[3], vec[3], g_row]), amax, amin)
        return out

    @staticmethod
    def backward(ctx, g):
        xq, vec, amax, amin = ctx.saved_tensors
        B, Cn, HW, K, count, chunk_len, training, bits_grad, has_gamma, has_beta = ctx.dims
        g = g.contiguous(memory_format=torch.channels_last) if ctx.nhwc else g.contiguous()
        gq = _quantize_grad_tensor(g, bits_grad, None, (1, -1), 0, False, True) if bits_grad is not None else g
        if ctx.nhwc and not _abi.is_nhwc(gq):
            gq = gq.contiguous(memory_format=torch.channels_last)
        dx, dgamma, dbeta = _abi.rangebn_backward(gq, xq, B, Cn, HW, K if training else 0, count, chunk_len, training, vec,
                                                  has_gamma, amax, amin, chunks_global=K, nhwc=ctx.nhwc)
        return (dx if ctx.needs_input_grad[0] else None, dgamma if has_gamma else None, dbeta if has_beta else None,
                None, None, None, None, None, None, None)


class RangeBN(nn.Module):
    """quantize2.py:272-349 ("normalized RangeBN", un-quantized affine parameters)."""

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
        self.quantize_input = QuantMeasure(self.num_bits, inplace=True, shape_measure=(1, 1, 1, 1), flatten_dims=(1, -1))
        self.eps = eps
        self.num_chunks = num_chunks
        self.reset_params()

    def reset_params(self):
        if self.weight is not None:
            self.weight.data.uniform_()
        if self.bias is not None:
            self.bias.data.zero_()

    def forward(self, x):
        x = self.quantize_input(x)
        if x.dim() == 2:
            x = x.unsqueeze(-1).unsqueeze(-1)
        out = _RangeBN2Fn.apply(x, self.weight, self.bias, self.running_mean, self.running_var, self.training,
                                self.momentum, self.num_chunks, self.eps, self.num_bits_grad)
        if out.size(3) == 1 and out.size(2) == 1:
            out = out.squeeze(-1).squeeze(-1)
        return out


class RangeBN1d(RangeBN):
    """quantize2.py:352-359."""

    def __init__(self, num_features, dim=1, momentum=0.1, affine=True, num_chunks=16, eps=1e-5, num_bits=8,
                 num_bits_grad=8):
        super(RangeBN1d, self).__init__(num_features, dim, momentum, affine, num_chunks, eps, num_bits, num_bits_grad)
        self.quantize_input = QuantMeasure(self.num_bits, inplace=True, shape_measure=(1, 1), flatten_dims=(1, -1))
