"""nn.Module wrappers around ``oracle.qref`` (CPU, fp32).  TEST INFRASTRUCTURE ONLY.

Same constructor arguments, parameter and buffer names as the reference's
QuantMeasure / QConv2d / QLinear / RangeBN (models/modules/quantize.py:198-505)
so that state dicts interchange with the reference and with the CUDA modules.
Used as (a) the checker in tests / smoke and (b) the timed CPU baseline of
``bench.py`` (the reference itself does not travel to the GPU box).
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import qref


class QuantMeasure(nn.Module):
    def __init__(self, num_bits=8, momentum=None, method="avg"):
        super().__init__()
        assert method in ("avg", "aciq")
        t = qref.RangeTracker(num_bits, momentum, method)
        for name in ("running_min", "running_max", "num_measurements", "running_var", "running_mean"):
            self.register_buffer(name, getattr(t, name))
        self.momentum, self.num_bits, self.method = momentum, num_bits, method
        self.enable_quant = True

    def _tracker(self):
        t = qref.RangeTracker.__new__(qref.RangeTracker)
        t.__dict__.update(running_min=self.running_min, running_max=self.running_max,
                          num_measurements=self.num_measurements, running_var=self.running_var,
                          running_mean=self.running_mean, momentum=self.momentum, bits=self.num_bits,
                          method=self.method, enable_quant=self.enable_quant, training=self.training)
        return t

    def forward(self, x):
        return self._tracker()(x)


class _QContract:
    """State shared by QConv2d / QLinear (quantize.py:280-299, 363-383)."""

    def _init_q(self, num_bits, num_bits_weight, num_bits_grad, biprecision, bias_quant, bias):
        self.num_bits = num_bits
        self.num_bits_weight = num_bits_weight or num_bits
        self.num_bits_grad = num_bits_grad
        self.biprecision = biprecision
        self.bias_quant = bias_quant and bias
        self.enable_quant = True
        self.quantize_input = QuantMeasure(num_bits)
        lo, hi = qref.weight_row_range(self.weight.detach())
        self.register_buffer("weight_min", lo)
        self.register_buffer("weight_max", hi)
        if self.bias_quant:
            self.register_buffer("bias_min", self.bias.detach().min())
            self.register_buffer("bias_max", self.bias.detach().max())

    def _contract(self, op, x):
        xq = self.quantize_input(x)
        if not self.enable_quant:
            return op(x, self.weight, self.bias)
        self.weight_min, self.weight_max = qref.weight_row_range(self.weight)
        if self.bias is not None:
            self.bias_min, self.bias_max = self.bias.min(), self.bias.max()
        return qref.qcontract(op, xq, self.weight, self.bias, self.num_bits_weight,
                              self.num_bits_grad, self.biprecision, self.bias_quant)


class QConv2d(nn.Conv2d, _QContract):
    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1,
                 bias=True, num_bits=8, num_bits_weight=None, num_bits_grad=None, biprecision=False,
                 bias_quant=True, per_channel=True, **kw):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, bias, **kw)
        assert per_channel, "per_channel=False raises in the reference too (SURVEY 0.8)"
        self._init_q(num_bits, num_bits_weight, num_bits_grad, biprecision, bias_quant, bias)

    def forward(self, x):
        return self._contract(
            lambda a, w, b: F.conv2d(a, w, b, self.stride, self.padding, self.dilation, self.groups), x)


class QLinear(nn.Linear, _QContract):
    def __init__(self, in_features, out_features, bias=True, num_bits=8, num_bits_weight=None,
                 num_bits_grad=None, biprecision=False, bias_quant=True, per_channel=True, **kw):
        super().__init__(in_features, out_features, bias, **kw)
        assert per_channel
        self._init_q(num_bits, num_bits_weight, num_bits_grad, biprecision, bias_quant, bias)

    def forward(self, x):
        return self._contract(F.linear, x)


class RangeBN(nn.Module):
    def __init__(self, num_features, dim=1, momentum=0.1, affine=True, num_chunks=16, eps=1e-5,
                 num_bits=8, num_bits_grad=8):
        super().__init__()
        self.register_buffer("running_mean", torch.zeros(num_features))
        self.register_buffer("running_var", torch.zeros(num_features))
        self.momentum, self.dim, self.eps, self.num_chunks = momentum, dim, eps, num_chunks
        self.num_bits, self.num_bits_grad = num_bits, num_bits_grad
        if affine:
            self.bias = nn.Parameter(torch.zeros(num_features))
            self.weight = nn.Parameter(torch.rand(num_features))     # quantize.py:455-459
        else:
            self.bias = self.weight = None
        self.quantize_input = QuantMeasure(num_bits)

    def forward(self, x):
        xq = self.quantize_input(x)
        return qref.rangebn_forward(xq, self.weight, self.bias, self.running_mean, self.running_var,
                                    self.training, self.momentum, self.num_chunks, self.eps,
                                    self.num_bits, self.num_bits_grad)
