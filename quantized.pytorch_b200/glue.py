"""stock-torch glue between the quantized modules, on native kernels.

``MaxPool2d`` is a drop-in for ``torch.nn.MaxPool2d`` (models/resnet_quantized.py:150): channels-last fp32 CUDA inputs with
C % 8 == 0 go through ``qb_maxpool2d_nhwc_fwd/bwd`` (same values, tie-breaking and gradient as ATen, one byte of index per
element instead of eight); every other input takes torch's own path, which IS the reference behaviour for this glue."""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _abi


class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, stride, pad):
        out, idx = _abi.maxpool2d_nhwc_fwd(x, k, stride, pad)
        ctx.save_for_backward(idx)
        ctx.geom = (tuple(x.shape), k, stride, pad)
        return out

    @staticmethod
    def backward(ctx, g):
        (idx,) = ctx.saved_tensors
        shape, k, stride, pad = ctx.geom
        g = _abi._f32c(g, "grad").contiguous(memory_format=torch.channels_last)
        return _abi.maxpool2d_nhwc_bwd(g, idx, shape, k, stride, pad), None, None, None


def _native_ok(x, k, stride, pad):
    return (x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and x.shape[1] % 8 == 0 and _abi.is_nhwc(x)
            and isinstance(k, int) and isinstance(stride, int) and isinstance(pad, int) and 1 <= k <= 15 and 2 * pad <= k
            and x.numel() // 8 < 2 ** 32 and x.data_ptr() % 32 == 0)


class MaxPool2d(nn.MaxPool2d):
    """torch.nn.MaxPool2d whose channels-last CUDA path runs on libqb200 (dilation 1, no ceil_mode / return_indices)."""

    def forward(self, x):
        k, s, p = self.kernel_size, self.stride, self.padding
        if self.dilation == 1 and not self.ceil_mode and not self.return_indices and _native_ok(x, k, s, p):
            return _MaxPoolFn.apply(x, k, s, p)
        return F.max_pool2d(x, k, s, p, self.dilation, self.ceil_mode, self.return_indices)
