"""stock-torch glue between the quantized modules, on native kernels.

``MaxPool2d`` is a drop-in for ``torch.nn.MaxPool2d`` (models/resnet_quantized.py:150): channels-last fp32 CUDA inputs with
C % 8 == 0 go through ``qb_maxpool2d_nhwc_fwd/bwd`` (same values, tie-breaking and gradient as ATen, one byte of index per
element instead of eight); every other input takes torch's own path, which IS the reference behaviour for this glue."""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _abi


# ReLU -> MaxPool2d (resnet_quantized.py:147-150).  glue.ReLU remembers its latest output; a glue.MaxPool2d that is fed
# exactly that tensor computes, in its backward, the gradient w.r.t. the ReLU's INPUT (windows whose maximum is not
# positive contribute nothing) together with its range, and leaves it marked; the ReLU backward that receives the marked
# tensor passes it on untouched.  If the ReLU output has other consumers autograd sums the gradients first, the mark no
# longer matches and the ReLU backward masks the sum - masking is idempotent, so the result is the same.
_relu_out = {"t": None}        # (tensor, version) of the latest glue.ReLU output produced under autograd
_premasked = {"t": None}       # (dx, version, range) left by the fused pooling backward


def _is_latest_relu_output(x):
    ent = _relu_out["t"]
    return (ent is not None and ent[0].data_ptr() == x.data_ptr() and ent[0].shape == x.shape
            and ent[0].stride() == x.stride() and x._version == ent[1])


class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, stride, pad, after_relu=False):
        out, idx = _abi.maxpool2d_nhwc_fwd(x, k, stride, pad)
        ctx.after_relu = bool(after_relu)
        if ctx.after_relu:
            ctx.save_for_backward(idx, out)
        else:
            ctx.save_for_backward(idx)
        ctx.geom = (tuple(x.shape), k, stride, pad)
        return out

    @staticmethod
    def backward(ctx, g):
        shape, k, stride, pad = ctx.geom
        g = _abi._f32c(g, "grad").contiguous(memory_format=torch.channels_last)
        if ctx.after_relu:
            idx, out = ctx.saved_tensors
            dx, rng = _abi.maxpool2d_relu_nhwc_bwd(g, idx, out, shape, k, stride, pad)
            _premasked["t"] = (dx, dx._version, rng)
            return dx, None, None, None, None
        (idx,) = ctx.saved_tensors
        return _abi.maxpool2d_nhwc_bwd(g, idx, shape, k, stride, pad), None, None, None, None


class _ReLUMaxPoolFn(torch.autograd.Function):
    """maxpool(relu(x)) computed as relu(maxpool(x)) - max and the monotone clamp commute, and whenever the two argmax
    choices could differ (a window without a positive element) the window's gradient is zero anyway - so the full-size ReLU
    output is never written: the clamp and the next QuantMeasure's statistics run on the 4x smaller pooled tensor, and the
    backward is the one-pass gather of qb_maxpool2d_relu_nhwc_bwd (gradient w.r.t. x with its range)."""

    @staticmethod
    def forward(ctx, x, k, stride, pad):
        out, idx = _abi.maxpool2d_nhwc_fwd(x, k, stride, pad)
        out, stats = _abi.relu_measure(out)
        ctx.save_for_backward(idx, out)
        ctx.geom = (tuple(x.shape), k, stride, pad)
        _last["stats"] = stats
        return out

    @staticmethod
    def backward(ctx, g):
        from . import quantize as Q
        idx, out = ctx.saved_tensors
        shape, k, stride, pad = ctx.geom
        g = _abi._f32c(g, "grad").contiguous(memory_format=torch.channels_last)
        dx, rng = _abi.maxpool2d_relu_nhwc_bwd(g, idx, out, shape, k, stride, pad)
        Q._stash_grad_range(dx, rng)
        return dx, None, None, None


def _native_ok(x, k, stride, pad):
    return (x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and x.shape[1] % 8 == 0 and _abi.is_nhwc(x)
            and isinstance(k, int) and isinstance(stride, int) and isinstance(pad, int) and 1 <= k <= 15 and 2 * pad <= k
            and x.numel() // 8 < 2 ** 32 and x.data_ptr() % 32 == 0)


class MaxPool2d(nn.MaxPool2d):
    """torch.nn.MaxPool2d whose channels-last CUDA path runs on libqb200 (dilation 1, no ceil_mode / return_indices)."""

    def forward(self, x):
        k, s, p = self.kernel_size, self.stride, self.padding
        if self.dilation == 1 and not self.ceil_mode and not self.return_indices and _native_ok(x, k, s, p):
            fused = torch.is_grad_enabled() and x.requires_grad and _is_latest_relu_output(x) and x.data_ptr() % 32 == 0
            return _MaxPoolFn.apply(x, k, s, p, fused)
        return F.max_pool2d(x, k, s, p, self.dilation, self.ceil_mode, self.return_indices)

    def relu_pool(self, x, relu):
        """``self(relu(x))`` (models/resnet_quantized.py:147-150) without materialising the ReLU output; falls back to
        exactly that composition whenever the native path does not apply."""
        from . import quantize as Q
        k, s, p = self.kernel_size, self.stride, self.padding
        if not (isinstance(relu, ReLU) and relu.training and self.dilation == 1 and not self.ceil_mode and not self.return_indices
                and _native_ok(x, k, s, p) and torch.is_grad_enabled() and x.requires_grad):
            return self(relu(x))
        out = _ReLUMaxPoolFn.apply(x, k, s, p)
        stats, _last["stats"] = _last["stats"], None
        if stats is not None:
            Q._stash_fwd_stats(out, stats)
        return out


class _AvgPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k):
        ctx.geom = (tuple(x.shape), k)
        return _abi.avgpool2d_nhwc_fwd(x, k)

    @staticmethod
    def backward(ctx, g):
        shape, k = ctx.geom
        g = _abi._f32c(g, "grad")
        if not g.is_contiguous(memory_format=torch.channels_last):
            g = g.contiguous(memory_format=torch.channels_last)
        return _abi.avgpool2d_nhwc_bwd(g, shape, k), None


class AvgPool2d(nn.AvgPool2d):
    """torch.nn.AvgPool2d (models/resnet_quantized.py:153, models/mobilenet_quantized.py:114) whose channels-last CUDA path
    runs on libqb200 for the geometry those models use: stride == kernel, no padding, floor mode."""

    def forward(self, x):
        k, s = self.kernel_size, self.stride
        if (isinstance(k, int) and s == k and self.padding == 0 and not self.ceil_mode and self.divisor_override is None
                and x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and x.shape[1] % 4 == 0
                and x.is_contiguous(memory_format=torch.channels_last) and x.shape[2] >= k and x.shape[3] >= k
                and x.data_ptr() % 16 == 0):
            return _AvgPoolFn.apply(x, k)
        return super().forward(x)


class _CrossEntropyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target, ignore_index):
        loss, row_state, n_valid = _abi.cross_entropy_fwd(logits, target, ignore_index)
        ctx.save_for_backward(logits, target, row_state, n_valid)
        ctx.ignore_index = ignore_index
        return loss

    @staticmethod
    def backward(ctx, g):
        logits, target, row_state, n_valid = ctx.saved_tensors
        g = g.to(torch.float32).contiguous()
        return _abi.cross_entropy_bwd(logits, target, row_state, n_valid, g, ctx.ignore_index), None, None


class CrossEntropyLoss(nn.CrossEntropyLoss):
    """torch.nn.CrossEntropyLoss as the reference's training loop uses it (main.py:362: ``criterion(output, target)`` with class
    index targets, mean reduction): forward and backward are one pass each over the logits on libqb200, the [B, C]
    log-softmax is never stored.  Any other configuration (class weights, label smoothing, probabilities as targets, other
    reductions, non-CUDA tensors) takes torch's own path."""

    def forward(self, input, target):
        if (input.is_cuda and input.dtype == torch.float32 and input.dim() == 2 and input.is_contiguous()
                and target.dtype == torch.int64 and target.dim() == 1 and target.is_contiguous() and self.weight is None
                and self.reduction == "mean" and self.label_smoothing == 0.0):
            return _CrossEntropyFn.apply(input, target, self.ignore_index)
        return super().forward(input, target)


def _relu_backward(g, out):
    """aten::threshold_backward(g, out, 0); on dense fp32 CUDA tensors of one layout the native pass also leaves the
    range of the result for the gradient quantizer that consumes it (RangeBN / QConv2d backward)."""
    ent, _premasked["t"] = _premasked["t"], None
    if (ent is not None and ent[0].data_ptr() == g.data_ptr() and ent[0].shape == g.shape and ent[0].stride() == g.stride()
            and g._version == ent[1] and g.shape == out.shape):
        from . import quantize as Q
        Q._stash_grad_range(g, ent[2])              # already masked by the fused pooling backward
        return g
    if (g.is_cuda and g.dtype == torch.float32 and out.dtype == torch.float32 and g.shape == out.shape
            and g.stride() == out.stride() and g.numel() > 0
            and (out.is_contiguous() or (out.dim() == 4 and out.is_contiguous(memory_format=torch.channels_last)))):
        from . import quantize as Q
        dg, rng = _abi.relu_bwd_minmax(g, out)
        Q._stash_grad_range(dg, rng)
        return dg
    return torch.ops.aten.threshold_backward(g, out, 0)


_last = {"stats": None}          # statistics block of the most recent _ReLUFn.forward (picked up right after apply)


class _ReLUFn(torch.autograd.Function):
    """relu in place + the statistics of the result for the QuantMeasure of the consuming layer (stashed by tensor)."""

    @staticmethod
    def forward(ctx, x, inplace):
        from . import quantize as Q
        out, stats = _abi.relu_measure(x, None if inplace else torch.empty_like(x))
        if inplace:
            ctx.mark_dirty(x)
        ctx.save_for_backward(out)
        _last["stats"] = stats
        return out

    @staticmethod
    def backward(ctx, g):
        (out,) = ctx.saved_tensors
        return _relu_backward(g, out), None


class _AddReLUFn(torch.autograd.Function):
    """relu(y + res) written into y, plus the statistics of the result; the masked gradient goes to both addends."""

    @staticmethod
    def forward(ctx, y, res):
        out, stats = _abi.add_relu_measure(y, res)
        ctx.mark_dirty(y)
        ctx.save_for_backward(out)
        _last["stats"] = stats
        return out

    @staticmethod
    def backward(ctx, g):
        (out,) = ctx.saved_tensors
        dg = _relu_backward(g, out)
        return dg, dg


class ReLU(nn.ReLU):
    """torch.nn.ReLU whose dense fp32 CUDA path also delivers the next QuantMeasure's statistics:
    the quantized layer that consumes the output then skips its own statistics pass.  Values and gradient are torch's."""

    def forward(self, x):
        from . import quantize as Q
        dense = x.is_contiguous() or (x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last))
        if not (x.is_cuda and x.dtype == torch.float32 and x.dim() >= 2 and dense and x.numel() > 0 and self.training):
            return super().forward(x)
        if torch.is_grad_enabled() and x.requires_grad:
            out = _ReLUFn.apply(x, self.inplace)
            stats, _last["stats"] = _last["stats"], None
            _relu_out["t"] = (out, out._version)
        else:
            out, stats = _abi.relu_measure(x, None if self.inplace else torch.empty_like(x))
        if stats is not None:
            Q._stash_fwd_stats(out, stats)
        return out

    def add_relu(self, y, res):
        """``y += res; return relu(y)`` - the tail of a residual block (models/resnet_quantized.py:62-68, 105-112) - as ONE
        pass that also delivers the next QuantMeasure's statistics.  Same values as the two torch ops (the sum is rounded
        to fp32 before the clamp).  Falls back to exactly those two ops whenever the fused kernel does not apply."""
        from . import quantize as Q
        dense = y.is_contiguous() or (y.dim() == 4 and y.is_contiguous(memory_format=torch.channels_last))
        if not (self.inplace and self.training and y.is_cuda and y.dtype == torch.float32 and res.dtype == torch.float32
                and y.dim() >= 2 and dense and y.numel() > 0 and y.shape == res.shape and y.stride() == res.stride()
                and y.data_ptr() != res.data_ptr()):
            y += res
            return self.forward(y)
        if torch.is_grad_enabled() and (y.requires_grad or res.requires_grad):
            out = _AddReLUFn.apply(y, res)
            stats, _last["stats"] = _last["stats"], None
        else:
            out, stats = _abi.add_relu_measure(y, res)
        if stats is not None:
            Q._stash_fwd_stats(out, stats)
        return out


# ---------------------------------------------------------------------------------------------------
# Optimizer / regulariser glue of the training step (main.py:362-380) on multi-tensor kernels (glue_opt.cu)
# ---------------------------------------------------------------------------------------------------
class _L2RegFn(torch.autograd.Function):
    """0.5 * wd * SUM_t SUM w_t^2 (models/resnet_quantized.py:157-165) in one launch pair; backward: one launch."""

    @staticmethod
    def forward(ctx, weight_decay, *tensors):
        ctx.wd = weight_decay
        ctx.save_for_backward(*tensors)
        return _abi.l2_sumsq([t.detach() for t in tensors], weight_decay)

    @staticmethod
    def backward(ctx, g):
        tensors = [t.detach() for t in ctx.saved_tensors]
        up = g.detach().reshape(1).float().contiguous()
        return (None,) + tuple(_abi.l2_grad(tensors, ctx.wd, upstream=up))


def regularization(model, weight_decay=1e-4):
    """``model.regularization()`` of the reference (resnet_quantized.py:157-165 / mobilenet_quantized.py:150-158): L2 penalty over
    the weights and biases of every Conv2d / Linear, as a differentiable device scalar - 3 launches instead of ~66 + ~66."""
    ts = []
    for m in model.modules():
        if isinstance(m, (nn.Conv2d, nn.Linear)):
            ts.append(m.weight)
            if m.bias is not None:
                ts.append(m.bias)
    if not ts:
        return 0.0
    if not all(t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() for t in ts):
        total = 0
        for t in ts:
            total = total + t.pow(2).sum()
        return weight_decay * 0.5 * total
    return _L2RegFn.apply(weight_decay, *ts)


class SGD(torch.optim.SGD):
    """torch.optim.SGD whose step() is ONE multi-tensor launch per parameter group (qb_sgd_step) - same arithmetic, same state
    dict ("momentum_buffer"), CUDA-graph capturable.  Groups using dampening / Nesterov / maximize, or non-CUDA parameters,
    take torch's own step."""

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            ps = [p for p in group["params"] if p.grad is not None]
            if not ps:
                continue
            native = (group["dampening"] == 0 and not group["nesterov"] and not group.get("maximize", False)
                      and all(p.is_cuda and p.dtype == torch.float32 and p.is_contiguous() and p.grad.is_contiguous()
                              and not p.grad.is_sparse for p in ps))
            if not native:
                for p in ps:
                    self._torch_single(p, group)
                continue
            mom = group["momentum"]
            first_flags = {("momentum_buffer" not in self.state[p]) for p in ps} if mom != 0 else {False}
            for first in sorted(first_flags):
                sel = [p for p in ps if mom == 0 or (("momentum_buffer" not in self.state[p]) == first)]
                bufs = None
                if mom != 0:
                    if first:
                        for p in sel:
                            self.state[p]["momentum_buffer"] = torch.empty_like(p)
                    bufs = [self.state[p]["momentum_buffer"] for p in sel]
                _abi.sgd_step([p.data for p in sel], [p.grad for p in sel], bufs, group["lr"], mom, group["weight_decay"], first)
        return loss

    def _torch_single(self, p, group):
        g = p.grad
        if group["weight_decay"] != 0:
            g = g.add(p, alpha=group["weight_decay"])
        if group["momentum"] != 0:
            buf = self.state[p].get("momentum_buffer")
            if buf is None:
                buf = self.state[p]["momentum_buffer"] = torch.clone(g).detach()
            else:
                buf.mul_(group["momentum"]).add_(g, alpha=1 - group["dampening"])
            g = g.add(buf, alpha=group["momentum"]) if group["nesterov"] else buf
        p.add_(g, alpha=-group["lr"])
