"""Consumers of the hot path: ResNet (ImageNet / CIFAR) and MobileNet-v1 built
from QConv2d + RangeBN + QLinear, all 8-bit with the bi-precision backward.

Architecture, initialisation, module names (hence state-dict keys) and training
regimes follow the reference's models/resnet_quantized.py:13-261 and
models/mobilenet_quantized.py:21-173, so reference checkpoints interchange.
The reference's own model files also run unchanged on this package (they only
import names from ``models.modules.quantize``) - these
builders exist because the reference tree is not present on the GPU box.

Every builder takes ``q``: the namespace supplying QConv2d / QLinear / RangeBN
(default: the CUDA-backed ``qtorch_b200.quantize``; the test-suite passes the CPU
oracle's modules to obtain the same network on the checker side).
"""
import math

import torch.nn as nn

BITS = dict(num_bits=8, num_bits_weight=8, num_bits_grad=8, biprecision=True)   # resnet_quantized.py:7-10


def _default_q():
    from . import quantize
    return quantize


def _qconv(q, cin, cout, k, stride=1, padding=0, groups=1, bias=False):
    return q.QConv2d(cin, cout, kernel_size=k, stride=stride, padding=padding, groups=groups, bias=bias, **BITS)


def _qbn(q, c):
    return q.RangeBN(c, num_bits=8, num_bits_grad=8)


def _maxpool(q):
    """nn.MaxPool2d (resnet_quantized.py:150); on the native module set its channels-last CUDA path runs on libqb200
    (glue.py) - identical values and gradients."""
    if getattr(q, "NATIVE_GLUE", False):
        from . import glue
        return glue.MaxPool2d
    return nn.MaxPool2d


def _prepare_weights(model, x):
    """Native module set on CUDA: every layer's weight kernels are issued on a side stream now (quantize.prepare_weights)."""
    if model._native_glue and x.is_cuda:
        from . import quantize as Q
        Q.prepare_weights(model)


def _join_prepared(model, out):
    if model._native_glue and out.is_cuda:
        from . import quantize as Q
        Q.join_prepared(out.device)
    return out


def _avgpool(q):
    """nn.AvgPool2d (resnet_quantized.py:153, mobilenet_quantized.py:114); native module set: glue.AvgPool2d, same values."""
    if getattr(q, "NATIVE_GLUE", False):
        from . import glue
        return glue.AvgPool2d
    return nn.AvgPool2d


def _relu(q, inplace=True):
    """nn.ReLU(inplace=True) (resnet_quantized.py:47,147) / nn.ReLU() (mobilenet_quantized.py:38-50); native module set:
    glue.ReLU (same values, plus the statistics the next quantized layer needs)."""
    if getattr(q, "NATIVE_GLUE", False):
        from . import glue
        return glue.ReLU(inplace=inplace)
    return nn.ReLU(inplace=inplace)


def _bn_relu(bn, relu, x, res=None, skip=None):
    """``relu(bn(x))`` / ``y = bn(x); y += res; relu(y)`` (models/resnet_quantized.py:47-75, 100-112); one fused pass when the
    modules are this package's (RangeBN.forward_relu), the plain composition for any other module namespace (the oracle)."""
    if hasattr(bn, "forward_relu"):
        return bn.forward_relu(x, relu, res, skip) if skip is not None else bn.forward_relu(x, relu, res)
    y = bn(x)
    if res is None:
        return relu(y)
    res = res() if callable(res) else res              # after bn(x), as in the reference (order of the noise draws in backward)
    if hasattr(relu, "add_relu"):
        return relu.add_relu(y, res)
    y += res
    return relu(y)


def _arm_skip(block, last_bn):
    """-> the holder to pass to the block's final fused RangeBN (identity skip), or None."""
    if not hasattr(last_bn, "forward_relu"):
        return None
    holder = {}
    block.conv1._skip_holder = holder
    if block.downsample is None:
        return holder
    ds0 = block.downsample[0] if isinstance(block.downsample, nn.Sequential) and len(block.downsample) else None
    if ds0 is not None and hasattr(ds0, "quantize_input"):
        ds0._skip_out_holder = holder
    return None


class _Identity(nn.Module):
    def forward(self, x):
        return x


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, q, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = _qconv(q, inplanes, planes, 3, stride, 1)
        self.bn1 = _qbn(q, planes)
        self.relu = _relu(q)
        self.conv2 = _qconv(q, planes, planes, 3, 1, 1)
        self.bn2 = _qbn(q, planes)
        self.downsample = downsample
        self._last_bn = "bn2"

    def forward(self, x):
        # native modules: conv1's input-gradient kernel accumulates into the other gradient of x - the skip gradient of an
        # identity block, or the downsample convolution's input gradient - instead of autograd adding the two tensors
        skip = _arm_skip(self, self.bn2)
        y = _bn_relu(self.bn1, self.relu, self.conv1(x))
        res = (lambda: x) if self.downsample is None else (lambda: self.downsample(x))
        return _bn_relu(self.bn2, self.relu, self.conv2(y), res, skip)   # native: normalise + add + ReLU + next layer's statistics


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, q, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = _qconv(q, inplanes, planes, 1)
        self.bn1 = _qbn(q, planes)
        self.conv2 = _qconv(q, planes, planes, 3, stride, 1)
        self.bn2 = _qbn(q, planes)
        self.conv3 = _qconv(q, planes, planes * 4, 1)
        self.bn3 = _qbn(q, planes * 4)
        self.relu = _relu(q)
        self.downsample = downsample
        self._last_bn = "bn3"

    def forward(self, x):
        skip = _arm_skip(self, self.bn3)
        y = _bn_relu(self.bn1, self.relu, self.conv1(x))
        y = _bn_relu(self.bn2, self.relu, self.conv2(y))
        res = (lambda: x) if self.downsample is None else (lambda: self.downsample(x))
        return _bn_relu(self.bn3, self.relu, self.conv3(y), res, skip)


class QResNet(nn.Module):
    """resnet_quantized.py:116-232."""

    def __init__(self, q, block, stages, num_classes, stem, pool, weight_decay=1e-4):
        super().__init__()
        self._q = (q.QConv2d, q.QLinear, q.RangeBN)
        self._native_glue = bool(getattr(q, "NATIVE_GLUE", False))
        width0, stem_k, stem_s, stem_p, use_maxpool = stem
        self.inplanes = width0
        self.conv1 = _qconv(q, 3, width0, stem_k, stem_s, stem_p)
        self.bn1 = _qbn(q, width0)
        self.relu = _relu(q)
        self.maxpool = _maxpool(q)(3, 2, 1) if use_maxpool else _Identity()
        for i in range(4):
            if i < len(stages):
                planes, n, stride = stages[i]
                layer = self._stage(q, block, planes, n, stride)
            else:
                layer = _Identity()
            setattr(self, "layer%d" % (i + 1), layer)
        self.avgpool = _avgpool(q)(pool)
        self.fc = q.QLinear(self.inplanes, num_classes, **BITS)
        self.weight_decay = weight_decay
        self._init()

    def _stage(self, q, block, planes, n, stride):
        down = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            down = nn.Sequential(_qconv(q, self.inplanes, planes * block.expansion, 1, stride),
                                 _qbn(q, planes * block.expansion))
        blocks = [block(q, self.inplanes, planes, stride, down)]
        self.inplanes = planes * block.expansion
        blocks += [block(q, self.inplanes, planes) for _ in range(1, n)]
        return nn.Sequential(*blocks)

    def _init(self):
        QConv2d, _, RangeBN = self._q
        for m in self.modules():                       # resnet_quantized.py:19-34
            if isinstance(m, QConv2d):
                fan = m.kernel_size[0] * m.kernel_size[1] * m.out_channels
                m.weight.data.normal_(0, math.sqrt(2.0 / fan))
            elif isinstance(m, RangeBN):
                m.weight.data.fill_(1)
                m.bias.data.zero_()
        for m in self.modules():
            if isinstance(m, (BasicBlock, Bottleneck)):
                nn.init.constant_(getattr(m, m._last_bn).weight, 0)
        self.fc.weight.data.normal_(0, 0.01)
        self.fc.bias.data.zero_()

    def forward(self, x):
        _prepare_weights(self, x)
        x = self.conv1(x)
        if hasattr(self.maxpool, "relu_pool") and hasattr(self.bn1, "forward_relu_pool"):
            # native modules: the pooling reads bn1's codes (the normalised full-size tensor never exists), ReLU on the pooled result
            x = self.bn1.forward_relu_pool(x, self.relu, self.maxpool)
        elif hasattr(self.maxpool, "relu_pool"):       # native glue: pool first, ReLU on the 4x smaller result (same values)
            x = self.maxpool.relu_pool(self.bn1(x), self.relu)
        else:
            x = self.maxpool(self.relu(self.bn1(x)))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = self.avgpool(x)
        return _join_prepared(self, self.fc(x.view(x.size(0), -1)))

    def regularization(self):
        """resnet_quantized.py:157-165: 0.5*wd*sum(w^2) over conv/linear weights and biases (native module set: one
        multi-tensor launch pair, glue.regularization)."""
        if self._native_glue and next(self.parameters()).is_cuda:
            from . import glue
            return glue.regularization(self, self.weight_decay)
        total = 0
        for m in self.modules():
            if isinstance(m, (nn.Conv2d, nn.Linear)):
                total = total + m.weight.pow(2).sum()
                if m.bias is not None:
                    total = total + m.bias.pow(2).sum()
        return self.weight_decay * 0.5 * total


_IMAGENET_LAYERS = {18: (BasicBlock, [2, 2, 2, 2]), 34: (BasicBlock, [3, 4, 6, 3]), 50: (Bottleneck, [3, 4, 6, 3]),
                    101: (Bottleneck, [3, 4, 23, 3]), 152: (Bottleneck, [3, 8, 36, 3])}


def resnet_quantized(q=None, depth=None, dataset="imagenet", num_classes=None):
    """resnet_quantized.py:235-261."""
    q = q or _default_q()
    if dataset == "imagenet":
        block, layers = _IMAGENET_LAYERS[depth or 50]
        stages = [(64 * 2 ** i, n, 1 if i == 0 else 2) for i, n in enumerate(layers)]
        model = QResNet(q, block, stages, num_classes or 1000, stem=(64, 7, 2, 3, True), pool=7)
        # resnet_quantized.py:187-201: linear warm-up of the learning rate over the first 5 epochs (5004 steps each at
        # batch 256), expressed - as in the reference - as the source of a step lambda for its regime scheduler
        scale, rate = 1.0, (0.1 - 0.0) / (5004 * 5)
        model.regime = [{"epoch": 0, "optimizer": "SGD", "momentum": 0.9,
                         "step_lambda": "lambda t: {'lr': %s + t * %s}" % (0, rate)},
                        {"epoch": 5, "lr": scale * 1e-1}, {"epoch": 30, "lr": scale * 1e-2},
                        {"epoch": 60, "lr": scale * 1e-3}, {"epoch": 80, "lr": scale * 1e-4}]
        return model
    if dataset == "cifar10":
        n = int(((depth or 56) - 2) / 6)
        stages = [(16, n, 1), (32, n, 2), (64, n, 2)]
        model = QResNet(q, BasicBlock, stages, num_classes or 10, stem=(16, 3, 1, 1, False), pool=8)
        model.regime = [{"epoch": 0, "optimizer": "SGD", "lr": 1e-1, "weight_decay": 1e-4, "momentum": 0.9},
                        {"epoch": 81, "lr": 1e-2}, {"epoch": 122, "lr": 1e-3, "weight_decay": 0},
                        {"epoch": 164, "lr": 1e-4}]
        return model
    raise ValueError("dataset must be 'imagenet' or 'cifar10'")


class _DWSeparable(nn.Module):
    """mobilenet_quantized.py:33-53: depthwise 3x3 (bias=True, so bias_quant is live) + RangeBN + ReLU,
    pointwise 1x1 + RangeBN + ReLU; children live under ``components`` to keep the state-dict keys."""

    def __init__(self, q, cin, cout, stride):
        super().__init__()
        self.components = nn.Sequential(
            _qconv(q, cin, cin, 3, stride, 1, groups=cin, bias=True), _qbn(q, cin), _relu(q, False),
            _qconv(q, cin, cout, 1), _qbn(q, cout), _relu(q, False))

    def forward(self, x):
        c = self.components
        x = _bn_relu(c[1], c[2], c[0](x))
        return _bn_relu(c[4], c[5], c[3](x))


class QMobileNet(nn.Module):
    """mobilenet_quantized.py:56-160."""

    def __init__(self, q, width=1.0, shallow=False, num_classes=1000):
        super().__init__()
        self._q = (q.QConv2d, q.QLinear, q.RangeBN)
        self._native_glue = bool(getattr(q, "NATIVE_GLUE", False))
        c = lambda n: int(round(width * n))
        plan = [(32, 64, 1), (64, 128, 2), (128, 128, 1), (128, 256, 2), (256, 256, 1), (256, 512, 2)]
        if not shallow:
            plan += [(512, 512, 1)] * 5
        plan += [(512, 1024, 2), (1024, 1024, 1)]
        layers = [_qconv(q, 3, c(32), 3, 2, 1), _qbn(q, c(32)), _relu(q, True)]
        layers += [_DWSeparable(q, c(a), c(b), s) for a, b, s in plan]
        self.features = nn.Sequential(*layers)
        self.avg_pool = _avgpool(q)(7)
        self.fc = q.QLinear(c(1024), num_classes, **BITS)
        self.regime = [{"epoch": 0, "optimizer": "SGD", "lr": 1e-1, "momentum": 0.9},
                       {"epoch": 30, "lr": 1e-2}, {"epoch": 60, "lr": 1e-3}, {"epoch": 80, "lr": 1e-4}]
        QConv2d, _, RangeBN = self._q
        for m in self.modules():                       # mobilenet_quantized.py:21-30
            if isinstance(m, QConv2d):
                fan = m.kernel_size[0] * m.kernel_size[1] * m.out_channels
                m.weight.data.normal_(0, math.sqrt(2.0 / fan))
            elif isinstance(m, RangeBN):
                m.weight.data.fill_(1)
                m.bias.data.zero_()
        self.fc.weight.data.normal_(0, 0.01)
        self.fc.bias.data.zero_()

    def forward(self, x):
        f = self.features
        _prepare_weights(self, x)
        x = _bn_relu(f[1], f[2], f[0](x))
        for i in range(3, len(f)):
            x = f[i](x)
        x = self.avg_pool(x)
        return _join_prepared(self, self.fc(x.view(x.size(0), -1)))

    def regularization(self, weight_decay=4e-5):
        if self._native_glue and next(self.parameters()).is_cuda:
            from . import glue
            return glue.regularization(self, weight_decay)
        total = 0
        for m in self.modules():
            if isinstance(m, (nn.Conv2d, nn.Linear)):
                total = total + m.weight.pow(2).sum()
                if m.bias is not None:
                    total = total + m.bias.pow(2).sum()
        return weight_decay * 0.5 * total


def mobilenet_quantized(q=None, num_classes=None, width=None, shallow=False):
    """mobilenet_quantized.py:163-173."""
    return QMobileNet(q or _default_q(), width or 1.0, bool(shallow), num_classes or 1000)
