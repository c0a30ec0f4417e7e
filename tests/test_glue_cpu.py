"""CPU-only checks of the host logic added around the native glue: on tensors the C-ABI kernels do
not take (CPU tensors, other geometries) the glue modules are exactly their torch base classes, the first-layer NCHW route and
the side-stream weight preparation stay out of the way without a GPU, and the native models build with the glue in place and
keep the reference's state-dict keys."""
import torch
import torch.nn as nn


def test_glue_modules_are_their_torch_base_classes_on_cpu():
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import glue
    gen = torch.Generator().manual_seed(0)
    x = torch.randn(4, 8, 7, 7, generator=gen)
    assert torch.equal(glue.AvgPool2d(7)(x), nn.AvgPool2d(7)(x))
    assert torch.equal(glue.AvgPool2d(3, stride=2, padding=1)(x), nn.AvgPool2d(3, stride=2, padding=1)(x))
    logits = torch.randn(6, 10, generator=gen, requires_grad=True)
    ref = logits.detach().clone().requires_grad_()
    t = torch.randint(0, 10, (6,), generator=gen)
    a, b = glue.CrossEntropyLoss()(logits, t), nn.CrossEntropyLoss()(ref, t)
    assert torch.equal(a, b)
    a.backward()
    b.backward()
    assert torch.equal(logits.grad, ref.grad)
    assert torch.equal(glue.CrossEntropyLoss(label_smoothing=0.1)(ref, t), nn.CrossEntropyLoss(label_smoothing=0.1)(ref, t))


def test_first_layer_route_and_weight_preparation_are_inert_without_cuda():
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import models, quantize as Q
    m = Q.QConv2d(3, 64, 7, 2, 3, bias=False)
    x = torch.randn(2, 3, 32, 32)
    assert not Q._stem_takes_nchw(m, x)                    # CPU tensor: the copying route (and, without a GPU, a loud failure later)
    net = models.resnet_quantized(Q, depth=20, dataset="cifar10")
    Q.prepare_weights(net)                                 # no CUDA / switched off: nothing is issued, nothing is stashed
    Q.join_prepared(torch.device("cpu"))
    assert all("_prepped" not in mod.__dict__ for mod in net.modules())
    assert not Q.PREP_WEIGHTS_ASIDE and not Q.FUSE_RANGEBN_FINALIZE and Q.FUSE_QUANT_ROWSUM and Q.STEM_NCHW_DIRECT


def test_native_models_use_the_glue_and_keep_the_reference_keys():
    import qtorch_b200  # noqa: F401
    from oracle import qref_modules
    from qtorch_b200 import glue, models, quantize as Q
    for build in (lambda q: models.resnet_quantized(q, depth=18, dataset="imagenet"), lambda q: models.mobilenet_quantized(q)):
        native, ref = build(Q), build(qref_modules)
        assert sorted(native.state_dict()) == sorted(ref.state_dict())
        pools = [m for m in native.modules() if isinstance(m, nn.AvgPool2d)]
        assert pools and all(isinstance(m, glue.AvgPool2d) for m in pools)
        assert not any(isinstance(m, glue.AvgPool2d) for m in ref.modules())
