"""SURVEY section 8(f) N1 glue: native channels-last MaxPool2d == torch.nn.MaxPool2d bit for bit (values, tie-breaking,
gradient), including post-ReLU inputs whose windows are full of equal zeros."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("shape,k,s,p", [((4, 64, 112, 112), 3, 2, 1), ((3, 16, 17, 23), 3, 2, 1), ((2, 8, 9, 9), 2, 2, 0),
                                          ((2, 24, 13, 11), 3, 1, 1), ((1, 8, 7, 7), 5, 3, 2), ((2, 8, 6, 6), 3, 3, 0)])
@pytest.mark.parametrize("relu", [True, False])
def test_maxpool_matches_torch_bitwise(lib, shape, k, s, p, relu):
    from qtorch_b200 import glue
    gen = torch.Generator().manual_seed(sum(shape) + k)
    x = torch.randn(shape, generator=gen)
    if relu:
        x = x.clamp_min(0)
        x = (x * 4).round() / 4                      # many exact ties besides the zeros
    xa = x.to(DEV).contiguous(memory_format=torch.channels_last).requires_grad_()
    xb = x.to(DEV).contiguous(memory_format=torch.channels_last).requires_grad_()
    n0 = lib.LAUNCHES
    out = glue.MaxPool2d(k, s, p)(xa)
    assert lib.LAUNCHES == n0 + 1                    # the native kernel ran
    ref = torch.nn.MaxPool2d(k, s, p)(xb)
    assert out.shape == ref.shape and out.is_contiguous(memory_format=torch.channels_last)
    assert torch.equal(out, ref)
    go = torch.randn(ref.shape, generator=gen).to(DEV).contiguous(memory_format=torch.channels_last)
    out.backward(go)
    ref.backward(go)
    assert torch.equal(xa.grad, xb.grad)


def test_maxpool_falls_back_to_torch_for_other_layouts(lib):
    from qtorch_b200 import glue
    x = torch.randn(2, 6, 9, 9, device=DEV)          # NCHW-contiguous, C % 8 != 0: torch's own kernel (reference behaviour)
    n0 = lib.LAUNCHES
    out = glue.MaxPool2d(3, 2, 1)(x)
    assert lib.LAUNCHES == n0
    assert torch.equal(out, torch.nn.functional.max_pool2d(x, 3, 2, 1))


@pytest.mark.parametrize("shape,nhwc", [((8, 64, 14, 14), True), ((4, 16, 9, 7), False), ((16, 40), False), ((3, 5, 3, 3), True)])
def test_relu_with_statistics(lib, shape, nhwc):
    """glue.ReLU: torch's values and gradient, in place; the statistics block it stashes equals qb_measure_stats of its
    output and is picked up by the next QuantMeasure (no second statistics launch)."""
    from qtorch_b200 import glue, quantize as Q
    gen = torch.Generator().manual_seed(sum(shape))
    x0 = torch.randn(shape, generator=gen)
    x0.view(-1)[3] = float("nan") if len(shape) == 2 else x0.view(-1)[3]
    xa, xb = x0.to(DEV), x0.to(DEV)
    if nhwc:
        xa, xb = xa.contiguous(memory_format=torch.channels_last), xb.contiguous(memory_format=torch.channels_last)
    xa.requires_grad_(); xb.requires_grad_()
    m = glue.ReLU(inplace=True).train()
    ya = m(xa * 1.0)
    yb = torch.nn.functional.relu(xb * 1.0)
    assert torch.equal(torch.nan_to_num(ya, nan=-7.0), torch.nan_to_num(yb, nan=-7.0))
    go = torch.randn(shape, generator=gen).to(DEV)
    ya.backward(go); yb.backward(go)
    assert torch.equal(torch.nan_to_num(xa.grad, nan=-7.0), torch.nan_to_num(xb.grad, nan=-7.0))
    if len(shape) == 2:
        return                                                       # NaN case: values only
    y = m(xa.detach().clone().requires_grad_() * 1.0)
    want = lib.measure_stats(y.detach())
    got = Q._take_fwd_stats(y.detach())
    assert got is not None and torch.equal(got, want)
    # consumer side: a QuantMeasure fed with the ReLU output launches only its finalize kernel
    qm = Q.QuantMeasure(8).to(DEV).train()
    y = m(xa.detach().clone().requires_grad_() * 1.0)
    lib.profile_begin()
    qm(y)
    names = [n for n, _, _ in lib.profile_end()]
    assert "qb_measure_stats" not in names and "qb_measure_finalize" in names, names


@pytest.mark.parametrize("shape,nhwc", [((4, 64, 20, 20), True), ((3, 16, 9, 7), False), ((5, 32, 1, 1), True), ((6, 40), False)])
def test_add_relu_with_statistics(lib, shape, nhwc):
    """glue.ReLU.add_relu: the block tail `y += res; y = relu(y)` (models/resnet_quantized.py:62-68) in one pass: torch's
    values bit for bit, the masked gradient on both addends, and the statistics of the result for the next QuantMeasure."""
    from qtorch_b200 import glue, quantize as Q
    gen = torch.Generator().manual_seed(sum(shape) + 1)
    a0, r0 = torch.randn(shape, generator=gen).to(DEV), torch.randn(shape, generator=gen).to(DEV)
    if nhwc:
        a0, r0 = a0.contiguous(memory_format=torch.channels_last), r0.contiguous(memory_format=torch.channels_last)
    leaves = [t.clone().requires_grad_() for t in (a0, r0, a0, r0)]
    m = glue.ReLU(inplace=True).train()
    lib.profile_begin()
    ya = m.add_relu(leaves[0] * 1.0, leaves[1] * 1.0)
    names = [n for n, _, _ in lib.profile_end()]
    assert names == ["qb_add_relu_measure"], names
    yb = leaves[2] * 1.0
    yb += leaves[3] * 1.0
    yb = torch.nn.functional.relu(yb)
    assert torch.equal(ya, yb)
    want = lib.measure_stats(ya.detach())
    got = Q._take_fwd_stats(ya.detach())
    assert got is not None and torch.equal(got, want)
    go = torch.randn(shape, generator=gen).to(DEV)
    if nhwc:
        go = go.contiguous(memory_format=torch.channels_last)
    ya.backward(go); yb.backward(go)
    assert torch.equal(leaves[0].grad, leaves[2].grad) and torch.equal(leaves[1].grad, leaves[3].grad)
    # layouts that differ fall back to the two torch ops, same values
    if len(shape) == 4 and shape[2] > 1:
        y2 = m.add_relu(a0.clone(), r0.contiguous())
        assert torch.equal(y2, torch.nn.functional.relu(a0 + r0))


@pytest.mark.parametrize("n", [1, 7, 8, 1000, 8 * 4096 + 3, 3_000_001])
def test_relu_backward_with_range(lib, n):
    """qb_relu_bwd_minmax: aten::threshold_backward's values bit for bit plus the exact {min, max} of the result."""
    gen = torch.Generator().manual_seed(n)
    g = torch.randn(n, generator=gen).to(DEV)
    out = torch.randn(n, generator=gen).clamp_min(0).to(DEV)
    dx, rng = lib.relu_bwd_minmax(g, out)
    want = torch.ops.aten.threshold_backward(g, out, 0)
    assert torch.equal(dx, want)
    assert float(rng[0]) == float(want.min()) and float(rng[1]) == float(want.max())
    dx2, rng2 = lib.relu_bwd_minmax(g, out)                       # workspace left clean
    assert torch.equal(dx2, want) and torch.equal(rng, rng2)


@pytest.mark.parametrize("n", [8, 8 * 1000, 8 * 123457])
def test_relu_backward_from_bit_mask(lib, n):
    """qb_relu_bwd_minmax_bits == qb_relu_bwd_minmax with the ReLU output reduced to one bit per element (bit e%8 of byte e/8),
    with and without the dx tensor."""
    torch.manual_seed(n)
    g = torch.randn(n, device=DEV)
    out = torch.relu(torch.randn(n, device=DEV))
    want, rng = lib.relu_bwd_minmax(g, out)
    keep = (~(out <= 0)).view(-1, 8).to(torch.int32)
    bits = (keep << torch.arange(8, device=DEV, dtype=torch.int32)).sum(1).to(torch.uint8)
    dx, rng2 = lib.relu_bwd_minmax_bits(g, bits)
    assert torch.equal(dx, want) and torch.equal(rng, rng2)
    none, rng3 = lib.relu_bwd_minmax_bits(g, bits, want_dx=False)
    assert none is None and torch.equal(rng, rng3)


def test_rangebn_backward_takes_the_relu_range(lib):
    """RangeBN -> glue.ReLU: the ReLU backward hands its result's range to the RangeBN gradient quantizer (no qb_minmax
    launch), and the gradients equal those of the stock nn.ReLU route under the same injected noise."""
    from qtorch_b200 import glue, quantize as Q
    torch.manual_seed(3)
    x = torch.randn(8, 32, 12, 12, device=DEV).contiguous(memory_format=torch.channels_last)
    go = torch.randn(8, 32, 12, 12, device=DEV).contiguous(memory_format=torch.channels_last)
    noise = (torch.rand(8, 32, 12, 12, device=DEV) - 0.5)
    res = []
    state = Q.RangeBN(32).state_dict()                              # reset_params draws gamma at random: share one draw
    for relu in (glue.ReLU(inplace=True), torch.nn.ReLU(inplace=True)):
        bn = Q.RangeBN(32)
        bn.load_state_dict(state)
        bn = bn.to(DEV).train()
        relu = relu.train()
        xi = x.clone().requires_grad_()
        prev = Q.set_noise_provider(lambda shape: noise.clone())
        try:
            y = relu(bn(xi))
            lib.profile_begin()
            y.backward(go)
            names = [r[0] for r in lib.profile_end()]
        finally:
            Q.set_noise_provider(prev)
        res.append((xi.grad.clone(), bn.weight.grad.clone(), names))
    assert "qb_relu_bwd_minmax" in res[0][2] and "qb_minmax" not in res[0][2], res[0][2]
    assert "qb_minmax" in res[1][2], res[1][2]
    assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])


@pytest.mark.parametrize("shape", [(16, 64, 14, 14), (32, 24, 6, 6), (16, 1024, 2, 2), (16, 48, 5, 4)])
@pytest.mark.parametrize("with_residual", [False, True])
def test_rangebn_forward_relu_equals_the_composition(lib, shape, with_residual):
    """RangeBN.forward_relu (qb_rangebn_apply_relu_measure: normalise [+ residual] + ReLU + next layer's statistics in one
    pass over the codes) == relu(bn(x)) / relu.add_relu(bn(x), res) bit for bit: output, stashed statistics, running
    buffers, and - under the same injected noise - every gradient.  Channel counts that divide the kernel's 1024-element
    thread row (64, 1024) and ones that do not (24, 48)."""
    from qtorch_b200 import glue, quantize as Q
    torch.manual_seed(11)
    C = shape[1]
    x = (torch.randn(shape, device=DEV) * 2).contiguous(memory_format=torch.channels_last)
    r = torch.randn(shape, device=DEV).contiguous(memory_format=torch.channels_last)
    go = torch.randn(shape, device=DEV).contiguous(memory_format=torch.channels_last)
    noise = torch.rand(shape, device=DEV) - 0.5
    state = Q.RangeBN(C).state_dict()
    res = []
    for fused in (True, False):
        bn = Q.RangeBN(C)
        bn.load_state_dict(state)
        bn = bn.to(DEV).train()
        relu = glue.ReLU(inplace=True).train()
        xi, ri = x.clone().requires_grad_(), r.clone().requires_grad_()
        prev = Q.set_noise_provider(lambda shp: noise.clone())
        try:
            lib.profile_begin()
            if fused:
                y = bn.forward_relu(xi, relu, ri if with_residual else None)
            else:
                y = relu.add_relu(bn(xi), ri) if with_residual else relu(bn(xi))
            names = [n[0] for n in lib.profile_end()]
            stats = Q._take_fwd_stats(y.detach())
            out = y.detach().clone()
            lib.profile_begin()
            y.backward(go)
            bwd_names = [n[0] for n in lib.profile_end()]
        finally:
            Q.set_noise_provider(prev)
        res.append(dict(out=out, stats=stats.clone(), names=names, bwd_names=bwd_names, dx=xi.grad.clone(), dr=ri.grad.clone() if with_residual else None,
                        dgamma=bn.weight.grad.clone(), dbeta=bn.bias.grad.clone(),
                        buffers={k: v.clone() for k, v in bn.state_dict().items()}))
    f, c = res
    assert "qb_rangebn_apply_relu_measure" in f["names"] and "qb_rangebn_apply_nhwc" not in f["names"], f["names"]
    assert "qb_rangebn_apply_nhwc" in c["names"]
    # backward of the fused ReLU from its 1-bit mask; without a residual the masked gradient is never written
    assert "qb_relu_bwd_minmax_bits" in f["bwd_names"] and "qb_relu_bwd_minmax" not in f["bwd_names"], f["bwd_names"]
    assert "qb_relu_bwd_minmax" in c["bwd_names"]
    assert torch.equal(f["out"], c["out"])
    assert torch.equal(f["stats"], c["stats"]), (f["stats"], c["stats"])
    for k in ("dx", "dgamma", "dbeta") + (("dr",) if with_residual else ()):
        assert torch.equal(f[k], c[k]), k
    for k, v in c["buffers"].items():
        assert torch.equal(f["buffers"][k], v), k


@pytest.mark.parametrize("shape", [(4, 64, 56, 56), (2, 16, 9, 11), (3, 8, 4, 4)])
def test_relu_maxpool_fused_backward(lib, shape):
    """glue.ReLU -> glue.MaxPool2d(3, 2, 1): the pooling backward yields the gradient w.r.t. the ReLU input directly (one
    kernel, with its range); values bit-identical to nn.ReLU -> nn.MaxPool2d."""
    from qtorch_b200 import glue, quantize as Q
    gen = torch.Generator().manual_seed(sum(shape) + 5)
    x0 = torch.randn(shape, generator=gen).to(DEV).contiguous(memory_format=torch.channels_last)
    xa, xb = x0.clone().requires_grad_(), x0.clone().requires_grad_()
    relu, pool = glue.ReLU(inplace=True).train(), glue.MaxPool2d(3, 2, 1)
    ya = pool(relu(xa * 1.0))
    yb = torch.nn.functional.max_pool2d(torch.relu(xb * 1.0), 3, 2, 1)
    assert torch.equal(ya, yb)
    go = torch.randn(ya.shape, generator=gen).to(DEV).contiguous(memory_format=torch.channels_last)
    lib.profile_begin()
    ya.backward(go)
    names = [r[0] for r in lib.profile_end()]
    yb.backward(go)
    assert "qb_maxpool2d_relu_nhwc_bwd" in names and "qb_relu_bwd_minmax" not in names and "qb_maxpool2d_nhwc_bwd" not in names, names
    assert torch.equal(xa.grad, xb.grad)
    # the range of that gradient was left for the consumer
    ent = Q._grad_ranges
    assert len(ent) >= 1
    rng = list(ent.values())[-1][2]
    assert float(rng[0]) == float(xb.grad.min()) and float(rng[1]) == float(xb.grad.max())
    # a second consumer of the ReLU output: autograd sums, the ReLU backward masks the sum - still torch's result
    xa2, xb2 = x0.clone().requires_grad_(), x0.clone().requires_grad_()
    ra = relu(xa2 * 1.0)
    (pool(ra).sum() + (ra * 0.5).sum()).backward()
    rb = torch.relu(xb2 * 1.0)
    (torch.nn.functional.max_pool2d(rb, 3, 2, 1).sum() + (rb * 0.5).sum()).backward()
    assert torch.allclose(xa2.grad, xb2.grad, rtol=0, atol=1e-6)


@pytest.mark.parametrize("shape", [(4, 64, 56, 56), (2, 16, 9, 11), (3, 8, 4, 4)])
def test_relu_pool_without_the_relu_output(lib, shape):
    """glue.MaxPool2d.relu_pool (the stem tail of models.py): maxpool(relu(x)) evaluated as relu(maxpool(x)) with the
    one-pass backward - values, gradient and the stashed statistics / gradient range equal torch's composition."""
    from qtorch_b200 import glue, quantize as Q
    gen = torch.Generator().manual_seed(sum(shape) + 9)
    x0 = torch.randn(shape, generator=gen)
    x0[0, :, :3, :3] = -x0[0, :, :3, :3].abs()                      # windows without a positive element
    x0[1, 0, 0, 0] = 0.0
    x0 = x0.to(DEV).contiguous(memory_format=torch.channels_last)
    xa, xb = x0.clone().requires_grad_(), x0.clone().requires_grad_()
    relu, pool = glue.ReLU(inplace=True).train(), glue.MaxPool2d(3, 2, 1)
    lib.profile_begin()
    ya = pool.relu_pool(xa * 1.0, relu)
    names = [r[0] for r in lib.profile_end()]
    assert names == ["qb_maxpool2d_nhwc_fwd", "qb_relu_measure"], names
    yb = torch.nn.functional.max_pool2d(torch.relu(xb * 1.0), 3, 2, 1)
    assert torch.equal(ya, yb)
    got = Q._take_fwd_stats(ya.detach())
    assert got is not None and torch.equal(got, lib.measure_stats(yb.detach().contiguous(memory_format=torch.channels_last)))
    go = torch.randn(ya.shape, generator=gen).to(DEV).contiguous(memory_format=torch.channels_last)
    ya.backward(go)
    yb.backward(go)
    assert torch.equal(xa.grad, xb.grad)
    rng = list(Q._grad_ranges.values())[-1][2]
    assert float(rng[0]) == float(xb.grad.min()) and float(rng[1]) == float(xb.grad.max())
    # eval / no-grad: plain composition
    with torch.no_grad():
        assert torch.equal(pool.relu_pool(x0.clone(), relu), torch.nn.functional.max_pool2d(torch.relu(x0), 3, 2, 1))


def test_l2_regularization_matches_torch(lib):
    """glue.regularization == the reference's model.regularization() (resnet_quantized.py:157-165), value and gradients."""
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import glue
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3), torch.nn.Conv2d(8, 16, 1, bias=False), torch.nn.Flatten(),
                              torch.nn.Linear(16, 10)).to(DEV)
    wd = 1e-4
    ref = 0
    for m in net.modules():
        if isinstance(m, (torch.nn.Conv2d, torch.nn.Linear)):
            ref = ref + m.weight.double().pow(2).sum()
            if m.bias is not None:
                ref = ref + m.bias.double().pow(2).sum()
    ref = wd * 0.5 * ref
    val = glue.regularization(net, wd)
    assert abs(float(val) - float(ref)) <= 1e-6 * abs(float(ref))
    (3.0 * val).backward()
    for p in net.parameters():
        assert torch.allclose(p.grad, 3.0 * wd * p.detach(), rtol=1e-6, atol=0)


def test_fused_sgd_matches_torch_bitwise(lib):
    """glue.SGD (qb_sgd_step) against torch.optim.SGD over 4 steps: momentum, weight decay, mixed tensor sizes (several
    chunks per tensor, scalars), same state-dict layout."""
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import glue
    gen = torch.Generator().manual_seed(5)
    shapes = [(64, 3, 7, 7), (64,), (512, 512, 3, 3), (1000, 512), (1000,), (1,), (17, 5)]
    for momentum, wd in ((0.9, 0.0), (0.9, 1e-4), (0.0, 5e-4)):
        pa = [torch.randn(s, generator=gen).to(DEV).requires_grad_() for s in shapes]
        pb = [p.detach().clone().requires_grad_() for p in pa]
        oa = glue.SGD(pa, lr=0.1, momentum=momentum, weight_decay=wd)
        ob = torch.optim.SGD(pb, lr=0.1, momentum=momentum, weight_decay=wd)
        for step in range(4):
            for a, b in zip(pa, pb):
                g = torch.randn(a.shape, generator=gen).to(DEV) * 0.01
                a.grad, b.grad = g.clone(), g.clone()
            oa.step()
            ob.step()
            for a, b in zip(pa, pb):
                assert torch.equal(a, b), (momentum, wd, step, tuple(a.shape), float((a - b).abs().max()))
        if momentum:
            for a, b in zip(pa, pb):
                assert torch.equal(oa.state[a]["momentum_buffer"], ob.state[b]["momentum_buffer"])
        assert sorted(oa.state_dict()["state"].keys()) == sorted(ob.state_dict()["state"].keys())


@pytest.mark.parametrize("shape,k", [((256, 512, 7, 7), 7), ((16, 1024, 7, 7), 7), ((8, 64, 8, 8), 2), ((4, 32, 9, 10), 3), ((2, 8, 4, 4), 4)])
def test_avgpool_matches_torch(lib, shape, k):
    """glue.AvgPool2d (qb_avgpool2d_nhwc_fwd/_bwd) against torch.nn.AvgPool2d on the same channels-last tensor: forward within
    1 ulp (same raster-order sum, one division), backward bit-exact; ragged borders (floor mode) get zero gradient."""
    import torch.nn as nn
    from qtorch_b200 import glue
    gen = torch.Generator().manual_seed(1)
    x = torch.randn(shape, generator=gen).to(DEV).contiguous(memory_format=torch.channels_last)
    xa, xb = x.clone().requires_grad_(), x.clone().requires_grad_()
    lib.profile_begin()
    ya = glue.AvgPool2d(k)(xa)
    names = [r[0] for r in lib.profile_end()]
    assert "qb_avgpool2d_nhwc_fwd" in names
    yb = nn.AvgPool2d(k)(xb)
    assert ya.shape == yb.shape and ya.is_contiguous(memory_format=torch.channels_last)
    assert float((ya - yb).abs().max()) <= 1.2e-7 * float(yb.abs().max())
    g = torch.randn(yb.shape, generator=gen).to(DEV)
    ya.backward(g)
    yb.backward(g)
    assert torch.equal(xa.grad, xb.grad)
    # other geometries keep torch's path
    z = glue.AvgPool2d(3, stride=2, padding=1)(x)
    assert torch.equal(z, nn.AvgPool2d(3, stride=2, padding=1)(x))


@pytest.mark.parametrize("B,C", [(256, 1000), (16, 10), (37, 1001), (5, 7), (1024, 32)])
def test_cross_entropy_matches_torch(lib, B, C):
    """glue.CrossEntropyLoss (qb_cross_entropy_fwd/_bwd) against torch.nn.CrossEntropyLoss on CUDA: loss to 2e-7 relative,
    gradient to 1e-6 of its largest entry (same log-softmax recipe; the row sum is the only reordering); ignore_index rows
    contribute nothing and are not counted; an upstream gradient scales the result."""
    import torch.nn as nn
    from qtorch_b200 import glue
    gen = torch.Generator().manual_seed(B + C)
    logits = (torch.randn(B, C, generator=gen) * 3).to(DEV)
    target = torch.randint(0, C, (B,), generator=gen).to(DEV)
    for ignore in (False, True):
        t = target.clone()
        if ignore:
            t[::3] = -100
        la, lb = logits.clone().requires_grad_(), logits.clone().requires_grad_()
        lib.profile_begin()
        a = glue.CrossEntropyLoss()(la, t)
        names = [r[0] for r in lib.profile_end()]
        assert "qb_cross_entropy_fwd" in names
        b = nn.CrossEntropyLoss()(lb, t)
        assert a.shape == b.shape == ()
        assert abs(float(a) - float(b)) <= 2e-7 * abs(float(b)) + 1e-7
        (a * 1.7).backward()
        (b * 1.7).backward()
        assert float((la.grad - lb.grad).abs().max()) <= 1e-6 * float(lb.grad.abs().max())
        if ignore:
            assert float(la.grad[::3].abs().max()) == 0.0
    # configurations outside the reference's use take torch's path and give torch's answer
    w = torch.rand(C, generator=gen).to(DEV)
    assert torch.equal(glue.CrossEntropyLoss(weight=w)(logits, target), nn.CrossEntropyLoss(weight=w)(logits, target))
    assert torch.equal(glue.CrossEntropyLoss(reduction="sum")(logits, target), nn.CrossEntropyLoss(reduction="sum")(logits, target))


@pytest.mark.parametrize("shape", [(8, 64, 28, 28), (4, 16, 9, 11), (2, 64, 112, 112), (3, 8, 5, 4)])
def test_rangebn_forward_relu_pool_equals_the_composition(lib, shape):
    """RangeBN.forward_relu_pool (qb_rangebn_apply_maxpool_nhwc: 3x3/2 max-pooling straight from the codes of the statistics pass,
    the normalised tensor never written) == maxpool.relu_pool(bn(x), relu) bit for bit: output, stashed statistics of the next
    layer, running buffers and - under the same injected noise - every gradient; and the kernel alone against pooling the
    materialised normalisation (values and window indices)."""
    from qtorch_b200 import glue, quantize as Q
    torch.manual_seed(7)
    B, C, H, W = shape
    x = (torch.randn(shape, device=DEV) * 2 + 0.3).contiguous(memory_format=torch.channels_last)
    # --- kernel level
    HW = H * W
    gamma, beta = torch.rand(C, device=DEV) + 0.5, torch.randn(C, device=DEV)
    rm, rv = torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
    qr = torch.tensor([float(x.min()) * 0.9, float(x.max()) * 0.8], device=DEV)
    xc = torch.empty(x.numel(), dtype=torch.uint8, device=DEV)
    K = 16 if (B * HW) % 16 == 0 else 1
    red, amax, amin = lib.rangebn_stats(x, B, C, HW, K, nhwc=True, q_range=qr, q_bits=8, x_codes_out=xc)
    vec = lib.rangebn_finalize(red, C, K, B * HW, B * HW / K, 1e-5, 0.1, True, 8, gamma, beta, rm, rv)
    # the packed-byte window search (maximum of the codes, one normalisation per channel) against pooling the materialised
    # normalisation: positive and negative slopes, a channel with gamma == 0 (every element equal: the first one wins) and one
    # whose normalised values collapse onto beta (that thread must take the element-wise form) - set in the quantized vectors
    vec = vec.clone()
    vec[3, 1] = -vec[3, 1]
    vec[3, 2] = 0.0
    vec[3, 3], vec[4, 3] = 1e-9, 1.0
    vec[3, 5] = -0.0
    for has_gamma, has_beta in ((True, True), (True, False), (False, True), (False, False)):
        full = lib.rangebn_apply(x, B, C, HW, vec, has_gamma, has_beta, nhwc=True, q_range=qr, q_bits=8, x_codes=xc)
        want, widx = lib.maxpool2d_nhwc_fwd(full, 3, 2, 1)
        got, gidx = lib.rangebn_apply_maxpool(xc, shape, vec, has_gamma, has_beta, qr, 8, 3, 2, 1)
        assert torch.equal(got.view(torch.int32), want.view(torch.int32)) and torch.equal(gidx, widx), (has_gamma, has_beta)
    # --- module level
    if (B * HW) % 16 != 0:
        return
    go_shape = want.shape
    go = torch.randn(go_shape, device=DEV).contiguous(memory_format=torch.channels_last)
    noise = torch.rand(shape, device=DEV) - 0.5
    state = Q.RangeBN(C).state_dict()
    res = []
    for fused in (True, False):
        bn = Q.RangeBN(C)
        bn.load_state_dict(state)
        bn = bn.to(DEV).train()
        relu, pool = glue.ReLU(inplace=True).train(), glue.MaxPool2d(3, 2, 1)
        xi = x.clone().requires_grad_()
        prev = Q.set_noise_provider(lambda shp: noise.clone())
        Q.FUSE_BN_POOL = fused
        try:
            lib.profile_begin()
            y = bn.forward_relu_pool(xi, relu, pool)
            names = [n[0] for n in lib.profile_end()]
            stats = Q._take_fwd_stats(y.detach())
            out = y.detach().clone()
            y.backward(go)
        finally:
            Q.FUSE_BN_POOL = True
            Q.set_noise_provider(prev)
        res.append(dict(out=out, stats=stats.clone(), names=names, dx=xi.grad.clone(), dgamma=bn.weight.grad.clone(),
                        dbeta=bn.bias.grad.clone(), buffers={k: v.clone() for k, v in bn.state_dict().items()}))
    f, c = res
    assert "qb_rangebn_apply_maxpool_nhwc" in f["names"] and "qb_rangebn_apply_nhwc" not in f["names"], f["names"]
    assert "qb_rangebn_apply_nhwc" in c["names"] and "qb_maxpool2d_nhwc_fwd" in c["names"], c["names"]
    assert torch.equal(f["out"], c["out"]) and torch.equal(f["stats"], c["stats"])
    for k in ("dx", "dgamma", "dbeta"):
        assert torch.equal(f[k], c[k]), k
    for k, v in c["buffers"].items():
        assert torch.equal(f["buffers"][k], v), k
