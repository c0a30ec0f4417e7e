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
