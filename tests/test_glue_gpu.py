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
