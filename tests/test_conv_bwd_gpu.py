"""Native backward contractions (conv_bwd.cu: qb_bwd_pack, qb_wt_pack, qb_conv_dgrad, qb_conv_wgrad) and the generic direct
kernels (conv_generic.cu) against fp64 references - SURVEY.md section 8 row A7 (models/modules/quantize.py:142-156, 342-346)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda"

CASES = [  # B, Cin, H, W, Cout, k, stride, pad
    (2, 64, 16, 16, 64, 1, 1, 0),
    (4, 64, 56, 56, 64, 3, 1, 1),        # ResNet-18 layer1
    (3, 64, 56, 56, 128, 3, 2, 1),       # stride 2: four parity classes
    (2, 64, 56, 56, 128, 1, 2, 0),       # 1x1 downsample: three classes are zero
    (4, 128, 28, 28, 128, 3, 1, 1),
    (4, 128, 28, 28, 256, 3, 2, 1),
    (5, 256, 14, 14, 256, 3, 1, 1),      # whole padded images per strip
    (5, 256, 14, 14, 512, 1, 2, 0),
    (7, 512, 7, 7, 512, 3, 1, 1),
    (2, 32, 112, 112, 64, 1, 1, 0),      # MobileNet pointwise
    (3, 512, 14, 14, 512, 1, 1, 0),
    (2, 1024, 7, 7, 1024, 1, 1, 0),
    (8, 16, 32, 32, 16, 3, 1, 1),        # ResNet-20 / CIFAR (32- and 64-byte operand rows)
    (8, 16, 32, 32, 32, 3, 2, 1),
    (8, 16, 32, 32, 32, 1, 2, 0),
    (8, 32, 16, 16, 64, 3, 2, 1),
    (2, 32, 20, 150, 48, 3, 1, 1),       # ragged everything
    (1, 96, 9, 11, 80, 3, 2, 1),
    (256, 512, 1, 1, 1000, 1, 1, 0),     # ResNet-18 fc
    (3, 64, 300, 5, 64, 3, 1, 1),        # strips taller than wide
]


def rel(a, b):
    return float((a.double().cpu() - b.double()).norm() / b.double().norm().clamp_min(1e-300))


def test_bwd_pack_codes_bit_exact_and_split(lib):
    torch.manual_seed(0)
    for n in (8 * 1000, 8 * 1000 + 5, 1 << 20):
        g = torch.randn(n, device=DEV) * 1e-3
        g[::7] = 0
        noise = torch.rand(n, device=DEV) - 0.5
        rng = lib.minmax(g)
        codes = torch.empty(n, dtype=torch.uint8, device=DEV)
        ref = lib.fakequant_grad(g, 8, range_dev=rng, noise=noise, stochastic=True, codes=codes)
        gq, hi, lo = lib.bwd_pack(g, rng, 8, noise=noise, stochastic=True)
        lo_, hi_ = float(rng[0]), float(rng[1])
        scale = max((hi_ - lo_) / 255.0, 1e-8)
        zp = int(min(max(0.0 - lo_ / scale, 0.0), 255.0))
        # the bf16 operand holds exactly (q - zp) of the reference's quantizer (quantize.py:77-98) with the same noise
        assert torch.equal(gq.float(), codes.float() - zp)
        assert torch.equal(gq.float() * torch.tensor(scale, dtype=torch.float64).float().to(DEV), ref)
        # hi + lo reproduces g to 2^-16
        assert float(((hi.double() + lo.double()) - g.double()).abs().max() / g.abs().max()) < 2.0 ** -16
        # Philox noise: same lattice, codes within one step
        gq2, _, _ = lib.bwd_pack(g, rng, 8, stochastic=True, seed=123, want_split=False)
        assert float((gq2.float() - (codes.float() - zp)).abs().max()) <= 1.0


@pytest.mark.parametrize("case", CASES)
def test_conv_dgrad(lib, case):
    B, Cin, H, W, Cout, k, stride, pad = case
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    gen = torch.Generator().manual_seed(sum(case))
    a = torch.randint(-255, 256, (B, Cout, Ho, Wo), generator=gen).float()
    a[torch.rand(a.shape, generator=gen) < 0.3] = 0
    w = torch.randn(Cout, Cin, k, k, generator=gen) * 0.05
    ad = a.to(DEV).contiguous(memory_format=torch.channels_last).to(torch.bfloat16)
    wt = lib.wt_pack(w.to(DEV))
    dx = lib.conv_dgrad(ad, None, wt, (B, Cin, H, W), Cout, k, k, stride, pad, scale_mul=0.5)
    ref = 0.5 * torch.nn.grad.conv2d_input((B, Cin, H, W), w.double(), a.double(), stride, pad)
    assert rel(dx, ref) < 2e-5, rel(dx, ref)                     # north_star bar: 1e-4
    # generic fp32 gradient: two planes on the A side as well (hi*hi + hi*lo + lo*hi)
    gf = torch.randn(B, Cout, Ho, Wo, generator=gen) * 1e-3
    gd = gf.to(DEV).contiguous(memory_format=torch.channels_last)
    _, hi, lo = lib.bwd_pack(gd, want_codes=False)
    dx2 = lib.conv_dgrad(hi, lo, wt, (B, Cin, H, W), Cout, k, k, stride, pad)
    ref2 = torch.nn.grad.conv2d_input((B, Cin, H, W), w.double(), gf.double(), stride, pad)
    assert rel(dx2, ref2) < 3e-5, rel(dx2, ref2)
    # accumulating form (the skip gradient of a residual block is already in the output): bit-identical to dx + skip, whether
    # the reduce-add kernel takes the geometry or the wrapper falls back to a separate add
    skip = torch.randn(B, Cin, H, W, generator=gen).to(DEV).contiguous(memory_format=torch.channels_last)
    want = dx + skip
    lib.profile_begin()
    got = lib.conv_dgrad(ad, None, wt, (B, Cin, H, W), Cout, k, k, stride, pad, scale_mul=0.5, accumulate_into=skip)
    names = [r[0] for r in lib.profile_end()]
    assert got.data_ptr() == skip.data_ptr()
    assert torch.equal(got, want), rel(got, want.cpu())
    if case in ((4, 64, 56, 56, 64, 3, 1, 1), (4, 128, 28, 28, 128, 3, 1, 1), (7, 512, 7, 7, 512, 3, 1, 1)):   # ResNet-18 identity blocks
        assert names == ["qb_conv_dgrad_acc"], names


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("lo_val", [0.0, -0.37])
def test_conv_wgrad(lib, case, lo_val):
    B, Cin, H, W, Cout, k, stride, pad = case
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    gen = torch.Generator().manual_seed(sum(case) + 1)
    q = torch.randint(0, 256, (B, Cin, H, W), generator=gen).float()
    g = torch.randn(B, Cout, Ho, Wo, generator=gen) * 1e-3
    hi_val = lo_val + 2.0
    rng = torch.tensor([lo_val, hi_val], device=DEV)
    scale32 = torch.tensor(max((hi_val - lo_val) / 255.0, 1e-8), dtype=torch.float64).float().double().item()
    lo32 = torch.tensor(lo_val).float().double().item()
    x_hat = q.double() * scale32 + lo32                          # zero padding stays TRUE zero: the la term is border-aware
    gd = g.to(DEV).contiguous(memory_format=torch.channels_last)
    x16 = q.to(DEV).contiguous(memory_format=torch.channels_last).to(torch.bfloat16)
    _, ghi, glo = lib.bwd_pack(gd, want_codes=False)
    dw, db = lib.conv_wgrad(ghi, glo, gd, x16, Cout, k, k, stride, pad, rng, 8, want_bias=True)
    dw2, _ = lib.conv_wgrad(ghi, glo, gd, x16, Cout, k, k, stride, pad, rng, 8, want_bias=True)
    assert torch.equal(dw, dw2)                                   # split-K with a fixed-order reduction: deterministic
    ref = torch.nn.grad.conv2d_weight(x_hat, (Cout, Cin, k, k), g.double(), stride, pad)
    assert rel(dw, ref) < 2e-5, rel(dw, ref)
    assert rel(db, g.double().sum((0, 2, 3))) < 2e-6


def test_unsupported_geometry_is_an_error_not_a_fallback(lib):
    a = torch.zeros(2, 12, 8, 8, device=DEV, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    wt = torch.zeros(2, 16, 9, 12, device=DEV, dtype=torch.bfloat16)
    with pytest.raises(lib.QB200Error):
        lib.conv_dgrad(a, None, wt, (2, 16, 8, 8), 12, 3, 3, 1, 1)          # C_out % 8 != 0


GENERIC = [  # B, Cin, H, W, Cout, (kh, kw), (sh, sw), (ph, pw), (dh, dw), groups, nhwc
    (2, 3, 9, 11, 8, (3, 3), (1, 1), (1, 1), (1, 1), 1, False),
    (2, 4, 9, 11, 6, (3, 3), (2, 2), (1, 1), (1, 1), 1, True),
    (3, 6, 8, 8, 6, (3, 3), (1, 1), (1, 1), (1, 1), 6, False),     # depthwise with C % 4 != 0
    (2, 8, 10, 7, 12, (3, 2), (2, 1), (0, 1), (1, 1), 2, True),    # grouped, rectangular everything
    (2, 5, 12, 12, 7, (3, 3), (1, 1), (2, 2), (2, 2), 1, False),   # dilation
    (2, 16, 9, 9, 16, (5, 5), (1, 1), (2, 2), (1, 1), 1, True),    # 25 taps
]


@pytest.mark.parametrize("case", GENERIC)
def test_generic_direct_kernels(lib, case):
    B, Cin, H, W, Cout, k, st, pd, dl, groups, nhwc = case
    gen = torch.Generator().manual_seed(sum(case[:5]))
    x = torch.randn(B, Cin, H, W, generator=gen)
    w = torch.randn(Cout, Cin // groups, k[0], k[1], generator=gen) * 0.1
    b = torch.randn(Cout, generator=gen)
    x64 = x.double().requires_grad_()
    w64 = w.double().requires_grad_()
    b64 = b.double().requires_grad_()
    ref = F.conv2d(x64, w64, b64, st, pd, dl, groups)
    go = torch.randn(ref.shape, generator=gen)
    ref.backward(go.double())
    fmt = torch.channels_last if nhwc else torch.contiguous_format
    xd = x.to(DEV).contiguous(memory_format=fmt)
    conv = (st, pd, dl, groups)
    out = lib.conv_generic_fwd(xd, w.to(DEV), b.to(DEV), conv)
    assert rel(out, ref.detach()) < 1e-6
    god = go.to(DEV).contiguous(memory_format=fmt)
    dx = lib.conv_generic_dgrad(god, w.to(DEV), (B, Cin, H, W), conv, like=xd)
    dw, db = lib.conv_generic_wgrad(god, xd, tuple(w.shape), conv, want_w=True, want_bias=True)
    assert rel(dx, x64.grad) < 1e-6 and rel(dw, w64.grad) < 1e-6 and rel(db, b64.grad) < 1e-6


def test_generic_linear(lib):
    torch.manual_seed(1)
    x, w, b = torch.randn(7, 24), torch.randn(10, 24), torch.randn(10)
    out = lib.conv_generic_fwd(x.to(DEV), w.to(DEV), b.to(DEV), None)
    assert rel(out, F.linear(x.double(), w.double(), b.double())) < 1e-6
    go = torch.randn(7, 10)
    dx = lib.conv_generic_dgrad(go.to(DEV), w.to(DEV), (7, 24), None)
    dw, db = lib.conv_generic_wgrad(go.to(DEV), x.to(DEV), (10, 24), None, want_w=True, want_bias=True)
    assert rel(dx, go.double() @ w.double()) < 1e-6
    assert rel(dw, go.double().t() @ x.double()) < 1e-6 and rel(db, go.double().sum(0)) < 1e-6


STEM16 = [  # B, H, W, Cout, k, stride, pad
    (4, 224, 224, 64, 7, 2, 3),          # ResNet conv1
    (2, 62, 46, 64, 7, 2, 3),            # ragged tiles
    (1, 8, 8, 64, 7, 2, 3),              # every pixel on the border
    (3, 64, 64, 32, 3, 2, 1),            # MobileNet stem
    (2, 224, 224, 32, 3, 2, 1),
]


@pytest.mark.parametrize("case", STEM16)
def test_conv_stem_wgrad16(lib, case):
    """First-layer weight gradient through the overlapping-window tensor map (qb_conv_stem_wgrad16) against fp64: image inputs
    have min != 0, so the border-aware zero-padding term (the 'ones' channel of the virtual operand) is exercised."""
    B, H, W, Cout, k, stride, pad = case
    gen = torch.Generator().manual_seed(sum(case) + 3)
    x = torch.randn(B, 3, H, W, generator=gen) * 1.1 + 0.3
    xd = x.to(DEV).contiguous(memory_format=torch.channels_last)
    rng = torch.tensor([float(x.min()), float(x.max())], device=DEV)
    codes = torch.empty_like(xd, dtype=torch.uint8)
    xq = lib.fakequant_affine(xd, 8, range_dev=rng, codes=codes)
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    g = torch.randn(B, Cout, Ho, Wo, generator=gen)
    gd = g.to(DEV).contiguous(memory_format=torch.channels_last)
    _, hi, lo = lib.bwd_pack(gd, want_codes=False)
    assert lib.conv_stem_wgrad16_supported(B, H, W, Cout, k, stride, pad)
    dw = lib.conv_stem_wgrad16(hi, lo, codes, rng, 8, Cout, k, stride, pad)
    dw2 = lib.conv_stem_wgrad16(hi, lo, codes, rng, 8, Cout, k, stride, pad)
    assert torch.equal(dw, dw2)
    ref = torch.nn.grad.conv2d_weight(xq.double().cpu(), (Cout, 3, k, k), g.double(), stride, pad)
    assert rel(dw, ref) < 2e-5, rel(dw, ref)
