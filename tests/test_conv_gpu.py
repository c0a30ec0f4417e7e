"""int8 tcgen05 implicit-GEMM convolution (qb_conv_i8_fwd) against fp64 / the reference's fp32 F.conv2d on the same
fake-quantized tensors: shapes of every ResNet-18 / MobileNet dense conv class, both signs of input range (the
border-aware zero-padding terms only matter when min != 0)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda"

CASES = [  # B, Cin, H, W, Cout, k, stride, pad
    (2, 64, 56, 56, 64, 3, 1, 1),
    (3, 64, 56, 56, 128, 3, 2, 1),
    (2, 64, 56, 56, 128, 1, 2, 0),
    (4, 128, 28, 28, 128, 3, 1, 1),
    (4, 128, 28, 28, 256, 3, 2, 1),
    (5, 256, 14, 14, 256, 3, 1, 1),
    (5, 256, 14, 14, 512, 1, 2, 0),
    (7, 512, 7, 7, 512, 3, 1, 1),
    (2, 32, 112, 112, 64, 1, 1, 0),
    (2, 32, 20, 150, 48, 3, 1, 1),
    (1, 96, 9, 11, 80, 3, 2, 1),
    (3, 64, 8, 8, 64, 3, 1, 1),
]


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("relu", [True, False])
@pytest.mark.parametrize("use_bias", [False, True])
def test_conv_i8_fwd(lib, case, relu, use_bias):
    B, Cin, H, W, Cout, k, stride, pad = case
    gen = torch.Generator().manual_seed(sum(case))
    x = torch.randn(B, Cin, H, W, generator=gen)
    if relu:
        x = x.clamp_min(0)
    else:
        x = x + 0.7                                        # min far from zero: padding correction terms are large
    w = torch.randn(Cout, Cin, k, k, generator=gen) * 0.05
    bias = torch.randn(Cout, generator=gen) * 0.1 if use_bias else None
    xd = x.to(DEV).contiguous(memory_format=torch.channels_last)
    rng = torch.tensor([float(x.min()), float(x.max())], device=DEV)
    codes = torch.empty((B, Cin, H, W), dtype=torch.uint8, device=DEV).contiguous(memory_format=torch.channels_last)
    xq = lib.fakequant_affine(xd, 8, range_dev=rng, codes=codes)
    pixsum = lib.rowsum_u8(codes, B * H * W, Cin)
    rmin, rmax = torch.empty(Cout, device=DEV), torch.empty(Cout, device=DEV)
    wq, wcodes, tapsum = lib.fakequant_rows_conv(w.to(DEV), 8, rmin, rmax)
    # weight prep agrees with the plain row kernel
    wq2, rmin2, rmax2 = lib.fakequant_rows(w.to(DEV), 8)
    assert torch.equal(wq, wq2) and torch.equal(rmin, rmin2) and torch.equal(rmax, rmax2)
    bd = bias.to(DEV) if use_bias else None
    stats = torch.empty(8, dtype=torch.float64, device=DEV)
    out = lib.conv_i8_fwd(codes, wcodes, B, H, W, Cin, Cout, k, k, stride, pad, rng, 8, rmin, rmax, 8, pixsum, tapsum, bias=bd,
                          out_stats=stats)
    torch.cuda.synchronize()
    # fused output statistics (the next layer's QuantMeasure block): extrema exact, sums to fp64 accumulation accuracy
    ref_stats = lib.measure_stats(out)
    assert torch.equal(stats[[0, 1, 4, 5, 6, 7]], ref_stats[[0, 1, 4, 5, 6, 7]]), (stats, ref_stats)
    assert torch.allclose(stats[2:4], ref_stats[2:4], rtol=1e-6, atol=1e-6 * float(ref_stats[3].sqrt())), (stats, ref_stats)
    stats2 = torch.empty(8, dtype=torch.float64, device=DEV)
    lib.conv_i8_fwd(codes, wcodes, B, H, W, Cin, Cout, k, k, stride, pad, rng, 8, rmin, rmax, 8, pixsum, tapsum, bias=bd,
                    out_stats=stats2)
    assert torch.equal(stats, stats2)                      # deterministic, workspace left clean
    assert out.is_contiguous(memory_format=torch.channels_last)
    ref64 = F.conv2d(xq.double().cpu(), wq.double().cpu(), bias.double() if use_bias else None, stride, pad)
    assert out.shape == ref64.shape
    rel = float((out.cpu().double() - ref64).norm() / ref64.norm())
    assert rel < 2e-5, rel      # fp32 epilogue; the reference itself is ~1e-6 from fp64
    ref32 = F.conv2d(xq.cpu(), wq.cpu(), bias, stride, pad)       # what the reference computes (fp32, CPU)
    rel32 = float((out.cpu() - ref32).norm() / ref32.norm())
    assert rel32 < 1e-4, rel32


def test_conv_i8_unsupported_is_an_error(lib):
    x = torch.zeros((1, 3, 8, 8), dtype=torch.uint8, device=DEV)
    z = torch.zeros(64, device=DEV)
    zi = torch.zeros(64, dtype=torch.int32, device=DEV)
    with pytest.raises(lib.QB200Error):
        lib.conv_i8_fwd(x, x, 1, 8, 8, 3, 64, 3, 3, 1, 1, z, 8, z, z, 8, zi, zi)


STEM_CASES = [  # B, Cin, H, W, Cout, k, stride, pad
    (3, 3, 224, 224, 64, 7, 2, 3),       # ResNet conv1
    (2, 3, 61, 45, 64, 7, 2, 3),         # ragged tiles, odd sizes
    (2, 3, 32, 32, 64, 7, 1, 3),         # stride 1 (cifar-style stem with a 7x7)
    (2, 3, 224, 224, 32, 3, 2, 1),       # MobileNet first conv
    (3, 3, 33, 47, 96, 3, 1, 1),
    (1, 3, 7, 7, 64, 7, 2, 3),           # every output pixel is a border pixel
    (2, 4, 30, 30, 128, 3, 2, 0),        # 4 input channels, no padding
]


@pytest.mark.parametrize("case", STEM_CASES)
@pytest.mark.parametrize("use_bias", [False, True])
def test_conv_stem_fwd(lib, case, use_bias):
    """First-layer convolution (qb_conv_stem_fwd, tcgen05 over a thread-built im2col operand): image-like input (both signs, so the zero-padding terms are
    live) against fp64 and the reference's fp32 F.conv2d on the same fake-quantized tensors."""
    B, Cin, H, W, Cout, k, stride, pad = case
    assert lib.conv_stem_supported(H, W, Cin, Cout, k, k, stride, pad)
    gen = torch.Generator().manual_seed(sum(case))
    x = torch.randn(B, Cin, H, W, generator=gen) * 1.1 + 0.3
    w = torch.randn(Cout, Cin, k, k, generator=gen) * 0.08
    bias = torch.randn(Cout, generator=gen) * 0.1 if use_bias else None
    xd = x.to(DEV).contiguous(memory_format=torch.channels_last)
    rng = torch.tensor([float(x.min()), float(x.max())], device=DEV)
    codes = torch.empty_like(xd, dtype=torch.uint8)
    assert codes.is_contiguous(memory_format=torch.channels_last) or Cin == 1
    xq = lib.fakequant_affine(xd, 8, range_dev=rng, codes=codes)
    rmin, rmax = torch.empty(Cout, device=DEV), torch.empty(Cout, device=DEV)
    wq, wcodes, tapsum = lib.fakequant_rows_conv(w.to(DEV), 8, rmin, rmax)
    out = lib.conv_stem_fwd(codes, wcodes, B, H, W, Cin, Cout, k, k, stride, pad, rng, 8, rmin, rmax, 8, tapsum,
                            bias=bias.to(DEV) if use_bias else None)
    torch.cuda.synchronize()
    assert out.is_contiguous(memory_format=torch.channels_last)
    ref64 = F.conv2d(xq.double().cpu(), wq.double().cpu(), bias.double() if use_bias else None, stride, pad)
    assert out.shape == ref64.shape
    err = (out.cpu().double() - ref64)
    rel = float(err.norm() / ref64.norm())
    assert rel < 2e-5, rel
    # borders are as accurate as the interior (per-border-case constants, not an approximation)
    assert float(err.abs().max()) < 2e-5 * float(ref64.abs().max()) * 10
    ref32 = F.conv2d(xq.cpu(), wq.cpu(), bias, stride, pad)
    assert float((out.cpu() - ref32).norm() / ref32.norm()) < 1e-4


def test_conv_stem_unsupported_is_an_error(lib):
    assert not lib.conv_stem_supported(8, 8, 16, 64, 3, 3, 1, 1)
    assert not lib.conv_stem_supported(224, 224, 3, 256, 7, 7, 2, 3)
    assert not lib.conv_stem_supported(224, 224, 3, 24, 7, 7, 2, 3)          # Cout % 16
    x = torch.zeros((1, 16, 8, 8), dtype=torch.uint8, device=DEV)
    z = torch.zeros(64, device=DEV)
    zi = torch.zeros(64 * 9, dtype=torch.int32, device=DEV)
    with pytest.raises(lib.QB200Error):
        lib.conv_stem_fwd(x, x, 1, 8, 8, 16, 64, 3, 3, 1, 1, z, 8, z, z, 8, zi)


def test_qconv2d_first_layer_uses_stem_kernel(lib):
    """QConv2d(3, 64, 7, 2, 3) routes to the native kernel and matches the cuDNN route on the same weights."""
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import quantize as Q
    torch.manual_seed(5)
    m = Q.QConv2d(3, 64, 7, 2, 3, bias=False, num_bits=8, num_bits_weight=8, num_bits_grad=None).to(DEV)   # no (stochastic) gradient quantizer: deterministic
    x = torch.randn(4, 3, 64, 64, device=DEV)
    before = lib.LAUNCHES
    lib.profile_begin()
    y1 = m(x)
    names = [r[0] for r in lib.profile_end()]
    y1.square().mean().backward()
    g1 = m.weight.grad.clone()
    assert lib.LAUNCHES > before
    assert "qb_conv_stem_fwd" in names, names
    m.weight.grad = None
    m2 = Q.QConv2d(3, 64, 7, 2, 3, bias=False, num_bits=8, num_bits_weight=8, num_bits_grad=None).to(DEV)
    m2.load_state_dict(m.state_dict())
    Q.USE_STEM_KERNEL = False
    try:
        y2 = m2(x)
        y2.square().mean().backward()
    finally:
        Q.USE_STEM_KERNEL = True
    assert float((y1 - y2).detach().norm() / y2.detach().norm()) < 1e-4
    assert float((g1 - m2.weight.grad).norm() / m2.weight.grad.norm()) < 1e-3


@pytest.mark.parametrize("shape", [(4, 3, 64, 64), (3, 3, 30, 34), (2, 3, 224, 224)])
def test_first_layer_quantizer_reads_nchw(lib, shape):
    """qb_fakequant_affine_nchw3: the codes of the NCHW image batch, written channels-last, are the codes qb_fakequant_affine
    gives on the channels-last copy (bit-exact), and QConv2d's first-layer route through it (no layout copy of the batch) gives
    the same output, weight gradient and running range as the route through the copy."""
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import quantize as Q
    torch.manual_seed(11)
    x = torch.randn(shape, device=DEV) * 1.7 + 0.3
    rng = torch.tensor([float(x.min()) * 0.9, float(x.max()) * 0.8], device=DEV)     # clamps on both sides
    for bits in (8, 5):
        xl = x.contiguous(memory_format=torch.channels_last)
        want = torch.empty_like(xl, dtype=torch.uint8)
        lib.fakequant_affine(xl, bits, range_dev=rng, codes=want, want_y=False)
        got = lib.fakequant_affine_nchw3(x, bits, rng)
        assert got.is_contiguous(memory_format=torch.channels_last) and got.shape == x.shape
        assert torch.equal(got, want)
    outs = []
    for direct in (True, False):
        torch.manual_seed(5)
        m = Q.QConv2d(3, 64, 7, 2, 3, bias=False, num_bits=8, num_bits_weight=8, num_bits_grad=None).to(DEV)
        Q.STEM_NCHW_DIRECT = direct
        try:
            lib.profile_begin()
            y = m(x)
            names = [r[0] for r in lib.profile_end()]
            y.square().mean().backward()
        finally:
            Q.STEM_NCHW_DIRECT = True
        assert ("qb_fakequant_affine_nchw3" in names) == direct, names
        outs.append((y.detach(), m.weight.grad.clone(), m.quantize_input.running_min.clone(), m.quantize_input.running_max.clone()))
    for i, (a, b) in enumerate(zip(*outs)):
        if i == 1:                                      # weight gradient: same codes in, fp32 partial sums combined atomically
            assert float((a - b).norm() / b.norm()) < 1e-5
        else:
            assert torch.equal(a, b)
    # an input that wants its gradient, or an odd plane size, keeps the copying route
    m = Q.QConv2d(3, 64, 7, 2, 3, bias=False).to(DEV)
    assert not Q._stem_takes_nchw(m, x.clone().requires_grad_())
    assert not Q._stem_takes_nchw(m, torch.randn(2, 3, 7, 9, device=DEV))
    with pytest.raises(lib.QB200Error):
        lib.load().qb_fakequant_affine_nchw3.restype                                   # symbol present
        lib._run("qb_fakequant_affine_nchw3", 0, x.data_ptr(), want.data_ptr(), 2, 63, 8, rng.data_ptr(), None)


@pytest.mark.parametrize("rows,cols", [(7, 16), (33, 32), (1000, 64), (12544, 64), (3001, 128), (777, 256), (200704, 64)])
def test_quantizer_with_row_sums(lib, rows, cols):
    """qb_fakequant_affine_rowsum: codes, bf16 codes and per-row code sums in one pass == qb_fakequant_affine16 + qb_rowsum_u8
    (bit-exact); row counts that leave a partial last warp included."""
    gen = torch.Generator().manual_seed(rows + cols)
    x = (torch.randn(rows, cols, generator=gen) * 1.3 + 0.2).to(DEV)
    rng = torch.tensor([float(x.min()) * 0.85, float(x.max()) * 0.9], device=DEV)
    for bits in (8, 4):
        assert lib.fakequant_affine_rowsum_supported(cols, bits)
        want_c = torch.empty(rows, cols, dtype=torch.uint8, device=DEV)
        want_h = torch.empty(rows, cols, dtype=torch.bfloat16, device=DEV)
        lib.fakequant_affine(x, bits, range_dev=rng, codes=want_c, codes16=want_h, want_y=False)
        want_s = lib.rowsum_u8(want_c, rows, cols)
        for with16 in (True, False):
            got_c = torch.zeros(rows, cols, dtype=torch.uint8, device=DEV)
            got_h = torch.zeros(rows, cols, dtype=torch.bfloat16, device=DEV) if with16 else None
            got_s = lib.fakequant_affine_rowsum(x, rows, cols, bits, rng, got_c, got_h)
            assert torch.equal(got_c, want_c) and torch.equal(got_s, want_s)
            assert not with16 or torch.equal(got_h, want_h)
            assert int(got_s.sum()) == int(want_c.sum(dtype=torch.int64))
    assert not lib.fakequant_affine_rowsum_supported(512, 8) and not lib.fakequant_affine_rowsum_supported(48, 8)
    with pytest.raises(lib.QB200Error):
        lib.fakequant_affine_rowsum(x[:, :8].contiguous(), rows, 8, 8, rng, got_c)


def test_qconv2d_quantizer_row_sum_route(lib):
    """QConv2d on the fused quantizer + row-sum route gives bit-identical output and gradients to the two-call route."""
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import quantize as Q
    x = torch.randn(8, 64, 14, 14, device=DEV).contiguous(memory_format=torch.channels_last)
    outs = []
    for fused in (True, False):
        torch.manual_seed(3)
        m = Q.QConv2d(64, 128, 3, 1, 1, bias=False, num_bits=8, num_bits_weight=8, num_bits_grad=None).to(DEV)
        xi = x.clone().requires_grad_()
        Q.FUSE_QUANT_ROWSUM = fused
        try:
            lib.profile_begin()
            y = m(xi)
            names = [r[0] for r in lib.profile_end()]
            y.square().mean().backward()
        finally:
            Q.FUSE_QUANT_ROWSUM = True
        assert ("qb_fakequant_affine_rowsum" in names) == fused and ("qb_rowsum_u8" in names) != fused, names
        outs.append((y.detach(), xi.grad.clone(), m.weight.grad.clone()))
    for a, b in zip(*outs):
        assert torch.equal(a, b)


WGRAD_CASES = [  # B, Cin, H, W, Cout, k, stride, pad
    (4, 3, 224, 224, 64, 7, 2, 3),       # ResNet conv1 (more tiles than SMs: several tiles per CTA)
    (2, 3, 61, 45, 64, 7, 2, 3),         # ragged tiles
    (1, 3, 7, 7, 64, 7, 2, 3),           # every pixel on the border
    (3, 3, 64, 64, 32, 3, 2, 1),         # MobileNet stem
    (2, 3, 32, 32, 64, 7, 1, 3),
]


@pytest.mark.parametrize("case", WGRAD_CASES)
def test_conv_stem_wgrad(lib, case):
    """qb_conv_stem_wgrad (tf32 tensor cores on the codes) against the fp64 weight gradient of F.conv2d on the dequantised
    input; the bar is the tf32 rounding of the gradient operand (what cuDNN's default path does too)."""
    B, Cin, H, W, Cout, k, stride, pad = case
    gen = torch.Generator().manual_seed(sum(case) + 1)
    x = torch.randn(B, Cin, H, W, generator=gen) * 1.1 + 0.3
    xd = x.to(DEV).contiguous(memory_format=torch.channels_last)
    rng = torch.tensor([float(x.min()), float(x.max())], device=DEV)
    codes = torch.empty_like(xd, dtype=torch.uint8)
    xq = lib.fakequant_affine(xd, 8, range_dev=rng, codes=codes)
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    g = torch.randn(B, Cout, Ho, Wo, generator=gen)
    gd = g.to(DEV).contiguous(memory_format=torch.channels_last)
    dw = lib.conv_stem_wgrad(gd, codes, rng, 8, Cout, k, k, stride, pad)
    dw2 = lib.conv_stem_wgrad(gd, codes, rng, 8, Cout, k, k, stride, pad)
    torch.cuda.synchronize()
    assert torch.equal(dw, dw2)                                  # deterministic
    w64 = torch.zeros(Cout, Cin, k, k, dtype=torch.float64, requires_grad=True)
    F.conv2d(xq.double().cpu(), w64, None, stride, pad).backward(g.double())
    ref = w64.grad
    rel = float((dw.cpu().double() - ref).norm() / ref.norm())
    assert rel < 5e-4, rel                                       # tf32 operand rounding: ~2^-12 per term, random sign
    # same class as the library route this replaces
    torch.backends.cudnn.allow_tf32 = True
    w32 = torch.zeros(Cout, Cin, k, k, device=DEV, requires_grad=True)
    F.conv2d(xq, w32, None, stride, pad).backward(gd)
    rel_cudnn = float((w32.grad.cpu().double() - ref).norm() / ref.norm())
    assert rel < max(2.0 * rel_cudnn, 5e-4), (rel, rel_cudnn)      # tiny inputs: cuDNN picks an fp32 kernel, tf32 noise does not average out


DW_CASES = [  # B, C, H, W, stride, pad
    (2, 32, 112, 112, 1, 1),      # MobileNet dw1
    (2, 64, 112, 112, 2, 1),      # dw2 (stride 2)
    (3, 128, 56, 56, 1, 1),
    (2, 256, 28, 28, 2, 1),
    (4, 512, 14, 14, 1, 1),
    (5, 1024, 7, 7, 1, 1),        # half of the pixels on the border
    (3, 1024, 7, 7, 2, 1),
    (2, 24, 9, 13, 1, 1),         # channel groups do not divide the CTA: weights rebuilt on the fly
    (1, 8, 3, 3, 1, 1),           # every pixel a border pixel
    (2, 20, 10, 10, 1, 0),        # no padding
    (300, 8, 6, 6, 1, 1),         # many samples per CTA: per-sample extrema slots
]


@pytest.mark.parametrize("case", DW_CASES)
@pytest.mark.parametrize("relu", [True, False])
@pytest.mark.parametrize("use_bias", [False, True])
def test_conv_dw_fwd(lib, case, relu, use_bias):
    """qb_conv_dw_fwd (depthwise 3x3 on the codes) against F.conv2d(groups=C) on the same fake-quantized tensors in fp64
    and in the reference's fp32 (north_star: dequantised conv outputs within 1e-4 relative L2), with the fused output
    statistics against qb_measure_stats."""
    B, C, H, W, stride, pad = case
    gen = torch.Generator().manual_seed(sum(case))
    x = torch.randn(B, C, H, W, generator=gen)
    x = x.clamp_min(0) if relu else x + 0.7
    w = torch.randn(C, 1, 3, 3, generator=gen) * 0.2
    bias = torch.randn(C, generator=gen) * 0.1 if use_bias else None
    xd = x.to(DEV).contiguous(memory_format=torch.channels_last)
    rng = torch.tensor([float(x.min()), float(x.max())], device=DEV)
    codes = torch.empty((B, C, H, W), dtype=torch.uint8, device=DEV).contiguous(memory_format=torch.channels_last)
    xq = lib.fakequant_affine(xd, 8, range_dev=rng, codes=codes)
    rmin, rmax = torch.empty(C, device=DEV), torch.empty(C, device=DEV)
    wq, wcodes, _ = lib.fakequant_rows_conv(w.to(DEV), 8, rmin, rmax)
    bd = bias.to(DEV) if use_bias else None
    stats = torch.empty(8, dtype=torch.float64, device=DEV)
    out = lib.conv_dw_fwd(codes, wcodes, B, H, W, C, 3, 3, stride, pad, rng, 8, rmin, rmax, 8, bias=bd, out_stats=stats)
    out_plain = lib.conv_dw_fwd(codes, wcodes, B, H, W, C, 3, 3, stride, pad, rng, 8, rmin, rmax, 8, bias=bd)
    torch.cuda.synchronize()
    assert torch.equal(out, out_plain)
    assert out.is_contiguous(memory_format=torch.channels_last)
    ref_stats = lib.measure_stats(out)
    assert torch.equal(stats[[0, 1, 4, 5, 6, 7]], ref_stats[[0, 1, 4, 5, 6, 7]]), (stats, ref_stats)
    assert torch.allclose(stats[2:4], ref_stats[2:4], rtol=1e-6, atol=1e-6 * float(ref_stats[3].sqrt())), (stats, ref_stats)
    stats2 = torch.empty(8, dtype=torch.float64, device=DEV)
    lib.conv_dw_fwd(codes, wcodes, B, H, W, C, 3, 3, stride, pad, rng, 8, rmin, rmax, 8, bias=bd, out_stats=stats2)
    assert torch.equal(stats[[0, 1, 4, 5, 6, 7]], stats2[[0, 1, 4, 5, 6, 7]])          # workspace left clean
    ref64 = F.conv2d(xq.double().cpu(), wq.double().cpu(), bias.double() if use_bias else None, stride, pad, 1, C)
    assert out.shape == ref64.shape
    rel = float((out.cpu().double() - ref64).norm() / ref64.norm())
    assert rel < 2e-6, rel
    ref32 = F.conv2d(xq.cpu(), wq.cpu(), bias, stride, pad, 1, C)       # what the reference computes (fp32, CPU)
    rel32 = float((out.cpu() - ref32).norm() / ref32.norm())
    assert rel32 < 1e-5, rel32


def test_conv_dw_unsupported_is_an_error(lib):
    x = torch.zeros((1, 8, 8, 8), dtype=torch.uint8, device=DEV)
    z = torch.zeros(8, device=DEV)
    with pytest.raises(lib.QB200Error):
        lib.conv_dw_fwd(x, x, 1, 8, 8, 8, 5, 5, 1, 2, z, 8, z, z, 8)        # 5x5
    with pytest.raises(lib.QB200Error):
        lib.conv_dw_fwd(x, x, 1, 8, 8, 6, 3, 3, 1, 1, z, 8, z, z, 8)        # C % 4 != 0


def test_qconv2d_depthwise_uses_dw_kernel_and_matches_oracle(lib):
    """QConv2d(C, C, 3, groups=C) (mobilenet_quantized.py:24) routes to qb_conv_dw_fwd, hands its output statistics to the
    following QuantMeasure, and agrees with the CPU oracle module and with the F.conv2d route, forward and backward."""
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import quantize as Q
    from oracle import qref_modules
    torch.manual_seed(11)
    kw = dict(kernel_size=3, stride=2, padding=1, groups=64, bias=False, num_bits=8, num_bits_weight=8, num_bits_grad=None)
    ref = qref_modules.QConv2d(64, 64, **kw).train()
    m = Q.QConv2d(64, 64, **kw)
    m.load_state_dict(ref.state_dict())
    m = m.to(DEV).train()
    x = torch.randn(4, 64, 30, 30).clamp_min(0)
    xr = x.clone().requires_grad_(True)
    yr = ref(xr)
    yr.square().mean().backward()
    xd = x.to(DEV).requires_grad_(True)
    lib.profile_begin()
    y1 = m(xd)
    names = [r[0] for r in lib.profile_end()]
    assert "qb_conv_dw_fwd" in names, names
    assert Q._take_fwd_stats(y1) is not None          # the next layer's range reduction came with the output
    y1.square().mean().backward()
    assert float((y1.detach().cpu() - yr.detach()).norm() / yr.detach().norm()) < 1e-4
    assert float((m.weight.grad.cpu() - ref.weight.grad).norm() / ref.weight.grad.norm()) < 1e-3
    assert float((xd.grad.cpu() - xr.grad).norm() / xr.grad.norm()) < 1e-3
    for k in ("quantize_input.running_min", "quantize_input.running_max"):
        np.testing.assert_allclose(m.state_dict()[k].cpu().numpy(), ref.state_dict()[k].numpy(), rtol=1e-6, atol=1e-7)
    m2 = Q.QConv2d(64, 64, **kw)
    m2.load_state_dict(ref.state_dict())
    m2 = m2.to(DEV).train()
    Q.USE_DW_KERNEL = False
    try:
        y2 = m2(x.to(DEV))
    finally:
        Q.USE_DW_KERNEL = True
    assert float((y1 - y2).detach().norm() / y2.detach().norm()) < 1e-4


DW_BWD_CASES = [  # B, C, H, W, stride, pad
    (2, 32, 112, 112, 1, 1),
    (2, 64, 112, 112, 2, 1),
    (3, 128, 57, 55, 2, 1),       # odd sizes: last row / column of the input receives no stride-2 tap
    (4, 512, 14, 14, 1, 1),
    (5, 1024, 7, 7, 1, 1),
    (3, 1024, 7, 7, 2, 1),
    (1, 8, 3, 3, 1, 1),
    (2, 16, 10, 10, 1, 0),
    (40, 8, 6, 6, 2, 1),
]


@pytest.mark.parametrize("case", DW_BWD_CASES)
def test_conv_dw_backward(lib, case):
    """qb_conv_dw_dgrad / qb_conv_dw_wgrad against aten::convolution_backward (groups = C) in fp64 on the same tensors."""
    B, C, H, W, stride, pad = case
    gen = torch.Generator().manual_seed(sum(case) + 3)
    x = torch.randn(B, C, H, W, generator=gen)
    w = torch.randn(C, 1, 3, 3, generator=gen) * 0.2
    Ho, Wo = (H + 2 * pad - 3) // stride + 1, (W + 2 * pad - 3) // stride + 1
    g = torch.randn(B, C, Ho, Wo, generator=gen)
    xd = x.to(DEV).contiguous(memory_format=torch.channels_last)
    gd = g.to(DEV).contiguous(memory_format=torch.channels_last)
    wd = w.to(DEV)
    assert lib.conv_dw_bwd_supported(gd, xd, wd, stride, pad) or H * W == 1
    dx = lib.conv_dw_dgrad(gd, wd, x.shape, stride, pad)
    dw, db = lib.conv_dw_wgrad(gd, xd, stride, pad, True)
    dw2, db2 = lib.conv_dw_wgrad(gd, xd, stride, pad, False)
    torch.cuda.synchronize()
    assert db2 is None and torch.equal(dw, dw2)                           # deterministic
    rx, rw, rb = torch.ops.aten.convolution_backward(g.double(), x.double(), w.double(), [C], [stride, stride], [pad, pad],
                                                     [1, 1], False, [0, 0], C, [True, True, True])
    assert dx.shape == rx.shape and dx.is_contiguous(memory_format=torch.channels_last)
    for got, ref in ((dx, rx), (dw, rw), (db, rb)):
        rel = float((got.cpu().double() - ref).norm() / ref.norm())
        assert rel < 2e-6, rel


def test_qconv2d_depthwise_backward_uses_native_kernels(lib):
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import quantize as Q
    torch.manual_seed(12)
    kw = dict(kernel_size=3, stride=1, padding=1, groups=32, bias=True, num_bits=8, num_bits_weight=8, num_bits_grad=None)
    m = Q.QConv2d(32, 32, **kw).to(DEV).train()
    m2 = Q.QConv2d(32, 32, **kw).to(DEV).train()
    m2.load_state_dict(m.state_dict())
    x = torch.randn(4, 32, 20, 20, device=DEV).clamp_min(0)
    xa, xb = x.clone().requires_grad_(), x.clone().requires_grad_()
    y = m(xa)
    lib.profile_begin()
    y.square().mean().backward()
    names = [r[0] for r in lib.profile_end()]
    assert "qb_conv_dw_dgrad" in names and "qb_conv_dw_wgrad" in names, names
    Q.USE_DW_KERNEL = False
    try:
        m2(xb).square().mean().backward()
    finally:
        Q.USE_DW_KERNEL = True
    for a, b in ((xa.grad, xb.grad), (m.weight.grad, m2.weight.grad), (m.bias.grad, m2.bias.grad)):
        assert float((a - b).norm() / b.norm()) < 1e-4
