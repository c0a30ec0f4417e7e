"""int8 tcgen05 GEMM (qb_gemm_i8) against the fp64 product of the dequantised operands, and against the reference's
F.linear on the same fake-quantized tensors (north_star: <= 1e-4 relative L2)."""
import numpy as np
import pytest
import torch

from oracle import qref

pytestmark = pytest.mark.gpu
DEV = "cuda"


def make_case(lib, M, N, K, seed, relu=True):
    gen = torch.Generator().manual_seed(seed)
    x = torch.randn(M, K, generator=gen)
    if relu:
        x = x.clamp_min(0)
    w = torch.randn(N, K, generator=gen) * 0.02
    b = torch.randn(N, generator=gen) * 0.1
    xd, wd = x.to(DEV), w.to(DEV)
    rng = torch.tensor([float(x.min()), float(x.max())], device=DEV)
    a_codes = torch.empty(M, K, dtype=torch.uint8, device=DEV)
    xq = lib.fakequant_affine(xd, 8, range_dev=rng, codes=a_codes)
    b_codes = torch.empty(N, K, dtype=torch.uint8, device=DEV)
    wq, rmin, rmax = lib.fakequant_rows(wd, 8, codes=b_codes)
    return dict(x=x, w=w, b=b, rng=rng, a_codes=a_codes, b_codes=b_codes, xq=xq, wq=wq, rmin=rmin, rmax=rmax)


@pytest.mark.parametrize("M,N,K", [(128, 128, 128), (128, 128, 1024), (256, 1000, 512), (100, 72, 80), (8, 10, 32),
                                   (1000, 300, 2000), (2048, 1024, 1024), (129, 257, 144),
                                   (2048, 2048, 2048), (1100, 1300, 1040), (4096, 512, 256), (777, 2052, 4112)])
@pytest.mark.parametrize("relu", [True, False])
def test_gemm_i8(lib, M, N, K, relu):
    c = make_case(lib, M, N, K, seed=M + N + K, relu=relu)
    ars = lib.rowsum_u8(c["a_codes"], M, K)
    brs = lib.rowsum_u8(c["b_codes"], N, K)
    assert torch.equal(ars.cpu().long(), c["a_codes"].cpu().long().sum(1))
    assert torch.equal(brs.cpu().long(), c["b_codes"].cpu().long().sum(1))
    bias = c["b"].to(DEV)
    out = lib.gemm_i8(c["a_codes"], c["b_codes"], M, N, K, c["rng"], 8, c["rmin"], c["rmax"], 8, ars, brs, bias=bias)
    torch.cuda.synchronize()
    ref64 = c["xq"].double().cpu() @ c["wq"].double().cpu().t() + c["b"].double()
    rel = float((out.cpu().double() - ref64).norm() / ref64.norm())
    assert rel < 5e-6, rel          # fp32 epilogue on centred integer sums (pair kernel) or fp64 epilogue (single)
    # the reference computes this product in fp32 on the CPU (quantize.py:424)
    lo, hi = qref.weight_row_range(c["w"])
    ref32 = torch.nn.functional.linear(qref.fq_affine(c["x"], 8, float(c["x"].min()), float(c["x"].max())),
                                       qref.fq_affine(c["w"], 8, lo, hi), c["b"])
    rel32 = float((out.cpu() - ref32).norm() / ref32.norm())
    assert rel32 < 1e-4, rel32
    # centred signed weight codes (what the modules use): same result, better conditioned
    rmin2, rmax2 = torch.empty(N, device=DEV), torch.empty(N, device=DEV)
    wq2, wc, tapsum = lib.fakequant_rows_conv(c["w"].to(DEV).view(N, K, 1, 1), 8, rmin2, rmax2)
    assert torch.equal(wq2.view(N, K), c["wq"])
    assert torch.equal(wc.view(N, K).view(torch.int8).cpu().long() + 128, c["b_codes"].cpu().long())
    out2 = lib.gemm_i8(c["a_codes"], wc, M, N, K, c["rng"], 8, rmin2, rmax2, 8, ars, tapsum.view(N), bias=bias, b_offset=128)
    rel2 = float((out2.cpu().double() - ref64).norm() / ref64.norm())
    assert rel2 < 5e-6, rel2


def test_gemm_i8_rejects_unaligned_k(lib):
    c = make_case(lib, 16, 16, 24, seed=1)
    ars, brs = lib.rowsum_u8(c["a_codes"], 16, 24), lib.rowsum_u8(c["b_codes"], 16, 24)
    with pytest.raises(lib.QB200Error):
        lib.gemm_i8(c["a_codes"], c["b_codes"], 16, 16, 24, c["rng"], 8, c["rmin"], c["rmax"], 8, ars, brs)
