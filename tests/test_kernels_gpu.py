"""Parity of the CUDA kernels (through the C ABI) against the CPU oracle and the reference goldens.
Integer codes and dequantized values must be BIT-EXACT given the same range / noise."""
import numpy as np
import pytest
import torch

from conftest import golden
from oracle import qref

pytestmark = pytest.mark.gpu
DEV = "cuda"


def T(a):
    return torch.from_numpy(np.asarray(a))


def bits_equal(a, b, what=""):
    a = a.detach().cpu().contiguous().numpy()
    b = b.detach().cpu().contiguous().numpy() if torch.is_tensor(b) else np.asarray(b)
    assert a.shape == b.shape, (a.shape, b.shape)
    bad = a.view(np.int32) != b.view(np.int32)
    # +0.0 / -0.0 are distinct bit patterns but equal values; the reference keeps whichever ATen makes
    bad &= ~((a == 0) & (b == 0))
    assert not bad.any(), "%s: %d / %d elements differ, max abs diff %g" % (what, bad.sum(), a.size, np.abs(a - b).max())


# ------------------------------------------------------------------------------------------------
# A1 affine
# ------------------------------------------------------------------------------------------------
def test_affine_golden_cases(lib):
    g = golden("fq_cases.npz")
    for i in range(int(g["count"])):
        x = T(g["x%d" % i]).to(DEV)
        bits, lo, hi = g["meta%d" % i]
        bits = int(bits)
        # Python-float range (host doubles)
        y = lib.fakequant_affine(x, bits, lo=float(lo), hi=float(hi))
        bits_equal(y, g["given%d" % i], "given case %d" % i)
        # range from data, tensor-scale branch
        r = lib.minmax(x)
        y = lib.fakequant_affine(x, bits, range_dev=r, scale_mode=lib.QB_SCALE_FLOAT)
        bits_equal(y, g["data%d" % i], "data case %d" % i)
        bits_equal(y, g["scalar_t%d" % i], "0-dim tensor range case %d" % i)
        # per-row ranges
        y, rmin, rmax = lib.fakequant_rows(x, bits)
        bits_equal(y, g["rows%d" % i], "rows case %d" % i)
        flat = T(g["x%d" % i]).flatten(1)
        bits_equal(rmin, flat.min(-1)[0])
        bits_equal(rmax, flat.max(-1)[0])


def test_affine_kat(lib):
    g = golden("kat.npz")
    x = T(g["x"]).to(DEV)
    bits_equal(lib.fakequant_affine(x, 8, lo=-1.0, hi=1.0), g["y8"])
    bits_equal(lib.fakequant_affine(x, 4, lo=-1.0, hi=1.0), g["y4"])
    c = torch.full((1, 4), 0.5, device=DEV)
    bits_equal(lib.fakequant_affine(c, 8, range_dev=lib.minmax(c), scale_mode=lib.QB_SCALE_FLOAT), g["const"])
    w = T(g["w"]).to(DEV)
    bits_equal(lib.fakequant_rows(w, 8)[0], g["w_rows"])


@pytest.mark.parametrize("n", [1, 3, 4, 5, 1023, 4096, 65537, (1 << 20) + 7])
@pytest.mark.parametrize("bits", [4, 8])
def test_affine_vs_oracle_sizes(lib, n, bits):
    gen = torch.Generator().manual_seed(n * 31 + bits)
    x = torch.randn(n, generator=gen) * 3
    lo, hi = float(x.min()) * 0.8, float(x.max()) * 0.9
    ref, ref_codes = qref.fq_affine(x, bits, lo, hi, want_codes=True)
    xd = x.to(DEV)
    codes = torch.empty(n, dtype=torch.uint8, device=DEV)
    y = lib.fakequant_affine(xd, bits, lo=lo, hi=hi, codes=codes)
    bits_equal(y, ref, "values")
    assert torch.equal(codes.cpu().float(), ref_codes), "integer codes differ"
    # device range pointer == host doubles of the same fp32 values
    r = torch.tensor([lo, hi], dtype=torch.float32)
    y2 = lib.fakequant_affine(xd, bits, range_dev=r.to(DEV))
    bits_equal(y2, qref.fq_affine(x, bits, float(r[0]), float(r[1])), "device range")
    # codes only
    codes2 = torch.zeros_like(codes)
    lib.fakequant_affine(xd, bits, lo=lo, hi=hi, codes=codes2, want_y=False)
    assert torch.equal(codes, codes2)


def test_affine_unaligned_view_and_inplace(lib):
    x = torch.randn(10007) * 2
    xd = x.to(DEV)
    view = xd[1:]                       # 4-byte aligned only -> scalar path
    y = lib.fakequant_affine(view, 8, lo=-1.5, hi=2.5)
    bits_equal(y, qref.fq_affine(x[1:], 8, -1.5, 2.5))
    buf = xd.clone()
    lib.fakequant_affine(buf, 8, lo=-1.5, hi=2.5, out=buf)
    bits_equal(buf, qref.fq_affine(x, 8, -1.5, 2.5))


def test_affine_empty(lib):
    x = torch.empty(0, device=DEV)
    assert lib.fakequant_affine(x, 8, lo=0.0, hi=1.0).numel() == 0


@pytest.mark.parametrize("rows,cols", [(1, 1), (64, 27), (64, 147), (512, 4608), (1000, 512), (7, 4099),
                                       (5, 8192), (3, 20004), (4, 9001), (2, 65536 + 6)])    # >= 8192: cluster-split rows
def test_rows_vs_oracle(lib, rows, cols):
    gen = torch.Generator().manual_seed(rows * 7 + cols)
    w = torch.randn(rows, cols, generator=gen) * 0.1
    if rows > 2:
        w[1] = 0.25                                           # constant row -> scale clamp 1e-8
    lo, hi = qref.weight_row_range(w)
    ref, ref_codes = qref.fq_affine(w, 8, lo, hi, want_codes=True)
    codes = torch.empty(rows, cols, dtype=torch.uint8, device=DEV)
    y, rmin, rmax = lib.fakequant_rows(w.to(DEV), 8, codes=codes)
    bits_equal(y, ref)
    assert torch.equal(codes.cpu().float(), ref_codes)
    bits_equal(rmin, lo.flatten())
    bits_equal(rmax, hi.flatten())
    # frozen ranges
    y2, _, _ = lib.fakequant_rows(w.to(DEV), 8, row_min=rmin, row_max=rmax, compute_range=False)
    bits_equal(y2, ref)


# ------------------------------------------------------------------------------------------------
# A2 gradient quantizer
# ------------------------------------------------------------------------------------------------
def test_grad_golden(lib):
    g = golden("fq_grad.npz")
    for i in range(int(g["count"])):
        gr, noise = T(g["g%d" % i]).to(DEV), T(g["noise%d" % i]).to(DEV)
        bits = int(g["bits%d" % i])
        r = lib.minmax(gr)
        out = lib.fakequant_grad(gr, bits, range_dev=r, noise=noise)
        bits_equal(out, g["out%d" % i], "two-call case %d" % i)
        out2, r2 = lib.fakequant_grad_fused(gr, bits, noise=noise)
        bits_equal(out2, g["out%d" % i], "fused case %d" % i)
        bits_equal(r2, r)


@pytest.mark.parametrize("n", [1, 5, 4096, 100003, (1 << 21) + 3])
@pytest.mark.parametrize("shift", [0.0, 4.0, -4.0])
def test_grad_vs_oracle(lib, n, shift):
    gen = torch.Generator().manual_seed(n + int(shift * 10) + 100)
    gr = (torch.randn(n, generator=gen) + shift) * 1e-3
    noise = torch.rand(n, generator=gen) - 0.5
    lo, hi = float(gr.min()), float(gr.max())
    ref, ref_codes = qref.fq_zero_point(gr, 8, lo, hi, noise, want_codes=True)
    codes = torch.empty(n, dtype=torch.uint8, device=DEV)
    out, r = lib.fakequant_grad_fused(gr.to(DEV), 8, noise=noise.to(DEV), codes=codes)
    bits_equal(out, ref)
    assert torch.equal(codes.cpu().float(), ref_codes)
    assert r.cpu().tolist() == [lo, hi]
    # deterministic (stochastic=False)
    out = lib.fakequant_grad(gr.to(DEV), 8, lo=lo, hi=hi, stochastic=False)
    bits_equal(out, qref.fq_zero_point(gr, 8, lo, hi, None))


def test_grad_philox_statistics_and_replay(lib):
    """Production noise: Philox.  Same (seed, offset) => same draw; unbiased stochastic rounding;
    the noise actually used can be recovered and fed to the oracle for a bit-exact match."""
    n = 1 << 20
    gen = torch.Generator().manual_seed(5)
    gr = (torch.randn(n, generator=gen) * 1e-3).to(DEV)
    a, r = lib.fakequant_grad_fused(gr, 8, seed=1234)
    b, _ = lib.fakequant_grad_fused(gr, 8, seed=1234)
    c, _ = lib.fakequant_grad_fused(gr, 8, seed=1235)
    assert torch.equal(a, b) and not torch.equal(a, c)
    lo, hi = r.cpu().tolist()
    s = qref.affine_scale(lo, hi, 8)
    err = (a - gr).double()
    assert abs(float(err.mean())) < 3 * s / np.sqrt(12 * n) * 2 + 1e-12     # unbiased
    assert float(err.abs().max()) <= s * 1.0001                             # within one step
    # recover the noise: quantize a zero tensor shifted so that codes reveal u -- simpler: compare the
    # two-call path with the fused path under the same Philox stream
    d = lib.fakequant_grad(gr, 8, range_dev=r, seed=1234)
    assert torch.equal(a, d)
    # step counter on the device changes the stream (CUDA-graph replays)
    step = torch.tensor([3], dtype=torch.int64, device=DEV)
    e, _ = lib.fakequant_grad_fused(gr, 8, seed=1234, step_dev=step)
    assert not torch.equal(a, e)


def test_fast_division_is_ieee(lib):
    """The FMA division every quantizer uses (common.cuh qb_div) against __fdiv_rn, bit for bit: dense random
    numerators over the magnitudes gradients/activations take, exact zeros (post-ReLU tensors; they must not change
    the answer), values around code boundaries, and the fallback window (huge / inf)."""
    gen = torch.Generator().manual_seed(77)
    n = 1 << 22
    for b in [1e-8, 3.1e-7, 1.7e-4, 0.0123, 0.9999, 1.0, 3.0, 255.0, 1e5]:
        mag = torch.exp2(torch.randint(-40, 20, (n,), generator=gen).float())
        a = torch.randn(n, generator=gen) * mag
        a[::7] = 0.0
        a[1::1013] = -0.0
        k = torch.randint(0, 256, (n // 4,), generator=gen).float()
        a[: n // 4] = (k + 0.5) * b                                   # quotients next to rounding ties
        a[n // 4: n // 2] = torch.nextafter(a[: n // 4], torch.zeros(()))
        a[-4:] = torch.tensor([3e38, -3e38, float("inf"), 1e30])
        fast, ieee = lib.selftest_div(a.to(DEV), b)
        bits_equal(fast, ieee, "b=%g" % b)
        ref = a / torch.tensor(b, dtype=torch.float32)               # ATen CPU true division
        sel = a.abs() >= 2.0 ** -100                                 # CPU may flush nothing; compare where normal
        assert torch.equal(ieee.cpu()[sel].view(torch.int32), ref[sel].view(torch.int32))


def test_philox_noise_lattice_and_moments(lib):
    """Recover the in-kernel noise u through the quantizer itself: with g = 0, range [-128, 127] (scale 1, zp 128)
    the code is rne(128 + u) - which only reveals the sign - so use 16 bits: scale 1/257.. simpler and exact:
    quantize g = k + 0.5 (k integer) with scale 1: the code is k + [u >= 0 rounding], giving P(u > 0); and the
    mean/variance of the rounding error of a constant tensor give E[u], Var[u]."""
    n = 1 << 22
    # scale exactly 1: lo = -128, hi = 127 -> s = 255/255 = 1, zp = 128
    frac = 0.25
    g = torch.full((n,), 3.0 + frac, device=DEV)
    g[0], g[1] = -128.0, 127.0                                     # pin the range
    out, r = lib.fakequant_grad_fused(g, 8, seed=99)
    assert r.cpu().tolist() == [-128.0, 127.0]
    up = (out[2:] == 4.0).double().mean().item()                   # P(round up) must equal the fraction
    assert abs(up - frac) < 4 * np.sqrt(frac * (1 - frac) / n)
    assert bool(((out[2:] == 3.0) | (out[2:] == 4.0)).all())
    # different octets / lanes are uncorrelated: lag-1 and lag-8 autocorrelation of the up/down pattern
    z = (out[2:] == 4.0).double() - up
    for lag in (1, 2, 8, 256):
        c = float((z[:-lag] * z[lag:]).mean() / (z * z).mean())
        assert abs(c) < 5 / np.sqrt(n), (lag, c)


def test_minmax(lib):
    for n in [1, 2, 1000, 1 << 22]:
        x = torch.randn(n)
        r = lib.minmax(x.to(DEV))
        assert r.cpu().tolist() == [float(x.min()), float(x.max())]
        r = lib.minmax(x.to(DEV))       # workspace left clean -> second call identical
        assert r.cpu().tolist() == [float(x.min()), float(x.max())]


# ------------------------------------------------------------------------------------------------
# A5 QuantMeasure statistics
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape", [(4, 3, 5, 5), (16, 64, 8, 8), (256, 512), (3, 10001), (2, 40000), (8192, 64)])
def test_measure_stats(lib, shape):
    gen = torch.Generator().manual_seed(sum(shape))
    x = torch.randn(shape, generator=gen) * 2 + 0.3
    rows = x.reshape(shape[0], -1)
    for _ in range(2):                                    # second call checks the workspace was left clean
        st = lib.measure_stats(x.to(DEV)).cpu()
        assert st[0].item() == rows.min(-1)[0].double().sum().item()
        assert st[1].item() == rows.max(-1)[0].double().sum().item()
        assert abs(st[2].item() - x.double().sum().item()) <= 1e-7 * x.double().abs().sum().item()
        assert abs(st[3].item() - (x.double() ** 2).sum().item()) <= 1e-7 * (x.double() ** 2).sum().item()
        assert st[4].item() == float(x.min()) and st[5].item() == float(x.max())
        assert st[6].item() == x.numel() and st[7].item() == shape[0]


@pytest.mark.parametrize("method", ["avg", "aciq"])
@pytest.mark.parametrize("mom", [None, 0.9])
def test_measure_golden(lib, method, mom):
    """Buffers after three training steps and the outputs, against the unmodified reference."""
    g = golden("measure.npz")
    tag = "%s_%s" % (method, "cum" if mom is None else "ema")
    bufs = [torch.zeros(1, device=DEV), torch.zeros(1, device=DEV), torch.zeros(1, device=DEV),
            torch.ones(1, device=DEV), torch.zeros(1, device=DEV)]          # min, max, n, var, mean
    m = 1 if method == "aciq" else 0
    for step in range(3):
        x = T(g["%s_x%d" % (tag, step)]).to(DEV)
        st = lib.measure_stats(x)
        r = lib.measure_finalize(st, bufs, mom, True, m, 8)
        y = lib.fakequant_affine(x, 8, range_dev=r)
        got = torch.cat(bufs).cpu().numpy()
        want = g["%s_buf%d" % (tag, step)]
        # the batch-mean of per-sample extrema is rounded once from fp64 here, ATen sums in fp32: <= 1 ulp
        np.testing.assert_allclose(got, want, rtol=3e-7, atol=1e-7)
        if np.array_equal(got[:2], want[:2]) and method == "avg":
            bits_equal(y, g["%s_y%d" % (tag, step)], "measure output step %d" % step)
        else:
            np.testing.assert_allclose(y.cpu().numpy(), g["%s_y%d" % (tag, step)], rtol=0, atol=2e-6 * (1 + step))
    r = lib.measure_finalize(None, bufs, mom, False, m, 8)
    y = lib.fakequant_affine(T(g["%s_xe" % tag]).to(DEV), 8, range_dev=r)
    np.testing.assert_allclose(y.cpu().numpy(), g["%s_ye" % tag], rtol=0, atol=5e-6)


@pytest.mark.parametrize("world", [1, 2, 4])
def test_allreduce_small_protocol_virtual_ranks(lib, world):
    """qb_allreduce_small (the one-shot peer-memory all-reduce of the statistic blocks): `world` virtual ranks on ONE GPU,
    each with its own exchange buffer and stream, run a sequence of calls of every kind; every rank must end up with the
    rank-ordered reduction, bit-identical across ranks, and the call counters / error words must be clean."""
    import ctypes
    slot = 4096
    nbytes = lib.load().qb_allreduce_small_buffer_bytes(world, slot)
    bufs = [torch.zeros(nbytes, dtype=torch.uint8, device=DEV) for _ in range(world)]
    ptrs = (ctypes.c_uint64 * world)(*[b.data_ptr() for b in bufs])
    streams = [torch.cuda.Stream() for _ in range(world)]
    gen = torch.Generator().manual_seed(world)
    torch.cuda.synchronize()
    calls = [(torch.float64, 8, 1), (torch.float64, 3 * 64, 0), (torch.float32, 2, 2), (torch.float64, 2 * 64, 0),
             (torch.float64, 8, 1), (torch.float32, 2, 2), (torch.float64, 512, 0)]
    for dtype, n, mode in calls * 3:
        src = [torch.randn(n, generator=gen, dtype=torch.float64).to(dtype).to(DEV) for _ in range(world)]
        work = [s.clone() for s in src]
        torch.cuda.synchronize()
        for r in range(world):
            with torch.cuda.stream(streams[r]):
                lib.allreduce_small(work[r], mode, ptrs, r, world, slot)
        torch.cuda.synchronize()
        want = src[0].clone()
        for q in range(1, world):
            if mode == 0:
                want = want + src[q]
            elif mode == 1:
                w2 = want + src[q]
                w2[4] = torch.minimum(want[4], src[q][4])
                w2[5] = torch.maximum(want[5], src[q][5])
                want = w2
            else:
                want = torch.stack([torch.minimum(want[0], src[q][0]), torch.maximum(want[1], src[q][1])])
        for r in range(world):
            assert torch.equal(work[r], want), (dtype, n, mode, r)
    for b in bufs:
        assert int(b[0:8].view(torch.int64).item()) == len(calls) * 3
        assert int(b[8:12].view(torch.int32).item()) == 0
    with pytest.raises(lib.QB200Error):
        lib.allreduce_small(torch.zeros(slot, dtype=torch.float64, device=DEV), 0, ptrs, 0, world, slot)
