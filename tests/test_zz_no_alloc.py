"""qb_l2_sumsq_noalloc (include/qb200.h): the regulariser sum with its scratch in a static device array, for callers that
cannot allocate.  No counterpart in the reference.  Checked through the C-ABI
against the existing entry point (qb_l2_sumsq, caller workspace): bit-identical (both run the same fixed-order fp64
reduction), on two calls in a row that share the static array, the second over FEWER chunks than the first so stale partial
sums sit right behind the ones it reads.  Collected last (file name) so a failure here cannot hide other tests under -x."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _lists(tensors):
    n = len(tensors)
    ptrs = (C.c_void_p * n)(*[t.data_ptr() for t in tensors])
    sizes = (C.c_int64 * n)(*[t.numel() for t in tensors])
    return ptrs, sizes, n


def _with_workspace(lib, tensors, wd, out, accumulate):
    ptrs, sizes, n = _lists(tensors)
    ws = torch.empty(lib.qb_mt_ws_bytes(n, sizes), dtype=torch.uint8, device=DEV)
    rc = lib.qb_l2_sumsq(ptrs, sizes, n, wd, out.data_ptr(), accumulate, ws.data_ptr(), ws.numel(), None)
    assert rc == 0
    torch.cuda.synchronize()


def test_l2_sumsq_noalloc_matches_workspace_entry_point():
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import _abi
    lib = _abi.load()
    assert lib.qb_l2_sumsq_noalloc_max_blocks() == 8192
    torch.cuda.set_device(0)
    assert lib.qb_l2_sumsq_noalloc_setup() == 0
    gen = torch.Generator().manual_seed(2024)
    shapes_a = [(64, 64, 3, 3), (1000, 512), (1000,), (7,), (8192,), (8193,), (128, 64, 3, 3), (3, 5, 7)]
    shapes_b = [(10, 64), (10,), (16, 3, 3, 3)]                    # fewer chunks: the array still holds call A's partial sums
    first = [torch.randn(s, generator=gen).to(DEV) for s in shapes_a]
    second = [(torch.randn(s, generator=gen) * 3).to(DEV) for s in shapes_b]
    wd = 1e-4
    want = torch.zeros(4, device=DEV)
    got = torch.full((4,), float("nan"), device=DEV)
    # outputs and inputs exist before the calls; the no-alloc calls themselves must not allocate device memory
    _with_workspace(lib, first, wd, want[0:1], 0)
    _with_workspace(lib, second, wd, want[1:2], 0)
    want[2] = want[0]
    _with_workspace(lib, second, wd, want[2:3], 1)                 # accumulate form
    torch.cuda.synchronize()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    before = torch.cuda.memory_allocated()
    for tensors, slot, acc in ((first, 0, 0), (second, 1, 0)):      # two calls in a row on the shared scratch
        ptrs, sizes, n = _lists(tensors)
        rc = lib.qb_l2_sumsq_noalloc(ptrs, sizes, n, wd, got[slot:slot + 1].data_ptr(), acc, stream)
        assert rc == 0, _abi.last_error()
    got[2:3].copy_(got[0:1])
    ptrs, sizes, n = _lists(second)
    assert lib.qb_l2_sumsq_noalloc(ptrs, sizes, n, wd, got[2:3].data_ptr(), 1, stream) == 0
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    assert torch.equal(got[:3], want[:3]), (got, want)
    # against fp64 on the host as well (the tolerance of tests/test_glue_gpu.py)
    ref = 0.5 * wd * sum(float(t.double().pow(2).sum()) for t in first)
    assert abs(float(got[0]) - ref) <= 1e-6 * abs(ref)
    # the python wrapper drives the same entry point
    out = torch.empty((), device=DEV)
    _abi.l2_sumsq_noalloc(first, wd, out)
    assert float(out) == float(want[0])


def test_l2_sumsq_noalloc_rejects_sizes_beyond_the_static_scratch():
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import _abi
    lib = _abi.load()
    max_blocks = int(lib.qb_l2_sumsq_noalloc_max_blocks())
    chunk = 8192
    big = torch.zeros(max_blocks * chunk + 1, device=DEV)          # 256 MiB + 4 B: one chunk more than the array holds
    out = torch.zeros(1, device=DEV)
    ptrs, sizes, n = _lists([big])
    rc = lib.qb_l2_sumsq_noalloc(ptrs, sizes, n, 1.0, out.data_ptr(), 0, None)
    assert rc == -1                                                 # QB_ERR_ARG
    assert "static scratch" in _abi.last_error()
    torch.cuda.synchronize()
    assert float(out) == 0.0                                        # nothing was launched
    ptrs, sizes, n = _lists([big[:max_blocks * chunk]])             # exactly at the limit: accepted
    big[:16] = 2.0
    assert lib.qb_l2_sumsq_noalloc(ptrs, sizes, n, 1.0, out.data_ptr(), 0, None) == 0
    torch.cuda.synchronize()
    assert float(out) == 32.0
