"""Does tcgen05.mma accept shared-memory descriptors that start at a row inside an 8-row swizzle atom (128-byte swizzle)?
Prints the relative error of D for start rows 0..9, K-major and MN-major, with and without the base-offset field."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa
from qtorch_b200 import _abi
lib = _abi.load()
torch.manual_seed(0)
for mn in (0, 1):
    if not mn:
        a = torch.randint(-8, 9, (144, 64), device="cuda").float(); b = torch.randint(-8, 9, (64, 64), device="cuda").float()
    else:
        a = torch.randint(-8, 9, (80, 128), device="cuda").float(); b = torch.randint(-8, 9, (80, 64), device="cuda").float()
    a16, b16 = a.to(torch.bfloat16).contiguous(), b.to(torch.bfloat16).contiguous()
    for boff in (0, 1):
        res = []
        for r0 in range(0, 11):
            out = torch.zeros(128, 64, device="cuda")
            rc = lib.qb_debug_umma_shift(a16.data_ptr(), b16.data_ptr(), out.data_ptr(), mn, r0, boff, torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
            ref = a[r0:r0 + 128] @ b.t() if not mn else a[r0:r0 + 64].t() @ b[r0:r0 + 64]
            res.append("%d:%s" % (r0, "ok" if torch.equal(out, ref) else "%.2g" % float((out - ref).norm() / ref.norm())))
        print("mn-major" if mn else "k-major ", "base_offset" if boff else "plain      ", " ".join(res), flush=True)

# MN-major operand whose two 64-row column blocks are the SAME strip read `lbo` rows apart (descriptor LBO = lbo * 128 B):
# D[m < 64] = SUM_k A[r0 + k][m] B[r0 + k], D[64 + m] = SUM_k A[r0 + lbo + k][m] B[r0 + k]  (two filter taps in one MMA)
a = torch.randint(-8, 9, (80, 128), device="cuda").float(); b = torch.randint(-8, 9, (80, 64), device="cuda").float()
a16, b16 = a.to(torch.bfloat16).contiguous(), b.to(torch.bfloat16).contiguous()
for lbo in (1, 2, 3, 8, 9):
    res = []
    for r0 in range(0, 6):
        out = torch.zeros(128, 64, device="cuda")
        lib.qb_debug_umma_shift(a16.data_ptr(), b16.data_ptr(), out.data_ptr(), 2, r0, lbo, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        ref = torch.cat([a[r0:r0 + 64, :64].t() @ b[r0:r0 + 64], a[r0 + lbo:r0 + lbo + 64, :64].t() @ b[r0:r0 + 64]])
        res.append("%d:%s" % (r0, "ok" if torch.equal(out, ref) else "%.2g" % float((out - ref).norm() / ref.norm())))
    print("mn-major overlapping blocks, LBO = %d rows" % lbo, " ".join(res), flush=True)
