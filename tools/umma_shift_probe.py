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
