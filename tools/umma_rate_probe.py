"""What does the tensor pipe sustain for the MMA shapes / operand layouts of conv_bwd.cu?  One CTA per SM issues n
back-to-back tcgen05.mma kind::f16 (M 128, K 16) with no memory traffic; prints cycles per MMA and the implied TFLOP/s
against the 2.25 PFLOP/s datasheet figure."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa: E402,F401
from qtorch_b200 import _abi  # noqa: E402

lib = _abi.load()
sms = torch.cuda.get_device_properties(0).multi_processor_count
mhz = 1965.0
n = 4096
names = {0: "A K-major, B K-major", 1: "A MN, B K", 2: "A K, B MN", 3: "A MN, B MN"}
for blocks in (1, sms):
    for N in (64, 128, 256):
        for mode in (0, 3, 4, 7, 15):
            cyc = torch.zeros(blocks, dtype=torch.int64, device="cuda")
            for _ in range(2):
                rc = lib.qb_debug_umma_rate(cyc.data_ptr(), blocks, n, N, mode, torch.cuda.current_stream().cuda_stream)
                assert rc == 0, _abi.last_error()
                torch.cuda.synchronize()
            c = float(cyc.max()) / n
            tf = 2.0 * 128 * N * 16 / c * mhz * 1e6 * sms / 1e12
            print("CTAs %3d  N %3d  %-22s %s %s  %6.1f cycles/MMA  -> %6.0f TFLOP/s machine-wide (%.2f of 2250)"
                  % (blocks, N, names[mode & 3], "walk" if mode & 4 else "same", "shift3" if mode & 8 else "      ", c, tf, tf / 2250.0),
                  flush=True)
