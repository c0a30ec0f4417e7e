"""Config 5b of BASELINE.json: QLinear int8 GEMM, M=N=K in 1024..8192.  Pure-GEMM TOP/s (2*M*N*K ops) against the
int8 dense datasheet peak (4.5 POP/s) and, for context, torch's bf16 / tf32 matmul on the same shapes."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa: E402,F401
from qtorch_b200 import _abi  # noqa: E402
from kernel_sweep import timeit  # noqa: E402


def main():
    dev = torch.device("cuda")
    sizes = [int(a) for a in sys.argv[1:]] or [1024, 2048, 4096, 8192]
    for n in sizes:
        M = N = K = n
        a = torch.randint(0, 256, (M, K), dtype=torch.uint8, device=dev)
        b = torch.randint(0, 256, (N, K), dtype=torch.uint8, device=dev)
        rng = torch.tensor([0.0, 4.0], device=dev)
        rmin, rmax = -torch.rand(N, device=dev), torch.rand(N, device=dev)
        ars, brs = _abi.rowsum_u8(a, M, K), _abi.rowsum_u8(b, N, K)
        out = torch.empty(M, N, device=dev)
        ms = timeit(lambda: _abi.gemm_i8(a, b, M, N, K, rng, 8, rmin, rmax, 8, ars, brs, out=out), iters=20)
        tops = 2.0 * M * N * K / ms / 1e9
        x16, w16 = torch.randn(M, K, device=dev, dtype=torch.bfloat16), torch.randn(N, K, device=dev, dtype=torch.bfloat16)
        ms16 = timeit(lambda: torch.matmul(x16, w16.t()), iters=20)
        x32, w32 = x16.float(), w16.float()
        torch.backends.cuda.matmul.allow_tf32 = True
        ms32 = timeit(lambda: torch.matmul(x32, w32.t()), iters=20)
        print("M=N=K=%5d  qb_gemm_i8 %.4f ms  %7.1f TOP/s (%.3f of 4500 int8 dense)  | cuBLAS bf16 %.4f ms %7.1f TF/s | tf32 %.4f ms %6.1f TF/s"
              % (n, ms, tops, tops / 4500, ms16, 2.0 * M * N * K / ms16 / 1e9, ms32, 2.0 * M * N * K / ms32 / 1e9), flush=True)


if __name__ == "__main__":
    main()
