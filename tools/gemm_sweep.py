"""BASELINE.json configs[4] part b / SURVEY.md section 8(d) config 5b: QLinear int8 GEMM sweep, M=N=K in {1024..8192}.

    python tools/gemm_sweep.py [--sizes 1024 2048 4096 8192] [--iters 20]

x ~ N(0,1) then ReLU, W ~ N(0, 0.02).  Two numbers per size:
  pure      qb_gemm_i8 alone on ready-made codes                         -> TOP/s (2*M*N*K) vs the int8 dense peak
  e2e       activation statistics + fake-quant to codes + row sums + weight row quantizer + GEMM (what QLinear.forward
            launches per call)
Correctness of every size is checked against an fp64 matmul of the dequantised operands (relative L2)."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa: E402,F401
from qtorch_b200 import _abi  # noqa: E402

INT8_PEAK_TOPS = 4500.0          # B200 datasheet, dense


def timed(fn, iters, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(iters):
        flush.zero_()                                  # > L2: operands come from HBM every time
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms.sort()
    return ms[len(ms) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", type=int, nargs="+", default=[1024, 2048, 4096, 8192])
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    _abi.device_info(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    g = torch.Generator(device="cpu").manual_seed(0)
    for n in args.sizes:
        M = N = K = n
        x = torch.randn(M, K, generator=g).relu_().to(dev)
        w = (torch.randn(N, K, generator=g) * 0.02).to(dev)
        bias = (torch.randn(N, generator=g) * 0.01).to(dev)
        a_range = torch.stack([x.min(), x.max()]).float().contiguous()
        xc = torch.empty(M, K, dtype=torch.uint8, device=dev)
        wc = torch.empty(N, K, dtype=torch.uint8, device=dev)
        xq = _abi.fakequant_affine(x, 8, a_range, codes=xc)
        wq, rmin, rmax = _abi.fakequant_rows(w, 8, codes=wc)
        ars = _abi.rowsum_u8(xc, M, K)
        brs = _abi.rowsum_u8(wc, N, K)
        out = torch.empty(M, N, dtype=torch.float32, device=dev)

        def pure():
            _abi.gemm_i8(xc, wc, M, N, K, a_range, 8, rmin, rmax, 8, ars, brs, bias, out=out)

        def e2e():
            _abi.minmax(x, a_range)
            _abi.fakequant_affine(x, 8, a_range, codes=xc, want_y=False)
            _abi.rowsum_u8(xc, M, K, out=ars)
            _abi.fakequant_rows(w, 8, rmin, rmax, codes=wc, want_y=False)
            _abi.rowsum_u8(wc, N, K, out=brs)
            _abi.gemm_i8(xc, wc, M, N, K, a_range, 8, rmin, rmax, 8, ars, brs, bias, out=out)

        pure()
        torch.cuda.synchronize()
        rows = slice(0, min(M, 512))
        ref = xq[rows].double() @ wq.double().t() + bias.double()
        rel = float((out[rows].double() - ref).norm() / ref.norm())
        t_pure = timed(pure, args.iters, flush)
        t_e2e = timed(e2e, args.iters, flush)
        tops = 2.0 * M * N * K / t_pure / 1e9
        tops_e = 2.0 * M * N * K / t_e2e / 1e9
        print("QLinear int8 M=N=K=%-5d pure %8.4f ms %8.1f TOP/s (%.3f of int8 peak %d)   e2e(+quantizers) %8.4f ms %8.1f TOP/s"
              "   rel-L2 vs fp64 %.2e" % (n, t_pure, tops, tops / INT8_PEAK_TOPS, int(INT8_PEAK_TOPS), t_e2e, tops_e, rel), flush=True)
        del x, w, xq, wq, xc, wc, out


if __name__ == "__main__":
    main()
