"""Config 5a of BASELINE.json: quantize() kernel sweep (1 MB - 4 GB), achieved GB/s per kernel.
Timed with CUDA events on the launching stream; inputs above 2^25 elements exceed L2 (126 MB)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa: E402,F401
from qtorch_b200 import _abi  # noqa: E402


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in evs:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2]


def main():
    peak = 6537.3
    try:
        peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    dev = torch.device("cuda")
    rows = []
    exps = [int(a) for a in sys.argv[1:]] or [18, 20, 22, 24, 26, 28, 30]       # 1 MB ... 4 GB of fp32
    for e in exps:
        n = 1 << e
        x = torch.randn(n, device=dev) * 3
        y = torch.empty_like(x)
        codes = torch.empty(n, dtype=torch.uint8, device=dev)
        r = _abi.minmax(x)
        xr = torch.relu(x)                      # half exact zeros with min == 0: what every post-ReLU QConv2d input is
        rr = _abi.minmax(xr)
        cases = {
            "affine_relu_in+codes(9B)": (9, lambda: _abi.fakequant_affine(xr, 8, range_dev=rr, out=y, codes=codes)),
            "grad_fused_relu_masked(12B)": (12, lambda: _abi.fakequant_grad_fused(xr, 8, out=y, seed=1)),
            "affine_given(8B/elem)": (8, lambda: _abi.fakequant_affine(x, 8, range_dev=r, out=y)),
            "affine_given+codes(9B)": (9, lambda: _abi.fakequant_affine(x, 8, range_dev=r, out=y, codes=codes)),
            "affine_4bit(8B)": (8, lambda: _abi.fakequant_affine(x, 4, range_dev=r, out=y)),
            "minmax(4B)": (4, lambda: _abi.minmax(x, range_out=r)),
            "measure_stats(4B)": (4, lambda: _abi.measure_stats(x.view(256, -1))),
            "grad_given_philox(8B)": (8, lambda: _abi.fakequant_grad(x, 8, range_dev=r, out=y, seed=1)),
            "grad_fused_philox(12B)": (12, lambda: _abi.fakequant_grad_fused(x, 8, out=y, seed=1)),
            "grad_fused_nonoise(12B)": (12, lambda: _abi.fakequant_grad_fused(x, 8, out=y, stochastic=False)),
            "torch_copy(8B)": (8, lambda: y.copy_(x)),
        }
        for name, (bpe, fn) in cases.items():
            ms = timeit(fn)
            gbs = bpe * n / ms / 1e6
            rows.append(dict(kernel=name, log2n=e, ms=round(ms, 4), gbs=round(gbs, 1), frac=round(gbs / peak, 3)))
            print("%-28s n=2^%d  %8.4f ms  %8.1f GB/s  %.3f of measured %.0f  %.3f of the 8000 GB/s datasheet"
                  % (name, e, ms, gbs, gbs / peak, peak, gbs / 8000.0), flush=True)
        del x, y, codes, xr
    # per-channel ranges (quantize.py:318-334): C rows of `cols` elements, range from the data + quantize (8 B/elem)
    for rows_, cols in [(64, 1 << 16), (64, 1 << 20), (512, 1 << 17), (1000, 1 << 16), (512, 4608), (1000, 512)]:
        for bits in (8, 4):
            w = torch.randn(rows_, cols, device=dev)
            out = torch.empty_like(w)
            ms = timeit(lambda: _abi.fakequant_rows(w, bits, out=out))
            gbs = 8 * w.numel() / ms / 1e6
            rows.append(dict(kernel="rows_%dbit" % bits, rows=rows_, cols=cols, ms=round(ms, 4), gbs=round(gbs, 1), frac=round(gbs / peak, 3)))
            print("per-channel %d-bit  %5d x %-8d %8.4f ms  %8.1f GB/s  %.3f of measured  %.3f of datasheet"
                  % (bits, rows_, cols, ms, gbs, gbs / peak, gbs / 8000.0), flush=True)
    out = os.path.join(os.path.dirname(__file__), "..", "gpurun_out")
    os.makedirs(out, exist_ok=True)
    json.dump(rows, open(os.path.join(out, "kernel_sweep.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
