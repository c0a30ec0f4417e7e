"""qb_conv_i8_fwd on every int8-eligible ResNet-18 convolution shape at batch 256 (plus the QLinear GEMM sweep of
BASELINE config 5b with --gemm): time per layer, int8 TOP/s against the dense int8 peak, and the HBM floor
(codes in + fp32 out) next to it - most ResNet-18 layers are output-write bound, not tensor bound.
Timed with CUDA events on the launching stream, L2 flushed between iterations."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa: E402,F401
from qtorch_b200 import _abi  # noqa: E402

LAYERS = [  # name, Cin, H, Cout, k, stride, pad, count in ResNet-18
    ("layer1 3x3 64->64 @56", 64, 56, 64, 3, 1, 1, 4),
    ("layer2.0 3x3 64->128 s2", 64, 56, 128, 3, 2, 1, 1),
    ("layer2.0 1x1 64->128 s2", 64, 56, 128, 1, 2, 0, 1),
    ("layer2 3x3 128->128 @28", 128, 28, 128, 3, 1, 1, 3),
    ("layer3.0 3x3 128->256 s2", 128, 28, 256, 3, 2, 1, 1),
    ("layer3.0 1x1 128->256 s2", 128, 28, 256, 1, 2, 0, 1),
    ("layer3 3x3 256->256 @14", 256, 14, 256, 3, 1, 1, 3),
    ("layer4.0 3x3 256->512 s2", 256, 14, 512, 3, 2, 1, 1),
    ("layer4.0 1x1 256->512 s2", 256, 14, 512, 1, 2, 0, 1),
    ("layer4 3x3 512->512 @7", 512, 7, 512, 3, 1, 1, 3),
]
INT8_PEAK_TOPS = 4500.0      # dense int8, datasheet (no measured int8 figure in MEASURED_PEAKS.json)
HBM_GBS = 6551.0


def timeit(fn, flush, iters=8, warm=2):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    B = int(os.environ.get("B", 256))
    dev = torch.device("cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    tot = 0.0
    tot_floor = 0.0
    sel = [int(a) for a in sys.argv[1:]]
    for li, (name, Cin, H, Cout, k, stride, pad, count) in enumerate(LAYERS):
        if sel and li not in sel:
            continue
        x = torch.randn(B, Cin, H, H, device=dev).clamp_min_(0).contiguous(memory_format=torch.channels_last)
        w = torch.randn(Cout, Cin, k, k, device=dev) * 0.05
        rng = _abi.minmax(x)
        codes = torch.empty_like(x, dtype=torch.uint8)
        _abi.fakequant_affine(x, 8, range_dev=rng, codes=codes, want_y=False)
        pixsum = _abi.rowsum_u8(codes, B * H * H, Cin)
        rmin, rmax = torch.empty(Cout, device=dev), torch.empty(Cout, device=dev)
        _, wcodes, tapsum = _abi.fakequant_rows_conv(w, 8, rmin, rmax)
        Ho = (H + 2 * pad - k) // stride + 1
        fn = lambda: _abi.conv_i8_fwd(codes, wcodes, B, H, H, Cin, Cout, k, k, stride, pad, rng, 8, rmin, rmax, 8, pixsum,
                                      tapsum)
        ms = timeit(fn, flush)
        ops = 2.0 * B * Ho * Ho * Cout * Cin * k * k
        byts = B * H * H * Cin + 4.0 * B * Ho * Ho * Cout + Cout * Cin * k * k
        floor = max(ops / (INT8_PEAK_TOPS * 1e9), byts / (HBM_GBS * 1e6))
        print("%-28s x%d  %7.3f ms  %7.1f TOP/s (%.3f of int8 peak)  %7.1f GB/s  floor %.3f ms  (%.2fx floor)"
              % (name, count, ms, ops / ms / 1e9, ops / ms / 1e9 / INT8_PEAK_TOPS, byts / ms / 1e6, floor, ms / floor),
              flush=True)
        tot += ms * count
        tot_floor += floor * count
    print("ResNet-18 b%d int8 forward convs: %.3f ms per step, floor %.3f ms" % (B, tot, tot_floor))


if __name__ == "__main__":
    main()
