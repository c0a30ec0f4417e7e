"""qb_conv_stem_fwd on the ResNet-18 stem (256x3x224x224, 7x7/2 -> 64) and the MobileNet stem (3x3/2 -> 32): median time
with L2 flushed, against cuDNN's F.conv2d on the same fake-quantized channels-last tensors and the output-write floor."""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa: E402,F401
from qtorch_b200 import _abi  # noqa: E402
from conv_sweep import timeit, HBM_GBS  # noqa: E402


def main():
    B = int(os.environ.get("B", 256))
    dev = torch.device("cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for name, Cout, k, stride, pad in (("resnet stem 7x7/2 3->64", 64, 7, 2, 3), ("mobilenet stem 3x3/2 3->32", 32, 3, 2, 1)):
        x = (torch.randn(B, 3, 224, 224, device=dev) * 1.1).contiguous(memory_format=torch.channels_last)
        w = torch.randn(Cout, 3, k, k, device=dev) * 0.08
        rng = _abi.minmax(x)
        codes = torch.empty_like(x, dtype=torch.uint8)
        xq = _abi.fakequant_affine(x, 8, range_dev=rng, codes=codes)
        rmin, rmax = torch.empty(Cout, device=dev), torch.empty(Cout, device=dev)
        wq, wcodes, tapsum = _abi.fakequant_rows_conv(w, 8, rmin, rmax)
        wq_cl = wq.contiguous(memory_format=torch.channels_last)
        t = timeit(lambda: _abi.conv_stem_fwd(codes, wcodes, B, 224, 224, 3, Cout, k, k, stride, pad, rng, 8, rmin, rmax, 8, tapsum),
                   flush)
        t_ref = timeit(lambda: F.conv2d(xq, wq_cl, None, stride, pad), flush)
        Ho = (224 + 2 * pad - k) // stride + 1
        bytes_ = B * 224 * 224 * 3 + 4 * B * Ho * Ho * Cout
        ops = 2 * B * Ho * Ho * Cout * k * k * 3
        print("%-28s stem %.3f ms (%.1f TOP/s, %.0f GB/s of %.0f)   cuDNN fp32 %.3f ms   floor %.3f ms" %
              (name, t, ops / t / 1e9, bytes_ / t / 1e6, HBM_GBS, t_ref, bytes_ / HBM_GBS / 1e6))


if __name__ == "__main__":
    main()
