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
        Ho = (224 + 2 * pad - k) // stride + 1
        t = timeit(lambda: _abi.conv_stem_fwd(codes, wcodes, B, 224, 224, 3, Cout, k, k, stride, pad, rng, 8, rmin, rmax, 8, tapsum),
                   flush)
        t_ref = timeit(lambda: F.conv2d(xq, wq_cl, None, stride, pad), flush)
        bytes_ = B * 224 * 224 * 3 + 4 * B * Ho * Ho * Cout
        ops = 2 * B * Ho * Ho * Cout * k * k * 3
        gout = torch.randn(B, Cout, Ho, Ho, device=dev).contiguous(memory_format=torch.channels_last)
        t_wg = timeit(lambda: _abi.conv_stem_wgrad(gout, codes, rng, 8, Cout, k, k, stride, pad), flush)
        w32 = torch.zeros_like(wq_cl)

        def cudnn_wgrad(gg=gout, xx=xq, ww=w32):
            return torch.ops.aten.convolution_backward(gg, xx, ww, None, [stride, stride], [pad, pad], [1, 1], False, [0, 0], 1,
                                                       [False, True, False])[1]
        t_wg_ref = timeit(cudnn_wgrad, flush)
        dw = _abi.conv_stem_wgrad(gout, codes, rng, 8, Cout, k, k, stride, pad)
        dw_ref = cudnn_wgrad()
        g16 = gout[:16].contiguous(memory_format=torch.channels_last)
        c16 = codes[:16].contiguous(memory_format=torch.channels_last)
        ref64 = cudnn_wgrad(g16.double(), xq[:16].double(), w32.double())
        dw16 = _abi.conv_stem_wgrad(g16, c16, rng, 8, Cout, k, k, stride, pad)
        dw16_ref = cudnn_wgrad(g16, xq[:16], w32)
        print("%-28s wgrad %.3f ms   cuDNN %.3f ms   rel.diff to cuDNN %.2e;  batch 16 vs fp64: ours %.2e, cuDNN %.2e" %
              (name, t_wg, t_wg_ref, float((dw - dw_ref).norm() / dw_ref.norm()),
               float((dw16.double() - ref64).norm() / ref64.norm()), float((dw16_ref.double() - ref64).norm() / ref64.norm())))
        print("%-28s stem %.3f ms (%.1f TOP/s, %.0f GB/s of %.0f)   cuDNN fp32 %.3f ms   floor %.3f ms" %
              (name, t, ops / t / 1e9, bytes_ / t / 1e6, HBM_GBS, t_ref, bytes_ / HBM_GBS / 1e6))


if __name__ == "__main__":
    main()
