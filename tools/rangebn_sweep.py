"""RangeBN kernel family on the ResNet-18 b256 activation shapes, NCHW vs channels-last; achieved GB/s
(algorithmic bytes per SURVEY 8d: stats 4 B/elem, apply 8, bwd reduce 8, bwd apply 8)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa: E402,F401
from qtorch_b200 import _abi  # noqa: E402
from kernel_sweep import timeit  # noqa: E402


def main():
    dev = torch.device("cuda")
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    for C, H in [(64, 112), (64, 56), (128, 28), (256, 14), (512, 7)]:
        for nhwc in (False, True):
            x = torch.randn(B, C, H, H, device=dev)
            g = torch.randn(B, C, H, H, device=dev) * 1e-3
            if nhwc:
                x = x.contiguous(memory_format=torch.channels_last)
                g = g.contiguous(memory_format=torch.channels_last)
            HW, n = H * H, x.numel()
            rm, rv = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
            gamma, beta = torch.rand(C, device=dev) + 0.5, torch.randn(C, device=dev)
            red, amax, amin = _abi.rangebn_stats(x, B, C, HW, 16, nhwc=nhwc)
            vec = _abi.rangebn_finalize(red, C, 16, B * HW, B * HW / 16, 1e-5, 0.1, True, 8, gamma, beta, rm, rv)
            out = torch.empty_like(x)
            t_stats = timeit(lambda: _abi.rangebn_stats(x, B, C, HW, 16, nhwc=nhwc))
            t_apply = timeit(lambda: _abi.rangebn_apply(x, B, C, HW, vec, True, True, out=out, nhwc=nhwc))
            t_bwd = timeit(lambda: _abi.rangebn_backward(g, x, B, C, HW, 16, B * HW, B * HW / 16, True, vec, True, amax, amin,
                                                         nhwc=nhwc))
            t_meas = timeit(lambda: _abi.measure_stats(x))
            print("C=%4d H=%3d %s  stats %.3f ms %6.0f GB/s | apply %.3f ms %6.0f GB/s | bwd(reduce+coef+apply,16B) %.3f ms %6.0f GB/s"
                  " | measure %.3f ms %6.0f GB/s" % (C, H, "NHWC" if nhwc else "NCHW", t_stats, 4 * n / t_stats / 1e6, t_apply,
                                                    8 * n / t_apply / 1e6, t_bwd, 16 * n / t_bwd / 1e6, t_meas,
                                                    4 * n / t_meas / 1e6), flush=True)


if __name__ == "__main__":
    main()
