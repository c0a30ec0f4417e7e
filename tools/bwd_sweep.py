"""Per-layer timing of the native backward contractions (qb_conv_dgrad / qb_conv_wgrad) on the dense convolution shapes of
ResNet-18 (ImageNet, batch 256) [--model mobilenet: the pointwise convolutions of MobileNet].

    python tools/bwd_sweep.py [--model resnet18|mobilenet|resnet20] [--batch 256] > profiles/bwd_sweep_rNN.txt

Useful work = 2*B*Ho*Wo*Cout*Cin*kh*kw flop per contraction; the kernels execute twice that in bf16 (hi + lo planes), which is
what the fraction of the measured dense bf16 peak (MEASURED_PEAKS.json) is quoted on.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa: E402,F401
from qtorch_b200 import _abi  # noqa: E402

R18 = [("layer1 3x3 64->64 @56", 4, 64, 56, 64, 3, 1, 1), ("layer2.0 3x3 64->128 s2", 1, 64, 56, 128, 3, 2, 1),
       ("layer2.0 1x1 64->128 s2", 1, 64, 56, 128, 1, 2, 0), ("layer2 3x3 128->128 @28", 3, 128, 28, 128, 3, 1, 1),
       ("layer3.0 3x3 128->256 s2", 1, 128, 28, 256, 3, 2, 1), ("layer3.0 1x1 128->256 s2", 1, 128, 28, 256, 1, 2, 0),
       ("layer3 3x3 256->256 @14", 3, 256, 14, 256, 3, 1, 1), ("layer4.0 3x3 256->512 s2", 1, 256, 14, 512, 3, 2, 1),
       ("layer4.0 1x1 256->512 s2", 1, 256, 14, 512, 1, 2, 0), ("layer4 3x3 512->512 @7", 3, 512, 7, 512, 3, 1, 1)]
MBN = [("pw 32->64 @112", 1, 32, 112, 64, 1, 1, 0), ("pw 64->128 @56", 1, 64, 56, 128, 1, 1, 0), ("pw 128->128 @56", 1, 128, 56, 128, 1, 1, 0),
       ("pw 128->256 @28", 1, 128, 28, 256, 1, 1, 0), ("pw 256->256 @28", 1, 256, 28, 256, 1, 1, 0), ("pw 256->512 @14", 1, 256, 14, 512, 1, 1, 0),
       ("pw 512->512 @14", 5, 512, 14, 512, 1, 1, 0), ("pw 512->1024 @7", 1, 512, 7, 1024, 1, 1, 0), ("pw 1024->1024 @7", 1, 1024, 7, 1024, 1, 1, 0)]
R20 = [("3x3 16->16 @32", 6, 16, 32, 16, 3, 1, 1), ("3x3 16->32 s2", 1, 16, 32, 32, 3, 2, 1), ("3x3 32->32 @16", 5, 32, 16, 32, 3, 1, 1),
       ("3x3 32->64 s2", 1, 32, 16, 64, 3, 2, 1), ("3x3 64->64 @8", 5, 64, 8, 64, 3, 1, 1)]


def timeit(fn, reps=10):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        fn()
    tot = 0.0
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    return tot / reps * 1e3     # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet18")
    ap.add_argument("--batch", type=int, default=256)
    args = ap.parse_args()
    peak = 1681.3
    try:
        peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["bf16_tflops"]
    except Exception:  # noqa: BLE001
        pass
    layers = {"resnet18": R18, "mobilenet": MBN, "resnet20": R20}[args.model]
    B = args.batch
    tot_d = tot_w = tot_p = 0.0
    print("# %s batch %d; bf16 dense peak %.0f TFLOP/s (measured cuBLAS burst); L2 flushed between launches" % (args.model, B, peak))
    for name, count, Cin, HW, Cout, k, s, p in layers:
        H = W = HW
        Ho = (H + 2 * p - k) // s + 1
        g = (torch.randn(B, Cout, Ho, Ho, device="cuda") * 1e-3).contiguous(memory_format=torch.channels_last)
        x16 = torch.randint(0, 256, (B, Cin, H, W), device="cuda").float().contiguous(memory_format=torch.channels_last).to(torch.bfloat16)
        w = torch.randn(Cout, Cin, k, k, device="cuda") * 0.05
        wt = _abi.wt_pack(w)
        rng = _abi.minmax(g)
        arng = torch.tensor([0.0, 2.0], device="cuda")
        gq, hi, lo = _abi.bwd_pack(g, rng, 8, stochastic=True, seed=1)
        t_p = timeit(lambda: _abi.bwd_pack(g, rng, 8, stochastic=True, seed=1))
        t_d = timeit(lambda: _abi.conv_dgrad(gq, None, wt, (B, Cin, H, W), Cout, k, k, s, p, scale_range=rng, scale_bits=8))
        t_w = timeit(lambda: _abi.conv_wgrad(hi, lo, g, x16, Cout, k, k, s, p, arng, 8))
        flop = 2.0 * B * Ho * Ho * Cout * Cin * k * k
        print("%-26s x%d  pack %6.1f us | dgrad %7.1f us %6.1f TF/s useful (%.3f of bf16 peak executed) | wgrad %7.1f us %6.1f TF/s useful (%.3f)"
              % (name, count, t_p, t_d, flop / t_d * 1e-6, 2 * flop / t_d * 1e-6 / peak, t_w, flop / t_w * 1e-6, 2 * flop / t_w * 1e-6 / peak),
              flush=True)
        tot_d += count * t_d
        tot_w += count * t_w
        tot_p += count * t_p
    print("total per step: pack %.3f ms, dgrad %.3f ms, wgrad %.3f ms" % (tot_p / 1e3, tot_d / 1e3, tot_w / 1e3))


if __name__ == "__main__":
    main()
