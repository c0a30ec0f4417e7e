"""Probe of the native backward contractions (conv_bwd.cu) against fp64 references; prints one line per case.

    python tools/bwd_probe.py [pack|dgrad|wgrad|all] [--bench]
"""
import os
import sys
import time

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa: E402,F401
from qtorch_b200 import _abi  # noqa: E402

DEV = "cuda"

CASES = [  # B, Cin, H, W, Cout, k, stride, pad
    (2, 64, 16, 16, 64, 1, 1, 0),
    (2, 64, 16, 16, 64, 3, 1, 1),
    (4, 64, 56, 56, 64, 3, 1, 1),
    (3, 64, 56, 56, 128, 3, 2, 1),
    (2, 64, 56, 56, 128, 1, 2, 0),
    (4, 128, 28, 28, 128, 3, 1, 1),
    (4, 128, 28, 28, 256, 3, 2, 1),
    (5, 256, 14, 14, 256, 3, 1, 1),
    (5, 256, 14, 14, 512, 1, 2, 0),
    (7, 512, 7, 7, 512, 3, 1, 1),
    (2, 32, 112, 112, 64, 1, 1, 0),
    (8, 16, 32, 32, 16, 3, 1, 1),
    (8, 16, 32, 32, 32, 3, 2, 1),
    (8, 16, 32, 32, 32, 1, 2, 0),
    (8, 32, 16, 16, 64, 3, 2, 1),
    (2, 32, 20, 150, 48, 3, 1, 1),
    (1, 96, 9, 11, 80, 3, 2, 1),
    (256, 512, 1, 1, 1000, 1, 1, 0),
    (32, 64, 1, 1, 10, 1, 1, 0),
]


def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-300))


def probe_pack():
    torch.manual_seed(0)
    for n in (8 * 1000, 8 * 1000 + 5, 1 << 20):
        g = (torch.randn(n, device=DEV) * 1e-3)
        g[::7] = 0
        noise = torch.rand(n, device=DEV) - 0.5
        rng = _abi.minmax(g)
        codes = torch.empty(n, dtype=torch.uint8, device=DEV)
        ref = _abi.fakequant_grad(g, 8, range_dev=rng, noise=noise, stochastic=True, codes=codes)
        gq, hi, lo = _abi.bwd_pack(g, rng, 8, noise=noise, stochastic=True)
        torch.cuda.synchronize()
        lo_, hi_ = float(rng[0]), float(rng[1])
        scale = max((hi_ - lo_) / 255.0, 1e-8)
        zp = int(min(max(0.0 - lo_ / scale, 0.0), 255.0))
        want = codes.float() - zp
        ok_codes = bool(torch.equal(gq.float(), want))
        deq = gq.float() * torch.tensor(scale, dtype=torch.float64).float().item()
        split_err = float(((hi.double() + lo.double()) - g.double()).abs().max() / g.abs().max())
        print("pack n=%d codes_exact=%s deq_vs_ref=%.2e split_err=%.2e" % (n, ok_codes, rel(deq, ref), split_err))
        gq2, _, _ = _abi.bwd_pack(g, rng, 8, stochastic=True, seed=123, want_split=False)
        d = (gq2.float() - want).abs().max()
        print("   philox codes within one step of the injected-noise codes: max |d| = %g" % float(d))


def probe_dgrad(bench=False):
    for case in CASES:
        B, Cin, H, W, Cout, k, stride, pad = case
        Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
        gen = torch.Generator().manual_seed(sum(case))
        a = torch.randint(-255, 256, (B, Cout, Ho, Wo), generator=gen).float()
        a[torch.rand(a.shape, generator=gen) < 0.3] = 0
        w = torch.randn(Cout, Cin, k, k, generator=gen) * 0.05
        ad = a.to(DEV).contiguous(memory_format=torch.channels_last).to(torch.bfloat16)
        wt = _abi.wt_pack(w.to(DEV))
        try:
            dx = _abi.conv_dgrad(ad, None, wt, (B, Cin, H, W), Cout, k, k, stride, pad, scale_mul=0.5)
            torch.cuda.synchronize()
        except Exception as e:  # noqa: BLE001
            print("dgrad %s FAILED: %s" % (case, e))
            continue
        ref = 0.5 * torch.nn.grad.conv2d_input((B, Cin, H, W), w.double(), a.double(), stride, pad)
        e1 = rel(dx.cpu(), ref)
        # two-plane A operand (generic fp32 gradient)
        gf = torch.randn(B, Cout, Ho, Wo, generator=gen) * 1e-3
        gd = gf.to(DEV).contiguous(memory_format=torch.channels_last)
        _, hi, lo = _abi.bwd_pack(gd, want_codes=False)
        dx2 = _abi.conv_dgrad(hi, lo, wt, (B, Cin, H, W), Cout, k, k, stride, pad)
        torch.cuda.synchronize()
        ref2 = torch.nn.grad.conv2d_input((B, Cin, H, W), w.double(), gf.double(), stride, pad)
        e2 = rel(dx2.cpu(), ref2)
        msg = "dgrad %-40s codes %.2e  split %.2e" % (str(case), e1, e2)
        if bench:
            for _ in range(3):
                _abi.conv_dgrad(ad, None, wt, (B, Cin, H, W), Cout, k, k, stride, pad)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(20):
                _abi.conv_dgrad(ad, None, wt, (B, Cin, H, W), Cout, k, k, stride, pad)
            torch.cuda.synchronize()
            msg += "  %.1f us" % ((time.perf_counter() - t0) / 20 * 1e6)
        print(msg, flush=True)


def probe_wgrad(bench=False):
    for case in CASES:
        B, Cin, H, W, Cout, k, stride, pad = case
        Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
        gen = torch.Generator().manual_seed(sum(case) + 1)
        for lo_val in (0.0, -0.37):
            q = torch.randint(0, 256, (B, Cin, H, W), generator=gen).float()
            g = torch.randn(B, Cout, Ho, Wo, generator=gen) * 1e-3
            hi_val = lo_val + 2.0
            rng = torch.tensor([lo_val, hi_val], device=DEV)
            scale = max((hi_val - lo_val) / 255.0, 1e-8)
            scale32 = torch.tensor(scale, dtype=torch.float64).float().double().item()
            lo32 = torch.tensor(lo_val).float().double().item()
            x_hat = q.double() * scale32 + lo32
            gd = g.to(DEV).contiguous(memory_format=torch.channels_last)
            x16 = q.to(DEV).contiguous(memory_format=torch.channels_last).to(torch.bfloat16)
            _, ghi, glo = _abi.bwd_pack(gd, want_codes=False)
            try:
                dw, db = _abi.conv_wgrad(ghi, glo, gd, x16, Cout, k, k, stride, pad, rng, 8, want_bias=True)
                torch.cuda.synchronize()
            except Exception as e:  # noqa: BLE001
                print("wgrad %s FAILED: %s" % (case, e))
                break
            ref = torch.nn.grad.conv2d_weight(x_hat, (Cout, Cin, k, k), g.double(), stride, pad)
            refb = g.double().sum((0, 2, 3))
            msg = "wgrad %-40s lo=%5.2f dw %.2e  db %.2e" % (str(case), lo_val, rel(dw.cpu(), ref), rel(db.cpu(), refb))
            if bench and lo_val == 0.0:
                for _ in range(3):
                    _abi.conv_wgrad(ghi, glo, gd, x16, Cout, k, k, stride, pad, rng, 8)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                for _ in range(20):
                    _abi.conv_wgrad(ghi, glo, gd, x16, Cout, k, k, stride, pad, rng, 8)
                torch.cuda.synchronize()
                msg += "  %.1f us" % ((time.perf_counter() - t0) / 20 * 1e6)
            print(msg, flush=True)


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    bench = "--bench" in sys.argv
    if what in ("pack", "all"):
        probe_pack()
    if what in ("dgrad", "all"):
        probe_dgrad(bench)
    if what in ("wgrad", "all"):
        probe_wgrad(bench)
