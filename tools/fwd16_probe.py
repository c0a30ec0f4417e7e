"""Diagnostic: QConv2d forward on 16-channel inputs (K padded by TMA zero fill) against the CPU oracle and fp64."""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa: E402,F401
from oracle import qref_modules  # noqa: E402
from qtorch_b200 import quantize as Q  # noqa: E402

DEV = "cuda"
for cfg, shape in [(dict(cin=16, cout=32, k=1, stride=2, pad=0, bias=True), (8, 16, 32, 32)),
                   (dict(cin=16, cout=32, k=1, stride=2, pad=0, bias=False), (8, 16, 32, 32)),
                   (dict(cin=16, cout=32, k=1, stride=1, pad=0, bias=True), (8, 16, 32, 32)),
                   (dict(cin=32, cout=32, k=1, stride=2, pad=0, bias=True), (8, 32, 32, 32)),
                   (dict(cin=16, cout=16, k=3, stride=1, pad=1, bias=True), (8, 16, 32, 32))]:
    for relu in (True, False):
        gen = torch.Generator().manual_seed(sum(shape) + cfg["cout"])

        def build(ns):
            return ns.QConv2d(cfg["cin"], cfg["cout"], cfg["k"], stride=cfg["stride"], padding=cfg["pad"], bias=cfg["bias"],
                              num_bits=8, num_bits_weight=8, num_bits_grad=8, biprecision=True)

        ref, mod = build(qref_modules).train(), build(Q)
        with torch.no_grad():
            ref.weight.copy_(torch.randn(ref.weight.shape, generator=gen) * 0.05)
            if ref.bias is not None:
                ref.bias.copy_(torch.randn(ref.bias.shape, generator=gen) * 0.1)
        mod.load_state_dict(ref.state_dict())
        mod = mod.to(DEV).train()
        x = torch.randn(shape, generator=gen)
        x = x.clamp_min(0) if relu else x + 0.5
        flat = x.view(shape[0], -1)
        flat[:, 0], flat[:, 1] = float(x.min()), float(x.max())
        outr = ref(x).detach()
        out = mod(x.to(DEV)).detach().cpu()
        # fp64 on the oracle's own fake-quantized tensors
        xq = ref.quantize_input(x).detach()
        from oracle import qref
        wq = qref.quantize(ref.weight.detach(), 8, ref.weight_min, ref.weight_max)
        bq = None
        if ref.bias is not None:
            bq = qref.quantize(ref.bias.detach(), 8, ref.bias_min, ref.bias_max).double()
        o64 = F.conv2d(xq.double(), wq.double(), bq, cfg["stride"], cfg["pad"])
        d = (out.double() - o64)
        dr = (outr.double() - o64)
        print(cfg, "relu" if relu else "shift", "| cuda vs fp64 %.2e | oracle vs fp64 %.2e | cuda vs oracle %.2e | max|d| %.2e at %s, |out| rms %.2e"
              % (float(d.norm() / o64.norm()), float(dr.norm() / o64.norm()), float((out - outr).norm() / outr.norm()),
                 float(d.abs().max()), tuple(int(i) for i in (d.abs() == d.abs().max()).nonzero()[0]), float(o64.pow(2).mean().sqrt())))
        bad = (d.abs() > 10 * d.abs().median().clamp_min(1e-9)).sum()
        print("    elements with |d| > 10 x median: %d of %d; per-channel err %s" % (int(bad), d.numel(),
              [round(float(v), 8) for v in d.abs().amax((0, 2, 3))[:8]]))
