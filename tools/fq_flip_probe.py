"""Diagnostic: QConv2d 16->32 1x1 (non-ReLU input, bias): capture what the module hands to qb_conv_i8_fwd."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa
from qtorch_b200 import _abi, quantize as Q
from oracle import qref, qref_modules
shape = (8, 16, 32, 32)
B, Cin, H, W = shape
Cout = 32
gen = torch.Generator().manual_seed(sum(shape) + 32)
def build(ns):
    return ns.QConv2d(16, 32, 1, stride=1, padding=0, bias=True, num_bits=8, num_bits_weight=8, num_bits_grad=8, biprecision=True)
ref, mod = build(qref_modules).train(), build(Q)
with torch.no_grad():
    ref.weight.copy_(torch.randn(ref.weight.shape, generator=gen) * 0.05)
    ref.bias.copy_(torch.randn(ref.bias.shape, generator=gen) * 0.1)
mod.load_state_dict(ref.state_dict())
mod = mod.cuda().train()
x = torch.randn(shape, generator=gen) + 0.5
flat = x.view(8, -1); flat[:, 0], flat[:, 1] = float(x.min()), float(x.max())
cap = {}
orig = _abi.conv_i8_fwd
def spy(codes, wcodes, *a, **k):
    cap["codes"], cap["wcodes"], cap["a"], cap["k"] = codes, wcodes, a, k
    return orig(codes, wcodes, *a, **k)
_abi.conv_i8_fwd = spy
for grad in (False, True):
    mod.weight.requires_grad_(grad); mod.bias.requires_grad_(grad)
    out = mod(x.cuda()).detach().cpu()
    outr = ref(x).detach()
    d = (out - outr).abs()
    print("requires_grad", grad, "rel", float((out - outr).norm() / outr.norm()), "max", float(d.max()))
    lo, hi = float(x.min()), float(x.max())
    rng = cap["a"][9]
    print("  range dev", rng.tolist(), "host", [lo, hi], "equal", float(rng[0]) == torch.tensor(lo).float().item(), float(rng[1]) == torch.tensor(hi).float().item())
    xq = qref.quantize(x, 8, lo, hi)
    scale = torch.tensor(max((hi - lo) / 255.0, 1e-8), dtype=torch.float64).float()
    cref = torch.round((xq - torch.tensor(lo).float()) / scale)
    cm = (cap["codes"].cpu().float() != cref)
    print("  code mismatches:", int(cm.sum()), cm.nonzero().tolist()[:5])
    pixsum = cap["a"][14]
    ps_ref = cap["codes"].permute(0, 2, 3, 1).reshape(-1, Cin).int().sum(1)
    print("  pixsum mismatches:", int((pixsum != ps_ref).sum()))
    bad = (d.amax(1) > 1e-5).nonzero()
    print("  bad pixels:", bad.tolist()[:10])
