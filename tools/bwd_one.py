"""One backward contraction of a chosen shape (for ncu captures): python tools/bwd_one.py Cin HW Cout k stride pad [B]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa
from qtorch_b200 import _abi
Cin, HW, Cout, k, s, p = (int(v) for v in sys.argv[1:7])
B = int(sys.argv[7]) if len(sys.argv) > 7 else 256
H = W = HW
Ho = (H + 2 * p - k) // s + 1
g = (torch.randn(B, Cout, Ho, Ho, device="cuda") * 1e-3).contiguous(memory_format=torch.channels_last)
x16 = torch.randint(0, 256, (B, Cin, H, W), device="cuda").float().contiguous(memory_format=torch.channels_last).to(torch.bfloat16)
wt = _abi.wt_pack(torch.randn(Cout, Cin, k, k, device="cuda") * 0.05)
rng = _abi.minmax(g)
arng = torch.tensor([0.0, 2.0], device="cuda")
gq, hi, lo = _abi.bwd_pack(g, rng, 8, stochastic=True, seed=1)
for _ in range(3):
    _abi.conv_dgrad(gq, None, wt, (B, Cin, H, W), Cout, k, k, s, p, scale_range=rng, scale_bits=8)
    _abi.conv_wgrad(hi, lo, g, x16, Cout, k, k, s, p, arng, 8)
torch.cuda.synchronize()
print("ok")
