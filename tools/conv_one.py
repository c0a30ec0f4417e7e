"""Forward, input-gradient and weight-gradient contraction of ONE convolution shape, three rounds (for ncu captures):
    python tools/conv_one.py Cin HW Cout k stride pad [B]
    ncu --set full --clock-control none --import-source on -k regex:k_conv -s 6 -c 3 -o X python tools/conv_one.py 128 28 128 3 1 1"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa: E402,F401
from qtorch_b200 import _abi  # noqa: E402

Cin, HW, Cout, k, s, p = (int(v) for v in sys.argv[1:7])
B = int(sys.argv[7]) if len(sys.argv) > 7 else 256
H = W = HW
Ho = (H + 2 * p - k) // s + 1
dev = "cuda"
x = torch.randn(B, Cin, H, W, device=dev).clamp_min_(0).contiguous(memory_format=torch.channels_last)
w = torch.randn(Cout, Cin, k, k, device=dev) * 0.05
xr = _abi.minmax(x)
codes = torch.empty_like(x, dtype=torch.uint8)
_abi.fakequant_affine(x, 8, range_dev=xr, codes=codes, want_y=False)
pixsum = _abi.rowsum_u8(codes, B * H * W, Cin)
rmin, rmax = torch.empty(Cout, device=dev), torch.empty(Cout, device=dev)
_, wcodes, tapsum = _abi.fakequant_rows_conv(w, 8, rmin, rmax)
g = (torch.randn(B, Cout, Ho, Ho, device=dev) * 1e-3).contiguous(memory_format=torch.channels_last)
x16 = codes.float().to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
wt = _abi.wt_pack(w)
rng = _abi.minmax(g)
gq, hi, lo = _abi.bwd_pack(g, rng, 8, stochastic=True, seed=1)
for _ in range(3):
    _abi.conv_i8_fwd(codes, wcodes, B, H, W, Cin, Cout, k, k, s, p, xr, 8, rmin, rmax, 8, pixsum, tapsum)
    _abi.conv_dgrad(gq, None, wt, (B, Cin, H, W), Cout, k, k, s, p, scale_range=rng, scale_bits=8)
    _abi.conv_wgrad(hi, lo, g, x16, Cout, k, k, s, p, xr, 8)
torch.cuda.synchronize()
print("ok")
