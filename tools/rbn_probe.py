"""RangeBN fused kernels at one ResNet-18 shape (default layer1: 256x64x56x56), for timing / ncu."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa
from qtorch_b200 import _abi

B, C, H = [int(a) for a in (sys.argv[1:4] if len(sys.argv) > 3 else (256, 64, 56))]
CODES = os.environ.get("RBN_CODES", "1") != "0"      # x_hat handed between the passes as uint8 codes
dev = torch.device("cuda")
HW = H * H
x = (torch.randn(B, C, H, H, device=dev) * 2).contiguous(memory_format=torch.channels_last)
g = (torch.randn(B, C, H, H, device=dev) * 1e-3).contiguous(memory_format=torch.channels_last)
g = g * (torch.rand_like(g) > 0.5)
gamma, beta = torch.rand(C, device=dev) + 0.5, torch.randn(C, device=dev)
rm, rv = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
qr = _abi.minmax(x)
xc = torch.empty(x.numel(), dtype=torch.uint8, device=dev) if CODES else None


def fwd():
    red, amax, amin = _abi.rangebn_stats(x, B, C, HW, 16, nhwc=True, q_range=qr, q_bits=8, x_codes_out=xc)
    vec = _abi.rangebn_finalize(red, C, 16, B * HW, B * HW / 16, 1e-5, 0.1, True, 8, gamma, beta, rm, rv)
    out = _abi.rangebn_apply(x, B, C, HW, vec, True, True, nhwc=True, q_range=qr, q_bits=8, x_codes=xc)
    return vec, amax, amin, out


def bwd(vec, amax, amin):
    r = _abi.minmax(g)
    gq = dict(range=r, bits=8, noise=None, stochastic=True, seed=7, step_dev=None)
    return _abi.rangebn_backward(g, x, B, C, HW, 16, B * HW, B * HW / 16, True, vec, True, amax, amin, nhwc=True, q_range=qr,
                                 q_bits=8, grad_quant=gq, want_dx_range=True, x_codes=xc)


def fused(vec, res=None):
    """normalise + (add) + ReLU + next layer's statistics + 1-bit mask from the codes (RangeBN.forward_relu)"""
    return _abi.rangebn_apply_relu_measure(xc, x, B, C, HW, vec, True, True, qr, 8, residual=res, want_bits=True)


codes8 = torch.empty_like(x, dtype=torch.uint8)
codes16 = torch.empty_like(x, dtype=torch.bfloat16)


def others(bits):
    """the quantizer / pack / ReLU-backward passes that sit between the RangeBN passes in a training step"""
    r = _abi.minmax(g)
    _abi.bwd_pack(g, range_dev=r, bits=8, seed=3)
    _abi.fakequant_affine_rowsum(x, B * HW, C, 8, qr, codes8, codes16)
    _abi.relu_bwd_minmax_bits(g, bits, want_dx=False)
    _abi.relu_bwd_minmax_bits(g, bits, want_dx=True)
    _abi.measure_stats(x)


vec, amax, amin, out = fwd()
for _ in range(3):
    fwd()
    bwd(vec, amax, amin)
    if CODES:
        fused(vec)
        others(fused(vec, g)[2])
torch.cuda.synchronize()
_abi.profile_begin()
for _ in range(5):
    fwd()
    bwd(vec, amax, amin)
    if CODES:
        fused(vec)
        others(fused(vec, g)[2])
torch.cuda.synchronize()
recs = _abi.profile_end()
agg = {}
for name, nbytes, ms in recs:
    a = agg.setdefault(name, [0.0, 0])
    a[0] += ms
    a[1] += 1
n = x.numel()
for k, (ms, c) in agg.items():
    print("%-36s %.4f ms/call   %.2f B/elem-equivalent at 6551 GB/s" % (k, ms / c, ms / c * 6551e6 / n))
