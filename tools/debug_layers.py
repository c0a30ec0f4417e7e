"""Debug aid: first layer whose forward output differs between the CUDA modules and the CPU oracle."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa
from qtorch_b200 import models, quantize as Q
from oracle import qref_modules, seeding

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
torch.backends.cudnn.allow_tf32 = False
ref = models.resnet_quantized(qref_modules, depth=20, dataset="cifar10").train()
seeding.reseed_parameters(ref, 1234)
mod = models.resnet_quantized(Q, depth=20, dataset="cifar10")
mod.load_state_dict(ref.state_dict())
mod = mod.cuda().train()
x, y = seeding.synthetic_batch((B, 3, 32, 32), 10, seed=123)
outs_r, outs_m = {}, {}
for n, m in ref.named_modules():
    if not list(m.children()) or n.endswith("bn1") or n.endswith("bn2"):
        m.register_forward_hook(lambda mod_, i, o, n=n: outs_r.__setitem__(n, o.detach().clone()))
for n, m in mod.named_modules():
    if not list(m.children()) or n.endswith("bn1") or n.endswith("bn2"):
        m.register_forward_hook(lambda mod_, i, o, n=n: outs_m.__setitem__(n, o.detach().cpu().clone()))
lr, lm = ref(x), mod(x.cuda())
for n in outs_r:
    if n in outs_m and outs_r[n].shape == outs_m[n].shape:
        a, b = outs_m[n].float(), outs_r[n].float()
        rel = float((a - b).norm() / (b.norm() + 1e-30))
        print("%-40s rel %.3e  max|d| %.3e" % (n, rel, float((a - b).abs().max())))
