"""How far apart are two runs of the CPU ORACLE itself on tests/test_loss_curve_gpu.py's task when only the seed of its
stochastic-rounding noise changes?  (Calibrates that test's late-step tolerance; not collected by pytest.)

    python tests/loss_curve_oracle_variance.py > profiles/loss_curve_oracle_variance_rNN.json      # ~5 min of CPU
"""
import json
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa: E402,F401
from oracle import qref_modules, seeding  # noqa: E402
from qtorch_b200 import models  # noqa: E402


def run(seed, steps=200):
    ref = models.resnet_quantized(qref_modules, depth=20, dataset="cifar10").train()
    seeding.reseed_parameters(ref, 1234)
    g = torch.Generator().manual_seed(77)
    teacher = torch.randn(10, 3 * 8 * 8, generator=g)
    data = []
    for _ in range(64):
        x = torch.randn(32, 3, 32, 32, generator=g)
        data.append((x, (F.avg_pool2d(x, 4).flatten(1) @ teacher.t()).argmax(1)))
    opt = torch.optim.SGD(ref.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    crit = torch.nn.CrossEntropyLoss()
    torch.manual_seed(seed)
    losses = []
    for s in range(steps):
        x, y = data[s % 64]
        loss = crit(ref(x), y)
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(float(loss))
    return losses


if __name__ == "__main__":
    res = {str(s): run(s) for s in (99, 7, 12345)}
    sm = {k: np.convolve(v, np.ones(20) / 20, "valid") for k, v in res.items()}
    keys = list(sm)
    spread = {a + "-" + b: float(np.max(np.abs(sm[a] - sm[b]) / sm[a])) for i, a in enumerate(keys) for b in keys[i + 1:]}
    print(json.dumps({"window": 20, "max_rel_smoothed_between_seeds": spread,
                      "final_smoothed": {k: float(v[-1]) for k, v in sm.items()}, "curves": res}))
