"""Golden loss curves of the CPU ORACLE for tests/test_loss_curve_gpu.py (test infrastructure; never on the product path).

north_star: "training loss curves must match over 200 steps".  The task is the one of tests/test_loss_curve_gpu.py
(ResNet-20 / CIFAR shape, 8-bit activations / weights / gradients, RangeBN, bi-precision backward, SGD momentum 0.9,
weight decay 1e-4, lr 0.01, batch 32, 64 distinct batches labelled by a random linear teacher, initial state
``seeding.reseed_parameters(.., 1234)``).  Only the seed of the stochastic-rounding noise differs between the runs, so the
spread of these curves IS the run-to-run band of the reference algorithm; the GPU test asserts that the curves of the
CUDA path (several Philox seeds) lie inside it.

    python -m oracle.gen_loss_curves            # ~10 min on 8 cores -> tests/golden/loss_curves_oracle.json
"""
import json
import multiprocessing as mp
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SEEDS = (99, 7, 12345, 1, 2, 3, 4, 5)
STEPS, BATCH, NBATCH, LR = 200, 32, 64, 0.01


def task_data():
    """The fixed synthetic task (same generator calls as tests/test_loss_curve_gpu.py)."""
    import torch
    import torch.nn.functional as F
    g = torch.Generator().manual_seed(77)
    teacher = torch.randn(10, 3 * 8 * 8, generator=g)
    data = []
    for _ in range(NBATCH):
        x = torch.randn(BATCH, 3, 32, 32, generator=g)
        data.append((x, (F.avg_pool2d(x, 4).flatten(1) @ teacher.t()).argmax(1)))
    return data


def run(seed, steps=STEPS):
    import torch
    import qtorch_b200  # noqa: F401
    from oracle import qref_modules, seeding
    from qtorch_b200 import models
    torch.set_num_threads(1)
    ref = models.resnet_quantized(qref_modules, depth=20, dataset="cifar10").train()
    seeding.reseed_parameters(ref, 1234)
    data = task_data()
    opt = torch.optim.SGD(ref.parameters(), lr=LR, momentum=0.9, weight_decay=1e-4)
    crit = torch.nn.CrossEntropyLoss()
    torch.manual_seed(seed)
    losses = []
    for s in range(steps):
        x, y = data[s % NBATCH]
        loss = crit(ref(x), y)
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(float(loss))
    return seed, losses


if __name__ == "__main__":
    with mp.get_context("spawn").Pool(min(len(SEEDS), os.cpu_count() or 1)) as pool:
        res = dict(pool.map(run, SEEDS))
    out = {"steps": STEPS, "batch": BATCH, "distinct_batches": NBATCH, "lr": LR, "init_seed": 1234,
           "noise_seeds": list(SEEDS), "curves": {str(k): res[k] for k in SEEDS}}
    path = os.path.join(ROOT, "tests", "golden", "loss_curves_oracle.json")
    with open(path, "w") as f:
        json.dump(out, f)
    print("wrote", path)
