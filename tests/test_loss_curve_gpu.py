"""north_star: "training loss curves must match over 200 steps".

ResNet-20 / CIFAR shape (BASELINE configs[0]), 8-bit activations / weights / gradients, RangeBN, bi-precision backward,
trained for QB_LOSS_STEPS (default 200) SGD steps (momentum 0.9, weight decay 1e-4) on a fixed synthetic set of 64
batches whose labels come from a random linear teacher (learnable, but not memorised within the run, so the loss stays
O(1) and a relative comparison means something), three times from the same initial state:

  * the CPU oracle (the reference's algorithm, oracle/qref_modules.py), drawing its own stochastic-rounding noise,
  * the CUDA path with the oracle's noise of the SAME step injected (recorded per step, replayed in backward order),
  * the CUDA path as shipped (in-kernel Philox noise, fused kernels).

Single codes flip on ~1e-6 differences of the contractions, so the trajectories are compared the way the oracle is
compared with itself when only its thread count changes (tests/test_oracle_golden.py): tightly over the first steps,
and as smoothed curves over the whole run.  The three curves are written to gpurun_out/loss_curve_200.json when that
directory exists (committed copy: profiles/loss_curve_200_r01s.json)."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import qref, qref_modules, seeding

pytestmark = pytest.mark.gpu
DEV = "cuda"
STEPS = int(os.environ.get("QB_LOSS_STEPS", "200"))
BATCH, NBATCH, LR = 32, 64, 0.01


def _smooth(v, w):
    v = np.asarray(v, dtype=np.float64)
    return np.convolve(v, np.ones(w) / w, mode="valid")


class _StepReplay:
    def __init__(self):
        self.q = []

    def load(self, tensors):
        self.q = [t.to(DEV) for t in tensors]

    def __call__(self, shape):
        t = self.q.pop(0)
        assert t.numel() == int(np.prod(shape)), (t.shape, shape)
        return t.reshape(shape)


def test_resnet20_loss_curve_200_steps():
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import models, quantize as Q
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        ref = models.resnet_quantized(qref_modules, depth=20, dataset="cifar10").train()
        seeding.reseed_parameters(ref, 1234)
        state = {k: v.clone() for k, v in ref.state_dict().items()}
        nets = {}
        for name in ("replay", "philox"):
            m = models.resnet_quantized(Q, depth=20, dataset="cifar10")
            m.load_state_dict(state)
            nets[name] = m.to(DEV).train()
        gen = torch.Generator().manual_seed(77)
        teacher = torch.randn(10, 3 * 8 * 8, generator=gen)
        data = []
        for _ in range(NBATCH):
            xb = torch.randn(BATCH, 3, 32, 32, generator=gen)
            data.append((xb, (torch.nn.functional.avg_pool2d(xb, 4).flatten(1) @ teacher.t()).argmax(1)))
        crit = torch.nn.CrossEntropyLoss()
        opts = {"oracle": torch.optim.SGD(ref.parameters(), lr=LR, momentum=0.9, weight_decay=1e-4)}
        for name, m in nets.items():
            opts[name] = torch.optim.SGD(m.parameters(), lr=LR, momentum=0.9, weight_decay=1e-4)
        curves = {"oracle": [], "replay": [], "philox": []}
        log, replay = [], _StepReplay()

        def recording(shape):
            n = torch.rand(shape) - 0.5
            log.append(n)
            return n

        def one(model, opt, x, y):
            loss = crit(model(x), y)
            opt.zero_grad()
            loss.backward()
            opt.step()
            return float(loss)

        torch.manual_seed(99)
        Q.manual_seed(99)
        for s in range(STEPS):
            x, y = data[s % NBATCH]
            log.clear()
            prev = qref.set_noise_provider(recording)
            try:
                curves["oracle"].append(one(ref, opts["oracle"], x, y))
            finally:
                qref.set_noise_provider(prev)
            xd, yd = x.to(DEV), y.to(DEV)
            replay.load(log)
            prev = Q.set_noise_provider(replay)
            try:
                curves["replay"].append(one(nets["replay"], opts["replay"], xd, yd))
            finally:
                Q.set_noise_provider(prev)
            assert not replay.q, "the CUDA path drew fewer noise tensors than the oracle"
            curves["philox"].append(one(nets["philox"], opts["philox"], xd, yd))
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old

    o, r, p = (np.asarray(curves[k]) for k in ("oracle", "replay", "philox"))
    w = max(1, min(20, STEPS // 4))
    so, sr, sp = _smooth(o, w), _smooth(r, w), _smooth(p, w)
    report = {"steps": STEPS, "batch": BATCH, "distinct_batches": NBATCH, "lr": LR, "window": w, "curves": curves,
              "max_rel_first5_replay": float(np.max(np.abs(r[:5] - o[:5]) / o[:5])),
              "max_rel_smoothed_replay": float(np.max(np.abs(sr - so) / so)),
              "max_rel_smoothed_philox": float(np.max(np.abs(sp - so) / so)),
              "final_smoothed": {"oracle": float(so[-1]), "replay": float(sr[-1]), "philox": float(sp[-1])}}
    if os.path.isdir("gpurun_out"):
        with open("gpurun_out/loss_curve_%d.json" % STEPS, "w") as f:
            json.dump(report, f)
    print({k: v for k, v in report.items() if k != "curves"})
    assert np.all(np.isfinite(o)) and np.all(np.isfinite(r)) and np.all(np.isfinite(p))
    # same noise, same data: the first steps agree to code-flip accuracy (this initialisation produces logits of
    # magnitude ~10, so a few flipped 8-bit codes move the loss by ~1e-2 relative; tests/test_modules_gpu.py holds the
    # tight single-step comparisons)
    assert abs(r[0] - o[0]) <= 2e-2 * abs(o[0])
    np.testing.assert_allclose(r[:5], o[:5], rtol=6e-2)
    # first half: the smoothed curves coincide (measured: within 2 %).  Second half: the nets enter the fast-learning phase
    # at slightly different steps and the trajectories separate the way the ORACLE's own do when only its noise seed
    # changes (tests/loss_curve_oracle_variance.py, profiles/loss_curve_oracle_variance_r01k.json: three seeds spread by
    # up to 6 % of the smoothed loss, 1.82 .. 1.93 at the end), so the bound there is that band with a safety factor
    half = min(len(so), 100)
    np.testing.assert_allclose(sr[:half], so[:half], rtol=0.05)
    np.testing.assert_allclose(sp[:half], so[:half], rtol=0.05)
    np.testing.assert_allclose(sr, so, rtol=0.2)
    np.testing.assert_allclose(sp, so, rtol=0.2)
    if STEPS >= 100:
        assert so[-1] < 0.8 * so[0] and sr[-1] < 0.8 * sr[0] and sp[-1] < 0.8 * sp[0], "no learning in 200 steps"
