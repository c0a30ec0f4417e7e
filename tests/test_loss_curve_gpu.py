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


def _oracle_band(w):
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "loss_curves_oracle.json")
    if not os.path.isfile(path):
        return None, None
    with open(path) as f:
        gold = json.load(f)
    if gold["batch"] != BATCH or gold["lr"] != LR or gold["steps"] < STEPS:
        return None, None
    sm = np.stack([_smooth(np.asarray(c[:STEPS]), w) for c in gold["curves"].values()])
    return sm.min(0), sm.max(0)


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
    # the late-step comparison of the Philox arm is a statement about distributions: test_philox_loss_curves_inside_oracle_band
    lo_b, hi_b = _oracle_band(w)
    if lo_b is not None and len(sr) <= len(lo_b):
        c, h = 0.5 * (lo_b + hi_b)[:len(sr)], 0.5 * (hi_b - lo_b)[:len(sr)]
        assert np.all(np.abs(sr - c)[w:] <= 3.0 * np.maximum(h, 1e-3 * c)[w:] + 0.02 * c[w:]), "replayed-noise run leaves the oracle band"
    if STEPS >= 100:
        assert so[-1] < 0.8 * so[0] and sr[-1] < 0.8 * sr[0] and sp[-1] < 0.8 * sp[0], "no learning in 200 steps"


def test_philox_loss_curves_inside_oracle_band():
    """north_star: "training loss curves must match over 200 steps" - as a statement about DISTRIBUTIONS, because the
    stochastic-rounding noise of the CUDA path (in-kernel Philox) cannot be the oracle's (torch's CPU generator):

    tests/golden/loss_curves_oracle.json holds 8 runs of the CPU oracle on this task that differ only in the noise seed
    (oracle/gen_loss_curves.py).  Their smoothed curves span an envelope [min, max](t).  Five CUDA runs with different
    Philox seeds must have their MEAN curve inside that envelope widened by 1.5x about its centre at every step, each
    single run inside it widened by 3x, and the same mean final loss to within the oracle's own spread."""
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import models, quantize as Q
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "loss_curves_oracle.json")) as f:
        gold = json.load(f)
    assert gold["steps"] >= STEPS and gold["batch"] == BATCH and gold["lr"] == LR
    w = max(1, min(20, STEPS // 4))
    oracle = np.stack([_smooth(np.asarray(c[:STEPS]), w) for c in gold["curves"].values()])
    lo, hi = oracle.min(0), oracle.max(0)
    centre, half = 0.5 * (lo + hi), 0.5 * (hi - lo)
    ref = models.resnet_quantized(qref_modules, depth=20, dataset="cifar10").train()
    seeding.reseed_parameters(ref, gold["init_seed"])
    state = {k: v.clone() for k, v in ref.state_dict().items()}
    gen = torch.Generator().manual_seed(77)
    teacher = torch.randn(10, 3 * 8 * 8, generator=gen)
    data = []
    for _ in range(NBATCH):
        xb = torch.randn(BATCH, 3, 32, 32, generator=gen)
        data.append((xb.to(DEV), (torch.nn.functional.avg_pool2d(xb, 4).flatten(1) @ teacher.t()).argmax(1).to(DEV)))
    crit = torch.nn.CrossEntropyLoss()
    curves = {}
    for seed in (11, 22, 33, 44, 55):
        m = models.resnet_quantized(Q, depth=20, dataset="cifar10")
        m.load_state_dict(state)
        m = m.to(DEV).train()
        opt = torch.optim.SGD(m.parameters(), lr=LR, momentum=0.9, weight_decay=1e-4)
        Q.manual_seed(seed)
        losses = []
        for s in range(STEPS):
            x, y = data[s % NBATCH]
            loss = crit(m(x), y)
            opt.zero_grad()
            loss.backward()
            opt.step()
            losses.append(loss.detach())
        curves[seed] = [float(v) for v in torch.stack(losses).cpu()]
    sm = np.stack([_smooth(np.asarray(c), w) for c in curves.values()])
    mean = sm.mean(0)
    report = {"window": w, "oracle_final": [float(v) for v in oracle[:, -1]], "philox_final": [float(v) for v in sm[:, -1]],
              "max_mean_excursion_in_half_widths": float(np.max(np.abs(mean - centre) / np.maximum(half, 1e-3 * centre))),
              "max_single_excursion_in_half_widths": float(np.max(np.abs(sm - centre) / np.maximum(half, 1e-3 * centre))),
              "max_rel_mean_vs_oracle_mean": float(np.max(np.abs(mean - oracle.mean(0)) / oracle.mean(0))), "curves": curves}
    if os.path.isdir("gpurun_out"):
        with open("gpurun_out/loss_curves_philox_%d.json" % STEPS, "w") as f:
            json.dump(report, f)
    print({k: v for k, v in report.items() if k != "curves"})
    assert np.all(np.isfinite(sm))
    hw = np.maximum(half, 1e-3 * centre)                     # the first steps are deterministic given the data: zero-width band
    first = w                                                # ... so the band test starts once the curves have decorrelated
    assert np.all(np.abs(mean - centre)[first:] <= 1.5 * hw[first:] + 0.01 * centre[first:]), "mean Philox curve leaves the oracle band"
    assert np.all(np.abs(sm - centre)[:, first:] <= 3.0 * hw[first:] + 0.02 * centre[first:]), "a Philox run leaves the 3x band"
    if STEPS >= 100:
        spread = oracle[:, -1].max() - oracle[:, -1].min()
        assert abs(sm[:, -1].mean() - oracle[:, -1].mean()) <= spread, "final loss biased against the oracle"
