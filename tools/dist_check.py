"""Batch-sharded training == single-process training on the global batch, checked on real GPUs.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tools/dist_check.py [--steps 3] [--model resnet20_cifar|resnet18]

Every rank first runs the SINGLE-process step sequence on the global batch (reducer off), then the sharded sequence
(reducer on, its contiguous batch slice, gradient all-reduce) from the same initial state, and compares per-step
losses, a set of gradients and all QuantMeasure / RangeBN buffers.  Stochastic-rounding noise is injected from a
seeded generator over the GLOBAL tensor shape and sliced per rank, so both runs round with the same draw."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qtorch_b200  # noqa: E402,F401
from qtorch_b200 import dist as qdist, models, quantize as Q  # noqa: E402


class GlobalNoise:
    """noise(shape_local) = slice of a generator draw over the global shape; the call counter keys the draw."""

    def __init__(self, dev, rank, world, sharded):
        self.dev, self.rank, self.world, self.sharded, self.calls = dev, rank, world, sharded, 0

    def __call__(self, shape):
        self.calls += 1
        gshape = (shape[0] * self.world,) + tuple(shape[1:]) if self.sharded else tuple(shape)
        gen = torch.Generator(device="cpu").manual_seed(1000 + self.calls)
        n = (torch.rand(gshape, generator=gen) - 0.5).to(self.dev)
        if self.sharded:
            per = shape[0]
            n = n[self.rank * per:(self.rank + 1) * per]
        return n.contiguous()


def run(model, x, y, steps, noise, sharded, world):
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9)
    crit = torch.nn.CrossEntropyLoss(reduction="sum")
    losses = []
    prev = Q.set_noise_provider(noise)
    gbatch = x.shape[0] * (world if sharded else 1)
    try:
        for _ in range(steps):
            loss = crit(model(x), y) / gbatch          # sum / global batch: shard gradients ADD to the global one
            opt.zero_grad(set_to_none=False)
            loss.backward()
            if sharded:
                qdist.allreduce_gradients(model, average=False)
                l = loss.detach().clone()
                dist.all_reduce(l)
                losses.append(float(l))
            else:
                losses.append(float(loss))
            grads = {n: p.grad.detach().clone() for n, p in model.named_parameters()}
            opt.step()
    finally:
        Q.set_noise_provider(prev)
    return losses, grads


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--model", default="resnet20_cifar")
    ap.add_argument("--batch", type=int, default=32, help="GLOBAL batch")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    if args.model == "resnet18":
        mk, shape, ncls = (lambda: models.resnet_quantized(Q, depth=18, dataset="imagenet")), (3, 224, 224), 1000
    else:
        mk, shape, ncls = (lambda: models.resnet_quantized(Q, depth=20, dataset="cifar10")), (3, 32, 32), 10
    torch.manual_seed(5)
    ref = mk().to(dev).train()
    shd = mk().to(dev).train()
    shd.load_state_dict(ref.state_dict())
    gen = torch.Generator().manual_seed(11)
    x = torch.randn(args.batch, *shape, generator=gen).to(dev)
    y = torch.randint(0, ncls, (args.batch,), generator=gen).to(dev)

    Q.set_stat_reducer(None)
    l_ref, g_ref = run(ref, x, y, args.steps, GlobalNoise(dev, rank, world, False), False, world)
    qdist.init(dev)
    l_shd, g_shd = run(shd, qdist.shard_batch(x), qdist.shard_batch(y), args.steps, GlobalNoise(dev, rank, world, True),
                       True, world)
    qdist.shutdown()

    def rel(a, b):
        return float((a - b).norm() / (b.norm() + 1e-30))

    grad_rel = max(rel(g_shd[n], g_ref[n]) for n in g_ref)
    worst = max(g_ref, key=lambda n: rel(g_shd[n], g_ref[n]))
    buf_rel = max(rel(b.float(), dict(ref.named_buffers())[n].float()) for n, b in shd.named_buffers())
    par_rel = max(rel(p, dict(ref.named_parameters())[n]) for n, p in shd.named_parameters())
    out = {"world": world, "rank": rank, "model": args.model, "global_batch": args.batch, "steps": args.steps,
           "loss_single": l_ref, "loss_sharded": l_shd,
           "loss_rel": max(abs(a - b) / abs(b) for a, b in zip(l_shd, l_ref)),
           "last_step_grad_rel_l2_max": grad_rel, "worst_grad": worst, "buffers_rel_l2_max": buf_rel,
           "params_rel_l2_max": par_rel}
    if rank == 0:
        print(json.dumps(out))
    ok = out["loss_rel"] < 1e-4 and buf_rel < 1e-4
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
