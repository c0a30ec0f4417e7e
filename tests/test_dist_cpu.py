"""World-size-2 `gloo` tests (CPU) of the cross-rank statistic reducers in quantized.pytorch_b200/dist.py.

The batch-sharded run is DEFINED to equal the single-process run on the global batch: each
rank produces its statistic block from its contiguous batch shard, the reducer combines them, and the result must be
the block a single process would have produced from the whole batch.  The blocks are built here with plain torch in
exactly the layout the kernels emit (measure.cu: double[8]; rangebn*.cu: double[3C] / double[2C]; float[2] ranges)."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def measure_block(x):
    """double[8] of qb_measure_stats for x [B, ...]."""
    rows = x.reshape(x.shape[0], -1).double()
    return torch.tensor([rows.min(1)[0].sum(), rows.max(1)[0].sum(), rows.sum(), (rows * rows).sum(), rows.min(), rows.max(),
                         float(rows.numel()), float(rows.shape[0])], dtype=torch.float64)


def rangebn_block(x, chunks):
    """double[3C] of qb_rangebn_stats: per channel sum over chunks of chunk max, of chunk min, sum x."""
    B, C = x.shape[0], x.shape[1]
    y = x.transpose(0, 1).contiguous().view(C, chunks, -1).double()
    return torch.cat([y.max(-1)[0].sum(-1), y.min(-1)[0].sum(-1), y.sum((1, 2))])


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import qtorch_b200  # noqa: F401
        from qtorch_b200 import dist as qdist, quantize as Q

        r, w = qdist.init()
        assert (r, w) == (rank, world)
        red = Q._reducer
        assert red is not None
        gen = torch.Generator().manual_seed(7)
        x = torch.randn(8, 6, 5, 5, generator=gen) * 2 + 0.3          # the GLOBAL batch, identical on both ranks
        g = torch.randn(8, 6, 5, 5, generator=gen) * 1e-3
        xs, gs = qdist.shard_batch(x), qdist.shard_batch(g)
        assert xs.shape[0] == 4 and torch.equal(xs, x[rank * 4:(rank + 1) * 4])

        # QuantMeasure block: sums add, extrema combine
        st = measure_block(xs)
        red.measure(st)
        ref = measure_block(x)
        assert torch.allclose(st, ref, rtol=1e-14, atol=0), (st, ref)
        assert st[4] == ref[4] and st[5] == ref[5] and st[6] == ref[6] and st[7] == 8

        # RangeBN: 16 global chunks -> 8 whole chunks per rank; 8*25 = 200 elements per channel, 100 per rank
        K = 4                                                          # 200 % 4 == 0, 100 % 2 == 0
        local, count = red.rangebn_layout(xs.shape[0], 25, K)
        assert (local, count) == (2, 200.0)
        blk = rangebn_block(xs, local)
        red.rangebn(blk)
        assert torch.allclose(blk, rangebn_block(x, K), rtol=1e-14, atol=0)
        with pytest.raises(Q.QB200Error):
            red.rangebn_layout(4, 25, 3)                               # 3 chunks do not split over 2 ranks

        # RangeBN backward sums
        sums = torch.cat([gs.double().sum((0, 2, 3)), (gs.double() * xs.double()).sum((0, 2, 3))])
        red.rangebn_bwd(sums)
        assert torch.allclose(sums, torch.cat([g.double().sum((0, 2, 3)), (g.double() * x.double()).sum((0, 2, 3))]),
                              rtol=1e-13, atol=0)

        # gradient-quantizer range: exact min / max over ranks
        rng = torch.tensor([float(gs.min()), float(gs.max())])
        red.grad_range(rng)
        assert rng.tolist() == [float(g.min()), float(g.max())]

        # flat-bucket gradient all-reduce (average) and parameter broadcast
        torch.manual_seed(100 + rank)
        m = torch.nn.Sequential(torch.nn.Linear(5, 3), torch.nn.Linear(3, 2))
        qdist.broadcast_parameters(m)
        flat = torch.cat([p.detach().reshape(-1) for p in m.parameters()])
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        assert torch.equal(gathered[0], gathered[1])
        for i, p in enumerate(m.parameters()):
            p.grad = torch.full_like(p, float(rank + 1) * (i + 1))
        qdist.allreduce_gradients(m)
        for i, p in enumerate(m.parameters()):
            assert torch.allclose(p.grad, torch.full_like(p, 1.5 * (i + 1)))
        # the same reduction overlapped with backward: buckets launched from post-accumulate hooks as their gradients arrive
        torch.manual_seed(5)
        net = torch.nn.Sequential(torch.nn.Linear(6, 16), torch.nn.ReLU(), torch.nn.Linear(16, 16), torch.nn.ReLU(),
                                  torch.nn.Linear(16, 4), Q.RangeBN(4))
        qdist.broadcast_parameters(net)
        twin = torch.nn.Sequential(torch.nn.Linear(6, 16), torch.nn.ReLU(), torch.nn.Linear(16, 16), torch.nn.ReLU(),
                                   torch.nn.Linear(16, 4), Q.RangeBN(4))
        twin.load_state_dict(net.state_dict())
        ovl = qdist.OverlappedGradReducer(net, bucket_bytes=200)       # small buckets: several of them, in backward order
        assert len(ovl.buckets) >= 3 and ovl.buckets[0][0] is net[4].bias
        xin = torch.randn(8, 6, generator=torch.Generator().manual_seed(40 + rank))
        for step in range(2):                                          # counters re-arm between steps
            for mdl in (net, twin):
                for p_ in mdl.parameters():
                    p_.grad = None
            for mdl in (net, twin):
                out = mdl[:5](xin * (step + 1))
                gam = (out * mdl[5].weight + mdl[5].bias).square().mean()   # gamma / beta get (replicated-style) gradients too
                gam.backward()
            ovl.finish()
            qdist.allreduce_gradients(twin)
            for a, b in zip(net.parameters(), twin.parameters()):
                assert torch.equal(a.grad, b.grad), (step, a.shape)
        ovl.remove()
        qdist.shutdown()
        assert Q._reducer is None
        ret[rank] = "ok"
    except BaseException as e:  # noqa: BLE001 - report to the parent
        ret[rank] = "%s: %s" % (type(e).__name__, e)
        raise
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_stat_reducers_world2_gloo():
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
    for p in procs:
        if p.is_alive():
            p.terminate()
    assert dict(ret) == {0: "ok", 1: "ok"}, dict(ret)
