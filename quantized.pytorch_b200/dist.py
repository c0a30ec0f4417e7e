"""Batch-sharded training: one process per GPU, torch.distributed (NCCL over NVLink 5 / NVSwitch).

The reference runs multi-GPU through ``nn.DataParallel`` (main.py:344-345): every replica computes its
QuantMeasure / RangeBN / gradient ranges on its own shard and the replicas' buffer updates are dropped, so its
numerics change with the GPU count.  Here the N-rank run is DEFINED to equal the single-process run on the
global batch: besides the gradient all-reduce, the tiny per-layer statistics are
all-reduced so that every rank derives the same ranges:

  QuantMeasure   double[8]   {sum_b min_b, sum_b max_b, sum x, sum x^2 | min | max | n, B}   SUM / MIN / MAX
  RangeBN fwd    double[3C]  {sum_k chunk-max, sum_k chunk-min, sum x}                      SUM (ranks own whole chunks)
  RangeBN bwd    double[2C]  {sum gq, sum gq*xq}                                             SUM
  grad quantizer float[2]    {min, max}  -> MAX over {-min, max}                             MAX

These are latency-bound messages (16 B ... 12 KB): one NCCL all-reduce each, enqueued on the compute stream so
no host synchronisation is introduced.  Gradients are all-reduced (SUM, then scaled by 1/world) in one flat
bucket per dtype after backward.
"""
import torch
import torch.distributed as dist

from . import quantize as Q

_state = {"device": None, "world": 1, "rank": 0}


class StatReducer:
    """All-reduces the statistic blocks IN PLACE (interface expected by quantize.set_stat_reducer)."""

    def __init__(self, group=None):
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)

    # QuantMeasure: entries 0-3 and 6-7 are sums, 4 is a min, 5 a max.  One SUM plus one MAX message.
    def measure(self, stats):
        ext = torch.stack([-stats[4], stats[5]])
        stats[4:6].zero_()                      # (a Python-scalar setitem is a host->device copy: illegal under capture)
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=self.group)
        dist.all_reduce(ext, op=dist.ReduceOp.MAX, group=self.group)
        stats[4:5].copy_(-ext[0:1])
        stats[5:6].copy_(ext[1:2])

    def rangebn_layout(self, B_local, HW, num_chunks):
        """-> (chunks owned by this rank, global element count per channel).  Contiguous batch shards hold whole
        chunks when num_chunks % world == 0 (16 chunks on 1/2/4/8 ranks)."""
        if num_chunks % self.world != 0:
            raise Q.QB200Error("RangeBN: num_chunks=%d must be divisible by world size %d" % (num_chunks, self.world))
        local = num_chunks // self.world
        if (B_local * HW) % local != 0:
            raise Q.QB200Error("RangeBN: local B*H*W=%d not divisible by the %d local chunks" % (B_local * HW, local))
        return local, float(B_local * HW * self.world)

    def rangebn(self, red):
        dist.all_reduce(red, op=dist.ReduceOp.SUM, group=self.group)

    def rangebn_bwd(self, sums):
        dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=self.group)

    def grad_range(self, rng):
        ext = torch.stack([-rng[0], rng[1]])
        dist.all_reduce(ext, op=dist.ReduceOp.MAX, group=self.group)
        rng[0:1].copy_(-ext[0:1])
        rng[1:2].copy_(ext[1:2])


class PeerReducer(StatReducer):
    """The same reductions as ONE one-shot kernel each over NVLink peer memory (csrc/allreduce_small.cu) instead of one or
    two NCCL all-reduces plus packing kernels: these messages are latency bound and on every layer's critical path.
    The exchange buffer is torch symmetric memory (peer-mapped on every rank of the node)."""

    SLOT_BYTES = 64 * 1024          # largest block: RangeBN double[3*C], C <= 2730

    def __init__(self, device, group=None):
        super().__init__(group)
        import ctypes
        import torch.distributed._symmetric_memory as symm
        from . import _abi
        self._abi = _abi
        nbytes = _abi.load().qb_allreduce_small_buffer_bytes(self.world, self.SLOT_BYTES)
        self.buf = symm.empty(nbytes, dtype=torch.uint8, device=device)
        self.handle = symm.rendezvous(self.buf, dist.group.WORLD if group is None else group)
        self.buf.zero_()
        torch.cuda.synchronize(device)
        dist.barrier(group=group)                      # every buffer is zero before anyone's first exchange
        self.ptrs = (ctypes.c_uint64 * self.world)(*[int(p) for p in self.handle.buffer_ptrs])

    def _ar(self, t, mode):
        if t.numel() * t.element_size() > self.SLOT_BYTES:
            return False
        self._abi.allreduce_small(t, mode, self.ptrs, self.rank, self.world, self.SLOT_BYTES)
        return True

    def measure(self, stats):
        if not self._ar(stats, 1):
            super().measure(stats)

    def rangebn(self, red):
        if not self._ar(red, 0):
            super().rangebn(red)

    def rangebn_bwd(self, sums):
        if not self._ar(sums, 0):
            super().rangebn_bwd(sums)

    def grad_range(self, rng):
        if not self._ar(rng, 2):
            super().grad_range(rng)

    def error(self):
        """True if an exchange timed out waiting for a peer.  The kernel traps in that case (a timeout is fatal: reducing
        stale slots would de-synchronise the ranks), so this is only ever seen by a process inspecting a peer's buffer."""
        return bool(self.buf[8:12].view(torch.int32).item())

    def check(self):
        """Host-side watchdog, call once per step (synchronises): raises if any exchange of this rank failed."""
        if self.error():
            raise Q.QB200Error("cross-rank statistic exchange timed out on rank %d (a peer did not arrive within 120 s)"
                               % self.rank)


def init(device=None, group=None, peer=None):
    """Install the cross-rank statistic reducer.  torch.distributed must already be initialised.  On CUDA the
    peer-memory reducer is used (``peer=False`` or QB_PEER_REDUCE=0 selects plain NCCL all-reduces)."""
    import os
    if not dist.is_initialized():
        raise RuntimeError("torch.distributed is not initialised")
    _state.update(device=device, world=dist.get_world_size(group), rank=dist.get_rank(group), group=group)
    if _state["world"] > 1:
        if peer is None:
            peer = os.environ.get("QB_PEER_REDUCE", "1") != "0"
        red = None
        if peer and device is not None and torch.device(device).type == "cuda" and dist.get_backend(group) == "nccl":
            try:
                red = PeerReducer(torch.device(device), group)
            except Exception as e:                     # no symmetric memory on this system: NCCL path, same semantics
                import sys
                sys.stderr.write("qtorch_b200.dist: peer-memory reducer unavailable (%s); using NCCL all-reduces\n"
                                 % (str(e).splitlines()[0] if str(e) else type(e).__name__))
        Q.set_stat_reducer(red if red is not None else StatReducer(group))
        Q.set_noise_rank(_state["rank"])
        _state["reducer"] = Q._reducer if hasattr(Q, "_reducer") else None
    return _state["rank"], _state["world"]


def check():
    """Raise if the peer-memory exchange of this rank reported a failure (PeerReducer.check)."""
    red = getattr(Q, "_reducer", None)
    if isinstance(red, PeerReducer):
        red.check()


def shutdown():
    Q.set_stat_reducer(None)
    Q.set_noise_rank(0)


def broadcast_parameters(model, src=0, group=None):
    """Replicas start identical (what DataParallel's per-step replicate achieves in the reference)."""
    with torch.no_grad():
        for t in list(model.parameters()) + list(model.buffers()):
            dist.broadcast(t, src=src, group=group)


def allreduce_gradients(model, group=None, average=True):
    """Flat-bucket all-reduce of the batch-sharded gradients (46.8 MB for ResNet-18): one NCCL call, NVLS-eligible.

    ``average=True``  - every rank's loss is the MEAN over its own shard: gradients are summed and divided by world.
    ``average=False`` - every rank's loss is (sum over its shard) / GLOBAL batch: gradients are summed only.
    RangeBN's gamma / beta gradients are different: they are computed from the all-reduced backward sums
    (StatReducer.rangebn_bwd), i.e. they are already the GLOBAL gradient on every rank and must not be summed again;
    with a local-mean loss they come out world x too large (the upstream gradients are), so they are divided instead."""
    world = dist.get_world_size(group)
    replicated = set()
    for m in model.modules():
        if isinstance(m, Q.RangeBN):
            replicated.update(id(p) for p in (m.weight, m.bias) if p is not None)
    grads = [p.grad for p in model.parameters() if p.grad is not None and id(p) not in replicated]
    if grads:
        flat = torch._utils._flatten_dense_tensors(grads)
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        if average:
            flat.div_(world)
        for g, f in zip(grads, torch._utils._unflatten_dense_tensors(flat, grads)):
            g.copy_(f)
    if average:
        for p in model.parameters():
            if p.grad is not None and id(p) in replicated:
                p.grad.div_(world)


class OverlappedGradReducer:
    """Gradient all-reduce overlapped with backward (what DistributedDataParallel's buckets do, for this package's explicit
    reduction): the parameters are grouped, in reverse registration order (the order backward produces their gradients), into
    buckets of ~``bucket_bytes``; a post-accumulate hook on every parameter counts its bucket down and the last one to arrive
    launches the bucket's asynchronous all-reduce (NCCL: on the process group's own stream, concurrent with the rest of
    backward); ``finish()`` after ``loss.backward()`` waits for the handles, applies the 1/world scaling and writes the reduced
    values back into ``.grad``.  Same result as ``allreduce_gradients`` (a SUM all-reduce is element-wise: the grouping does not
    change any sum), including the treatment of RangeBN's gamma / beta.  Opt-in; not capturable in a CUDA graph together with
    the hooks' Python bookkeeping, so the benchmark's captured step keeps the flat-bucket form."""

    def __init__(self, model, group=None, average=True, bucket_bytes=8 << 20):
        self.group, self.average = group, average
        self.world = dist.get_world_size(group)
        replicated = set()
        for m in model.modules():
            if isinstance(m, Q.RangeBN):
                replicated.update(id(p) for p in (m.weight, m.bias) if p is not None)
        self.replicated = [p for p in model.parameters() if id(p) in replicated and p.requires_grad]
        params = [p for p in model.parameters() if id(p) not in replicated and p.requires_grad]
        self.buckets, cur, size = [], [], 0
        for p in reversed(params):
            cur.append(p)
            size += p.numel() * p.element_size()
            if size >= bucket_bytes:
                self.buckets.append(cur)
                cur, size = [], 0
        if cur:
            self.buckets.append(cur)
        self._where = {id(p): i for i, b in enumerate(self.buckets) for p in b}
        self._left = [len(b) for b in self.buckets]
        self._pending = []
        self._hooks = [p.register_post_accumulate_grad_hook(self._arrived) for p in params]

    def _arrived(self, p):
        i = self._where[id(p)]
        self._left[i] -= 1
        if self._left[i] == 0:
            grads = [q.grad for q in self.buckets[i]]
            flat = torch._utils._flatten_dense_tensors(grads)
            self._pending.append((i, flat, grads, dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)))

    def finish(self):
        """Call once after backward: every bucket is reduced, scaled and back in ``.grad``; counters are re-armed."""
        done = set()
        for i, flat, grads, work in self._pending:
            work.wait()
            if self.average:
                flat.div_(self.world)
            for g, f in zip(grads, torch._utils._unflatten_dense_tensors(flat, grads)):
                g.copy_(f)
            done.add(i)
        self._pending = []
        for i, b in enumerate(self.buckets):               # parameters that received no gradient this step leave a bucket open
            if i not in done:
                grads = [q.grad for q in b if q.grad is not None]
                if grads:
                    flat = torch._utils._flatten_dense_tensors(grads)
                    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)
                    if self.average:
                        flat.div_(self.world)
                    for g, f in zip(grads, torch._utils._unflatten_dense_tensors(flat, grads)):
                        g.copy_(f)
        self._left = [len(b) for b in self.buckets]
        if self.average:
            for p in self.replicated:
                if p.grad is not None:
                    p.grad.div_(self.world)

    def remove(self):
        for h in self._hooks:
            h.remove()
        self._hooks = []


def shard_batch(x, rank=None, world=None):
    """Contiguous slice of the global batch for this rank (so RangeBN chunks map to whole ranks)."""
    rank = _state["rank"] if rank is None else rank
    world = _state["world"] if world is None else world
    per = x.shape[0] // world
    return x[rank * per:(rank + 1) * per]
