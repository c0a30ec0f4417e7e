#!/usr/bin/env python
"""bench.py - the reference's headline workload on B200: ResNet-18 (ImageNet shape) 8-bit training step
(8-bit activations / weights / gradients, RangeBN, stochastic gradient rounding), images/s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch 256] [--impl ours|reference] [--no-graph]
                    [--model resnet18|mobilenet|resnet20_cifar] [--mode train|eval] [--scaling weak|strong]

N>1 is launched by torch.distributed.run (one rank per GPU, NCCL).  --scaling weak (default): per-GPU batch stays 256;
--scaling strong: the GLOBAL batch is --batch (256), sharded 256/N per GPU (BASELINE configs[2] as written).
--mode eval: BASELINE configs[1] - ResNet-18 batch-256 evaluation forward (model.eval(), no_grad, running ranges,
deterministic rounding) after one training step has populated the ranges.
One JSON line on stdout (rank 0).  See DESIGN.md section "Measurement" for how each field is obtained.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRICS = {"resnet18": "ResNet-18 8-bit train images/sec", "mobilenet": "MobileNet 8-bit train images/sec",
           "resnet20_cifar": "ResNet-20 (CIFAR) 8-bit train images/sec"}
EVAL_METRICS = {k: v.replace("train", "eval forward") for k, v in METRICS.items()}
UNIT = "images/s"
INT8_PEAK_TOPS = 4500.0        # B200 datasheet, dense int8 (the bound of the QConv2d/QLinear contractions)
DTYPE = ("u8 codes (activations) x s8 codes (weights) -> s32 on tcgen05 (kind::i8) for the forward contractions; bf16 integer "
         "codes x bf16 hi+lo splits -> f32 on tcgen05 (kind::f16) for the backward contractions (operand error 2^-17); fp32 "
         "carrier for fake-quantized tensors and statistics")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=256, help="per-GPU batch")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-graph", action="store_true", help="do not capture the step in a CUDA graph")
    ap.add_argument("--model", default="resnet18", choices=["resnet18", "resnet20_cifar", "mobilenet"])
    ap.add_argument("--mode", default="train", choices=["train", "eval"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--cpu-sample-batch", type=int, default=16)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def build_model(name, q):
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import models
    if name == "resnet18":
        return models.resnet_quantized(q, depth=18, dataset="imagenet"), (3, 224, 224), 1000
    if name == "mobilenet":
        return models.mobilenet_quantized(q), (3, 224, 224), 1000
    return models.resnet_quantized(q, depth=20, dataset="cifar10"), (3, 32, 32), 10


# ------------------------------------------------------------------------------------------------------
# CPU arm: the reference's algorithm restated on torch-CPU (oracle/), all host threads.  The reference tree
# itself does not exist on the GPU box, so kind = "port".
# ------------------------------------------------------------------------------------------------------
def host_cores():
    """Threads the CPU arm uses: every core this process may run on (torchrun exports OMP_NUM_THREADS=1, which would
    otherwise leave the oracle single-threaded on the multi-GPU launches)."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    return max(1, n)


def cpu_reference_run(model_name, batch, steps, warmup, mode="train"):
    from oracle import qref_modules
    torch.set_num_threads(host_cores())
    torch.manual_seed(0)
    model, shape, ncls = build_model(model_name, qref_modules)
    model.train()
    if mode == "eval":
        return cpu_reference_eval(model, shape, ncls, batch, steps, warmup)
    opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9)
    crit = torch.nn.CrossEntropyLoss()
    x = torch.randn(batch, *shape)
    y = torch.randint(0, ncls, (batch,))
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        loss = crit(model(x), y) + model.regularization()
        opt.zero_grad()
        loss.backward()
        opt.step()
        float(loss)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    return batch * len(times) / sum(times), sum(times) / len(times)


def cpu_reference_eval(model, shape, ncls, batch, steps, warmup):
    """configs[1] on the CPU oracle: one training step to populate the running ranges, then timed eval forwards."""
    crit = torch.nn.CrossEntropyLoss()
    x = torch.randn(batch, *shape)
    y = torch.randint(0, ncls, (batch,))
    crit(model(x), y).backward()
    model.eval()
    times = []
    with torch.no_grad():
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            float(model(x).sum())
            if i >= warmup:
                times.append(time.perf_counter() - t0)
    return batch * len(times) / sum(times), sum(times) / len(times)


def main_reference(args, rank):
    if rank != 0:
        return
    cores = host_cores()
    ips, spstep = cpu_reference_run(args.model, args.cpu_sample_batch, args.steps, args.warmup, args.mode)
    sample = "%s %s at batch %d (bounded sample of the batch-%d workload), torch-CPU oracle port, %d threads" % (
        args.model, "train step" if args.mode == "train" else "eval forward", args.cpu_sample_batch, args.batch, cores)
    print(json.dumps({
        "impl": "reference", "metric": (METRICS if args.mode == "train" else EVAL_METRICS)[args.model], "value": round(ips, 3), "unit": UNIT,
        "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(spstep * 1e3, 2), "higher_is_better": True,
        "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args), "per_gpu_batch": args.batch, "cpu_sample_batch": args.cpu_sample_batch},
        "cpu_baseline": {"value": round(ips, 3), "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": round(ips, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def workload_name(args):
    if args.mode == "eval":
        return {"resnet18": "resnet_quantized depth=18 ImageNet 224x224 batch 256 eval forward, deterministic rounding (BASELINE configs[1])",
                "mobilenet": "mobilenet_quantized ImageNet 224x224 eval forward", "resnet20_cifar": "resnet_quantized depth=20 CIFAR eval forward"}[args.model]
    return {"resnet18": "resnet_quantized depth=18 ImageNet 224x224 8-bit train step (BASELINE configs[2])",
            "mobilenet": "mobilenet_quantized ImageNet 224x224 8-bit train step (BASELINE configs[3])",
            "resnet20_cifar": "resnet_quantized depth=20 CIFAR 32x32 8-bit train step (BASELINE configs[0])"}[args.model]


# ------------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [s.strip() for s in out.strip().split(",")]
                self.samples.append(float(f[0]))
                self.max_mhz = float(f[1])
                for n, v in zip(names, f[2:]):
                    if v.lower().startswith("active"):
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def main_ours(args, rank, world, local_rank):
    import qtorch_b200  # noqa: F401
    from qtorch_b200 import _abi, quantize as Q
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    _abi.device_info(local_rank)           # raises unless an sm_100 device is there: no fallback
    if world > 1:
        from qtorch_b200 import dist as qdist
        qdist.init(dev)
    torch.manual_seed(0)
    model, shape, ncls = build_model(args.model, Q)
    model = model.to(dev).train()
    if world > 1:
        qdist.broadcast_parameters(model)
    from qtorch_b200 import glue
    opt = glue.SGD(model.parameters(), lr=0.1, momentum=0.9)     # torch.optim.SGD arithmetic, one multi-tensor launch
    crit = glue.CrossEntropyLoss()                               # nn.CrossEntropyLoss arithmetic, one pass each way on libqb200
    if args.scaling == "strong":
        if args.batch % world != 0:
            raise SystemExit("--scaling strong: global batch %d is not divisible by %d GPUs" % (args.batch, world))
        B = args.batch // world            # BASELINE configs[2]: global batch 256 sharded 256/128/64/32
    else:
        B = args.batch
    evalm = args.mode == "eval"
    gen = torch.Generator().manual_seed(1000 + rank)
    # two pinned host batches (alternated) for the end-to-end leg; device-resident copies for the `value` leg
    host_x = [torch.randn(B, *shape, generator=gen).pin_memory() for _ in range(2)]
    host_y = [torch.randint(0, ncls, (B,), generator=gen).pin_memory() for _ in range(2)]
    x_static = host_x[0].to(dev)
    y_static = host_y[0].to(dev)
    loss_static = torch.zeros((), device=dev)

    def eval_body():
        with torch.no_grad():
            loss_static.copy_(crit(model(x_static), y_static))

    def step_body():
        if evalm:
            return eval_body()
        out = model(x_static)
        loss = crit(out, y_static) + model.regularization()
        opt.zero_grad(set_to_none=True)      # torch default: no fill + accumulate kernels per parameter
        loss.backward()
        if world > 1:
            qdist.allreduce_gradients(model)
        opt.step()
        Q.noise_step(dev).add_(1)
        loss_static.copy_(loss.detach())

    # eager warm-up (also allocates every workspace outside any capture)
    side = torch.cuda.Stream(dev)
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        if evalm:                          # configs[1]: "after >= 1 train step to populate running ranges"
            evalm = False
            step_body()
            evalm = True
            model.eval()
        for _ in range(3):
            step_body()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = None
    use_graph = not args.no_graph
    n0 = _abi.LAUNCHES
    if use_graph:
        try:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                step_body()
        except Exception as e:                     # report, fall back to eager launches of the SAME kernels
            sys.stderr.write("CUDA-graph capture failed (%s); running eager\n" % (str(e).splitlines()[0],))
            graph = None
            torch.cuda.synchronize()
    run_step = graph.replay if graph is not None else step_body
    if graph is None:
        n0 = _abi.LAUNCHES
        step_body()
    step_body_launches = _abi.LAUNCHES - n0      # our kernels per step (captured once, replayed every step)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        run_step()
    # ---- device-resident leg (`value`) ----
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        run_step()
    e1.record()
    barrier()
    ms_dev = e0.elapsed_time(e1)
    # ---- end-to-end leg: pinned host batch -> device every step, loss read back every step ----
    copy_stream = torch.cuda.Stream(dev)
    stage_x = [torch.empty_like(x_static) for _ in range(2)]
    stage_y = [torch.empty_like(y_static) for _ in range(2)]
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    barrier()
    t0 = time.perf_counter()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    losses = []
    # Input pipeline of the public API's training loop: the pinned batch of step i+1 crosses PCIe on a copy stream
    # into one of two staging buffers while step i computes; every step's H2D copy (the first one included, fully
    # exposed) and loss read-back are inside the timed region.
    cur = torch.cuda.current_stream(dev)

    def prefetch(i):
        k = i % 2
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[k])                 # the staging buffer's previous batch has been taken
            stage_x[k].copy_(host_x[k], non_blocking=True)
            stage_y[k].copy_(host_y[k], non_blocking=True)
            ready[k].record(copy_stream)

    prefetch(0)
    for i in range(args.steps):
        k = i % 2
        if i + 1 < args.steps:
            prefetch(i + 1)
        cur.wait_event(ready[k])
        x_static.copy_(stage_x[k], non_blocking=True)           # device-to-device into the graph's input tensors
        y_static.copy_(stage_y[k], non_blocking=True)
        consumed[k].record(cur)
        run_step()
        losses.append(float(loss_static))           # D2H read of the step's result (syncs)
    e3.record()
    barrier()
    ms_e2e = e2.elapsed_time(e3)
    wall_e2e = time.perf_counter() - t0
    sampler.stop_flag = True
    sampler.join(timeout=2)
    del copy_stream, stage_x, stage_y
    # max over ranks
    if world > 1:
        t = torch.tensor([ms_dev, ms_e2e], device=dev, dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ms_dev, ms_e2e = t.tolist()
    # ---- roofline of our dominant kernel: one extra, instrumented eager step (events around every C-ABI call) ----
    roof = None
    cpu_base = None
    roof = roofline_pass(step_body, _abi)          # on EVERY rank: the step contains collectives
    if rank == 0:
        if not args.no_cpu_baseline and world == 1:
            ips, _ = cpu_reference_run(args.model, args.cpu_sample_batch, 2, 1, args.mode)
            cpu_base = {"value": round(ips, 3), "unit": UNIT, "cores": host_cores(), "kind": "port",
                        "sample": "%s %s at batch %d x 2 steps after 1 warm-up (oracle port on torch-CPU)" % (
                            args.model, "train step" if args.mode == "train" else "eval forward", args.cpu_sample_batch)}
    if world > 1:
        qdist.check()                      # the peer-memory statistic exchange reported no failure (host watchdog)
        torch.distributed.barrier()
    if rank != 0:
        return
    imgs = B * world * args.steps
    h2d = host_x[0].numel() * 4 + host_y[0].numel() * 8
    line = {
        "metric": (EVAL_METRICS if args.mode == "eval" else METRICS)[args.model], "value": round(imgs / (ms_dev / 1e3), 2), "unit": UNIT,
        "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_dev / args.steps, 3), "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": DTYPE,
        "data": "synthetic",
        "config": {"workload": workload_name(args), "per_gpu_batch": B, "global_batch": B * world,
                   "parallelism": "dp%d" % world, "cuda_graph": graph is not None,
                   "l2": "inputs+activations per step (>2 GB) exceed the 126 MB L2"},
        "e2e": {"value": round(imgs / (ms_e2e / 1e3), 2), "unit": UNIT, "h2d_bytes_per_step": h2d * world,
                "d2h_bytes_per_step": 4 * world, "ms_per_step": round(ms_e2e / args.steps, 3),
                "wall_ms_per_step": round(wall_e2e * 1e3 / args.steps, 3),
                "pipeline": "pinned host batch of step i+1 copied on a second stream into double-buffered staging while "
                            "step i computes; all K copies (the first fully exposed) and K loss reads are inside the timed "
                            "region"},
        "gpu_launches": int(step_body_launches * args.steps),
        "clocks": sampler.summary(), "roofline": roof, "cpu_baseline": cpu_base, "loss_last": losses[-1],
    }
    print(json.dumps(line))


TENSOR_FAMILIES = {"qb_conv_i8_fwd": "i8", "qb_gemm_i8": "i8", "qb_conv_stem_fwd": "i8", "qb_conv_dgrad": "bf16", "qb_conv_dgrad_acc": "bf16", "qb_conv_wgrad": "bf16",
                   "qb_conv_stem_wgrad": "tf32", "qb_conv_stem_wgrad16": "bf16"}


def roofline_pass(step_body, _abi):
    """Times every C-ABI kernel call of ONE eager step with CUDA events on the launching stream and reports the kernel family
    with the largest total time against the roofline that bounds it: HBM for the quantizers / RangeBN /
    glue (algorithmic bytes), the tensor pipe for the contractions (algorithmic flops = 2*M*N*K useful operations; the
    backward kernels execute twice that - hi + lo planes - and the int8 forward runs at twice the bf16 rate, so its peak is
    2 x the measured bf16 figure).  Peaks: MEASURED_PEAKS.json (sustained bf16 figure: these kernels are timed inside a step)."""
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak, src = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
    tpeak, tsrc = (peaks["bf16_tflops_sustained"], "measured (sustained)") if "bf16_tflops_sustained" in peaks else (1400.0, "fallback")
    _abi.profile_begin()
    step_body()
    torch.cuda.synchronize()
    recs = _abi.profile_end()
    fam = {}
    for name, nbytes, ms in recs:
        f = fam.setdefault(name, [0, 0.0, 0])
        f[0] += nbytes
        f[1] += ms
        f[2] += 1
    if not fam:
        return None
    top = max(fam.items(), key=lambda kv: kv[1][1])
    name, (nbytes, ms, calls) = top
    fams = {}
    for k, v in sorted(fam.items(), key=lambda kv: -kv[1][1]):
        fams[k] = {"GBps": round(v[0] / v[1] / 1e6, 1), "frac_hbm": round(v[0] / v[1] / 1e6 / peak, 3), "ms": round(v[1], 3), "calls": v[2]}
        ops = _abi.PROFILE_OPS.get(k)
        if ops and k in TENSOR_FAMILIES:
            kind = TENSOR_FAMILIES[k]
            pk = tpeak * (2.0 if kind == "i8" else (0.5 if kind == "tf32" else 1.0))
            fams[k].update({"TFLOPs": round(ops / v[1] / 1e9, 1), "tensor_kind": kind, "frac_tensor_peak": round(ops / v[1] / 1e9 / pk, 4)})
            if kind == "i8":
                fams[k]["frac_int8_datasheet"] = round(ops / v[1] / 1e9 / INT8_PEAK_TOPS, 4)
    traffic = load_traffic()
    common = {"kernel": name, "launches_per_step": calls, "ms_per_step": round(ms, 4), "traffic": traffic.get(name),
              "how": "CUDA events around each launch of one extra eager step after the timed region; traffic = ncu "
                     "dram__bytes_read.sum + dram__bytes_write.sum per launch, mean over the step's launches of this kernel "
                     "(profiles/traffic_r02.json from the committed ncu capture), null when none is committed",
              "families": fams}
    if name in TENSOR_FAMILIES and _abi.PROFILE_OPS.get(name):
        kind = TENSOR_FAMILIES[name]
        pk = tpeak * (2.0 if kind == "i8" else (0.5 if kind == "tf32" else 1.0))
        achieved = _abi.PROFILE_OPS[name] / ms / 1e9
        return dict(common, bound="tensor", achieved=round(achieved, 1), peak=round(pk, 1), peak_source=tsrc + " bf16 cuBLAS x %s" % (
            {"i8": "2 (int8 rate)", "bf16": "1", "tf32": "0.5 (tf32 rate)"}[kind]), unit="TFLOP/s", frac=round(achieved / pk, 4),
            algorithmic_flops_per_step=_abi.PROFILE_OPS[name])
    achieved = nbytes / ms / 1e6
    return dict(common, bound="hbm", achieved=round(achieved, 1), peak=peak, peak_source=src, unit="GB/s", frac=round(achieved / peak, 4),
                algorithmic_bytes_per_step=nbytes)


def load_traffic():
    """Per-launch DRAM bytes of each kernel family from the committed ncu --set full capture of this build
    (tools/ncu_step_summary.py writes profiles/traffic_r02.json)."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "traffic_r02.json")))
    except Exception:
        return {}


def main():
    args = parse()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        main_reference(args, rank)
        return
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        main_ours(args, rank, world, local_rank)
    finally:
        if world > 1:
            torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
