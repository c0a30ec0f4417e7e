"""Summarise an `ncu --set full --page raw --csv` export of one training step per kernel family.

    python tools/ncu_step_summary.py gpurun_out/step.raw.csv [> profiles/ncu_step_rNN_summary.txt]

Per family: launches, total / mean duration, DRAM bytes read + written per launch (the `traffic` figure bench.py's
roofline quotes), DRAM / L1 / issue / occupancy percentages (duration-weighted means).  ncu durations are cold-cache and
serialised: shares, not absolutes, are comparable with the CUDA-event timings of bench.py."""
import csv
import re
import sys
from collections import OrderedDict

UNIT_SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}


def short(name):
    name = re.sub(r"^void ", "", name)
    m = re.match(r"([A-Za-z0-9_]+)(<[^(]*>)?", name)
    base = m.group(1) if m else name
    targs = (m.group(2) or "") if m else ""
    targs = re.sub(r"\(int\)|\(bool\)", "", targs)
    return base + targs


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}

    def val(r, key):
        i = ix.get(key)
        if i is None or r[i] in ("", "n/a"):
            return None
        return float(r[i].replace(",", "")) * UNIT_SCALE.get(units[i], 1.0)

    fam = OrderedDict()
    for r in rows[2:]:
        f = fam.setdefault(short(r[ix["Kernel Name"]]), [])
        f.append(r)
    cols = [("dram%", "dram__throughput.avg.pct_of_peak_sustained_elapsed"),
            ("l1tex%", "l1tex__throughput.avg.pct_of_peak_sustained_active"),
            ("lts%", "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
            ("issue%", "sm__issue_active.avg.pct_of_peak_sustained_elapsed"),
            ("warps%", "sm__warps_active.avg.pct_of_peak_sustained_active"),
            ("alu%", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
            ("fma%", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
            ("tensor%", "sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_active")]
    cols = [(a, b) for a, b in cols if b in ix]
    tot = sum(val(r, "gpu__time_duration.sum") or 0.0 for f in fam.values() for r in f)
    print("# %s: %d kernel launches captured, %.1f us in total (ncu durations: cold cache, serialised)" % (
        sys.argv[1], sum(len(f) for f in fam.values()), tot))
    print("%-52s %4s %9s %8s %6s %12s %12s %5s  %s" % ("kernel", "n", "total_us", "mean_us", "share", "rdMB/launch",
                                                      "wrMB/launch", "regs", " ".join("%7s" % a for a, _ in cols)))
    for name, f in sorted(fam.items(), key=lambda kv: -sum(val(r, "gpu__time_duration.sum") or 0.0 for r in kv[1])):
        t = [val(r, "gpu__time_duration.sum") or 0.0 for r in f]
        rd = sum(val(r, "dram__bytes_read.sum") or 0.0 for r in f) / len(f) / 1e6
        wr = sum(val(r, "dram__bytes_write.sum") or 0.0 for r in f) / len(f) / 1e6
        regs = val(f[0], "launch__registers_per_thread")
        w = []
        for _, key in cols:
            num = sum((val(r, key) or 0.0) * tt for r, tt in zip(f, t))
            w.append(num / max(sum(t), 1e-9))
        print("%-52s %4d %9.1f %8.1f %5.1f%% %12.2f %12.2f %5d  %s" % (name[:52], len(f), sum(t), sum(t) / len(f),
                                                                     100 * sum(t) / tot, rd, wr, int(regs or 0),
                                                                     " ".join("%7.1f" % x for x in w)))


    # --json PATH: DRAM bytes (read + written) per launch of the dominant kernel of each C-ABI entry point, the `traffic`
    # figure of bench.py's roofline (profiles/traffic_rNN.json)
    if "--json" in sys.argv:
        import json
        entry = {"k_conv_wgrad": "qb_conv_wgrad", "k_conv_dgrad_flat": "qb_conv_dgrad", "k_conv_dgrad": "qb_conv_dgrad",
                 "k_conv_i8": "qb_conv_i8_fwd", "k_conv_stem": "qb_conv_stem_fwd", "k_rbn_stats_q8_nhwc": "qb_rangebn_stats_nhwc",
                 "k_rbn_bwd_reduce_nhwc": "qb_rangebn_bwd_reduce_nhwc", "k_rbn_bwd_apply_fused": "qb_rangebn_bwd_apply_fused_nhwc",
                 "k_rbn_apply_codes_nhwc": "qb_rangebn_apply_nhwc", "k_bwd_pack": "qb_bwd_pack", "k_fq_affine_v8": "qb_fakequant_affine16",
                 "k_relu_bwd_minmax": "qb_relu_bwd_minmax_bits", "k_maxpool_bwd_3s2": "qb_maxpool2d_relu_nhwc_bwd",
                 "k_maxpool_fwd_nhwc": "qb_maxpool2d_nhwc_fwd", "k_conv_dw3x3": "qb_conv_dw_fwd", "k_conv_dw_dgrad": "qb_conv_dw_dgrad",
                 "k_conv_dw_wgrad": "qb_conv_dw_wgrad", "k_gemm_i8": "qb_gemm_i8"}
        out, acc = {}, {}
        for name, f in fam.items():
            base = re.sub(r"<.*$", "", name)
            e = entry.get(base)
            if name.startswith("k_measure_stats"):
                e = "qb_rangebn_apply_relu_measure" if ("<3" in name or "<4" in name) else "qb_measure_stats"
            if e is None:
                continue
            b = sum((val(r, "dram__bytes_read.sum") or 0.0) + (val(r, "dram__bytes_write.sum") or 0.0) for r in f)
            a = acc.setdefault(e, [0.0, 0])
            a[0] += b
            a[1] += len(f)
        for e, (b, n) in acc.items():
            out[e] = round(b / n)
        json.dump(out, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
