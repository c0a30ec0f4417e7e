"""profiles/traffic_rNN.json (the `traffic` figure of bench.py's roofline) from an ncu metrics list of one eager training step:

    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file X.csv \\
        python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline
    python tools/traffic_from_metrics.py X.csv profiles/traffic_r02.json > profiles/step_dram_r02_summary.txt

Per C-ABI entry point: mean DRAM bytes (read + written) per launch of its dominant kernel; the text summary lists every kernel
with launches, time share and read / written MB per launch."""
import csv
import json
import re
import sys
from collections import OrderedDict, defaultdict

ENTRY = {"k_conv_wgrad": "qb_conv_wgrad", "k_conv_dgrad_flat": "qb_conv_dgrad", "k_conv_dgrad": "qb_conv_dgrad", "k_conv_i8": "qb_conv_i8_fwd",
         "k_conv_stem": "qb_conv_stem_fwd", "k_rbn_stats_q8_nhwc": "qb_rangebn_stats_nhwc", "k_rbn_bwd_reduce_nhwc": "qb_rangebn_bwd_reduce_nhwc",
         "k_rbn_bwd_apply_fused": "qb_rangebn_bwd_apply_fused_nhwc", "k_rbn_apply_codes_nhwc": "qb_rangebn_apply_nhwc", "k_bwd_pack": "qb_bwd_pack",
         "k_fq_affine_v8": "qb_fakequant_affine16", "k_relu_bwd_minmax": "qb_relu_bwd_minmax_bits", "k_maxpool_bwd_3s2": "qb_maxpool2d_relu_nhwc_bwd",
         "k_maxpool_fwd_nhwc": "qb_maxpool2d_nhwc_fwd", "k_conv_dw3x3": "qb_conv_dw_fwd", "k_conv_dw_dgrad": "qb_conv_dw_dgrad",
         "k_conv_dw_wgrad": "qb_conv_dw_wgrad", "k_gemm_i8": "qb_gemm_i8", "k_measure_stats": "qb_rangebn_apply_relu_measure", "k_fq_affine_rowsum": "qb_fakequant_affine_rowsum",
         "k_fq_affine_nchw3": "qb_fakequant_affine_nchw3", "k_avgpool_bwd_nhwc": "qb_avgpool2d_nhwc_bwd"}
ALIAS = {"qb_rangebn_bwd_reduce_nhwc": "qb_rangebn_bwd_reduce_coeffs_nhwc"}      # same kernel behind the one-call entry point
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}


def main():
    with open(sys.argv[1], newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    per = OrderedDict()                        # launch id -> {name, metrics}
    for r in csv.DictReader(lines):
        d = per.setdefault(r["ID"], {"name": r["Kernel Name"]})
        d[r["Metric Name"]] = float(r["Metric Value"].replace(",", "")) * SCALE.get(r["Metric Unit"], 1.0)
    fam = defaultdict(list)
    for d in per.values():
        name = re.sub(r"^void ", "", re.sub(r"\(.*$", "", d["name"]))
        fam[name].append(d)
    tot = sum(d.get("gpu__time_duration.sum", 0.0) for v in fam.values() for d in v)
    print("# %s: %d launches, %.1f us (ncu: cold cache, serialised)" % (sys.argv[1], len(per), tot))
    print("%-60s %4s %9s %6s %12s %12s" % ("kernel", "n", "total_us", "share", "rdMB/launch", "wrMB/launch"))
    out, acc = {}, defaultdict(lambda: [0.0, 0])
    for name, v in sorted(fam.items(), key=lambda kv: -sum(d.get("gpu__time_duration.sum", 0.0) for d in kv[1])):
        t = sum(d.get("gpu__time_duration.sum", 0.0) for d in v)
        rd = sum(d.get("dram__bytes_read.sum", 0.0) for d in v) / len(v)
        wr = sum(d.get("dram__bytes_write.sum", 0.0) for d in v) / len(v)
        print("%-60s %4d %9.1f %5.1f%% %12.2f %12.2f" % (name[:60], len(v), t, 100 * t / max(tot, 1e-9), rd / 1e6, wr / 1e6))
        e = ENTRY.get(re.sub(r"<.*$", "", name))
        if e:
            acc[e][0] += (rd + wr) * len(v)
            acc[e][1] += len(v)
    for e, (b, n) in acc.items():
        out[e] = round(b / n)
        if e in ALIAS:
            out[ALIAS[e]] = out[e]
    json.dump(out, open(sys.argv[2], "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
