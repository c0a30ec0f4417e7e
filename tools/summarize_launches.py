"""Summarise an ncu launch list (`--metrics gpu__time_duration.sum --csv`) per kernel name:
    python tools/summarize_launches.py gpurun_out/launches.csv [header text] > profiles/launches_rNN_summary.txt
ncu times are cold-cache and serialised: compare SHARES, not absolutes."""
import csv
import re
import sys
from collections import defaultdict


def short(name):
    name = re.sub(r"\(.*$", "", name)           # drop the argument list
    return name[:84]


def main():
    path = sys.argv[1]
    header = sys.argv[2] if len(sys.argv) > 2 else ""
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        ns = v * {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1.0)
        rows.append((short(r["Kernel Name"]), ns))
    tot = sum(ns for _, ns in rows)
    agg = defaultdict(lambda: [0, 0.0])
    for k, ns in rows:
        agg[k][0] += 1
        agg[k][1] += ns
    if header:
        print(header)
    print("total %.3f ms over %d launches" % (tot / 1e6, len(rows)))
    ours = sum(ns for k, ns in rows if re.search(r"\bk_[a-z]", k))
    print("ours (k_*) %.3f ms = %.1f%%" % (ours / 1e6, 100 * ours / tot))
    for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-84s %5d %8.3f ms %5.1f%%" % (k, n, ns / 1e6, 100 * ns / tot))


if __name__ == "__main__":
    main()
