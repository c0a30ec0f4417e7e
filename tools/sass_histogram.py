"""SASS opcode histogram of the tensor-core / TMA instructions per kernel of libqb200.so (cuobjdump -sass):
    python tools/sass_histogram.py > profiles/sass_histogram_rNN.txt
UTCIMMA / UTCHMMA = tcgen05.mma kind::i8 / kind::f16, UTCBAR = tcgen05.commit, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG /
UTMAREDG = TMA tensor load / store / reduce, SYNCS = mbarrier operations, UTCATOMSWS = tcgen05.alloc."""
import collections
import os
import re
import subprocess
import sys

so = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "quantized.pytorch_b200", "libqb200.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
OPS = ["UTCIMMA", "UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTCATOMSWS", "UTMALDG", "UTMASTG", "UTMAREDG", "UTMACCTL", "SYNCS",
       "ELECT", "IDP", "HMMA", "IMMA", "LDG", "STG", "ATOMG", "REDG", "RED"]
fn, hist, total = None, collections.OrderedDict(), {}
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        fn = re.sub(r"\(.*$", "", fn)
        hist[fn] = collections.Counter()
        total[fn] = 0
        continue
    m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if fn and m:
        total[fn] += 1
        op = m.group(1)
        if op in OPS:
            hist[fn][op] += 1
print("# cuobjdump -sass %s: instruction counts per kernel (static)" % os.path.basename(so))
print("%-58s %6s  %s" % ("kernel", "instr", "tensor-core / TMA / memory opcodes"))
for fn, h in hist.items():
    if not (fn.startswith("k_") or fn.startswith("void k_")):
        continue
    print("%-58s %6d  %s" % (fn[:58], total[fn], " ".join("%s:%d" % (k, h[k]) for k in OPS if h[k])))
