#!/bin/bash
# A/B of the launch-level fusions on ONE box: bench.py with each switch off in turn (device-timed and end-to-end images/s).
out=${1:-gpurun_out/ab_fusions.txt}
for i in 1 2; do
for cfg in "" "QB_PREP_ASIDE=1" "QB_FUSE_ROWSUM=0" "QB_FUSE_RBN_FINALIZE=1" "QB_FUSE_RBN_COEFFS=0" "QB_STEM_NCHW=0"; do
  v=$(env $cfg python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'])")
  echo "round $i [$cfg] $v" >> $out
done; done
cat $out
