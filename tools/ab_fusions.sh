for i in 1 2; do
for cfg in "" "QB_FUSE_ROWSUM=0" "QB_FUSE_RBN_FINALIZE=0" "QB_FUSE_RBN_COEFFS=0" "QB_FUSE_ROWSUM=0 QB_FUSE_RBN_FINALIZE=0 QB_FUSE_RBN_COEFFS=0"; do
  v=$(env $cfg python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'])")
  echo "round $i [$cfg] $v" >> gpurun_out/ab_r02q.txt
done; done
cat gpurun_out/ab_r02q.txt
