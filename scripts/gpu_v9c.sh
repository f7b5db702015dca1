#!/bin/bash
# v9 A/B: parity subset, bench at 512^3 with and without the df L2 prefetch, optional ncu: scripts/gpu_v9c.sh <tag> [ncu]
set -u
TAG=${1:-v9c}
OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q --timeout 120 -k "one_substep or tile_edges or three_substeps" > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log
tail -4 $OUT/pytest.log
for pf in 1 0; do
  PCB200_DF_PREFETCH=$pf timeout 200 python bench.py --kernel-variant 9 --no-cpu-baseline --steps 5 > $OUT/bench_pf$pf.json 2> $OUT/bench_pf$pf.err; echo "bench pf=$pf exit $?"
  python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_pf$pf.json")); print("pf=$pf", "%.3f G/s"%(d["value"]/1e9), "frac %.3f"%d["roofline"]["frac"], "kernel ms %.3f"%d["roofline"]["avg_kernel_ms"], d["clocks"])
except Exception as e: print("failed", e); print(open("$OUT/bench_pf$pf.err").read()[-2000:])
PY
done
if [ "${2:-}" = "ncu" ]; then
CMD="python bench.py --n 256 --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline --kernel-variant 9"
$CMD > $OUT/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rhs_update -s 9 -c 3 -o $OUT/prof $CMD > $OUT/ncu_full.log 2>&1
tail -2 $OUT/ncu_full.log
fi
