#!/bin/bash
# variant 9 bring-up: parity subset first, then A/B bench against variant 5 (every step under its own timeout)
# usage: scripts/gpu_v9.sh <tag> [full]
set -u
TAG=${1:-v9}
OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q --timeout 120 -k "one_substep or tile_edges or three_substeps" > $OUT/pytest_v9.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_v9.log
tail -15 $OUT/pytest_v9.log
for v in 9 5; do
  timeout 200 python bench.py --kernel-variant $v --no-cpu-baseline --steps 5 > $OUT/bench_v$v.json 2> $OUT/bench_v$v.err; echo "bench v$v exit $?"
  python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_v$v.json")); print("v$v", "%.3f G/s"%(d["value"]/1e9), "frac %.3f"%d["roofline"]["frac"], "kernel ms %.3f"%d["roofline"]["avg_kernel_ms"], d["clocks"])
except Exception as e: print("v$v failed", e); print(open("$OUT/bench_v$v.err").read()[-2000:])
PY
done
if [ "${2:-}" = "full" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q --timeout 300 > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log
  tail -5 $OUT/pytest_gpu.log
fi
