#!/bin/bash
# quick GPU iteration: parity tests + bench for the given kernel variants (every step under its own timeout)
# usage: scripts/gpu_quick.sh <tag> [variants...]
set -u
TAG=${1:-q}; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 300 python -m pytest tests -m gpu -x -q --timeout 60 > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log
tail -5 $OUT/pytest_gpu.log
for v in "$@"; do
  timeout 120 python bench.py --kernel-variant $v --no-cpu-baseline --steps 5 > $OUT/bench_v$v.json 2> $OUT/bench_v$v.err; echo "bench v$v exit $?"
  python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_v$v.json")); print("v$v", "%.3f G/s"%(d["value"]/1e9), "frac %.3f"%d["roofline"]["frac"], "kernel ms %.3f"%d["roofline"]["avg_kernel_ms"], d["clocks"])
except Exception as e: print("v$v failed", e); print(open("$OUT/bench_v$v.err").read()[-2000:])
PY
done
