#!/bin/bash
# sustained bench A/B of kernel variants: scripts/gpu_bench2.sh "9 10" [steps]
set -u
OUT=gpurun_out/bench2; mkdir -p $OUT
for kv in $1; do
  timeout 300 python bench.py --kernel-variant $kv --no-cpu-baseline --steps ${2:-10} --e2e-steps 1 > $OUT/bench_v$kv.json 2> $OUT/bench_v$kv.err; echo "bench v$kv exit $?"
  python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_v$kv.json")); print("v$kv", "%.3f G/s"%(d["value"]/1e9), "frac %.3f"%d["roofline"]["frac"], "kernel ms %.3f"%d["roofline"]["avg_kernel_ms"], d["clocks"], d.get("checksum"))
except Exception as e: print("failed", e); print(open("$OUT/bench_v$kv.err").read()[-2000:])
PY
done
