#!/bin/bash
# round-2 state check: full GPU test suite, default bench, A/B of variants 5 and 9 at 512^3
set -u
OUT=gpurun_out/r2a; mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 600 > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log
tail -5 $OUT/pytest.log
for kv in 9 5; do
  timeout 300 python bench.py --kernel-variant $kv --no-cpu-baseline --steps 10 > $OUT/bench_v$kv.json 2> $OUT/bench_v$kv.err; echo "bench v$kv exit $?"
  python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_v$kv.json")); print("v$kv", "%.3f G/s"%(d["value"]/1e9), "frac %.3f"%d["roofline"]["frac"], "kernel ms %.3f"%d["roofline"]["avg_kernel_ms"], d["clocks"], d.get("checksum"))
except Exception as e: print("failed", e); print(open("$OUT/bench_v$kv.err").read()[-2000:])
PY
done
