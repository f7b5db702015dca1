#!/bin/bash
# variant 8 check: parity tests of the kernel variants, bench at 512^3, with "ncu" as 2nd argument ncu --set full at 256^3
# usage: scripts/gpu_v8.sh <tag> [ncu]
set -u
OUT=gpurun_out/${1:-v8}; mkdir -p $OUT
timeout 200 python -m pytest tests/test_parity_gpu.py -m gpu -x -q --timeout 60 -k "one_substep or tile_edges or loopback" > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log
tail -3 $OUT/pytest_gpu.log
timeout 120 python bench.py --kernel-variant 8 --no-cpu-baseline --steps 5 > $OUT/bench_v8.json 2> $OUT/bench_v8.err; echo "bench exit $?"
cut -c1-300 $OUT/bench_v8.json
if [ "${2:-}" = ncu ]; then timeout 200 bash scripts/gpu_ncu.sh ${1:-v8} 8 256; fi
