#!/bin/bash
# final check of a round: all GPU tests, smoke(), the default bench line and the reference arm
# usage: scripts/gpu_final.sh <tag>
set -u
OUT=gpurun_out/${1:-final}; mkdir -p $OUT
timeout 300 python -m pytest tests -m gpu -x -q --timeout 60 > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log
tail -3 $OUT/pytest_gpu.log
timeout 120 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $OUT/smoke.log
timeout 200 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; cut -c1-600 $OUT/bench.json
