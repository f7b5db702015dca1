#!/bin/bash
# ncu --set full of the fused kernel at n^3: scripts/gpu_ncu.sh <tag> <variant> [n]
set -u
TAG=$1; V=$2; N=${3:-256}
OUT=gpurun_out/$TAG; mkdir -p $OUT
CMD="python bench.py --n $N --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline --kernel-variant $V"
$CMD > $OUT/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rhs_update -s 9 -c 3 -o $OUT/prof_v${V}_$N $CMD > $OUT/ncu_full.log 2>&1
tail -3 $OUT/ncu_full.log; ls -la $OUT
