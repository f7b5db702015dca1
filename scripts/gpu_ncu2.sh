#!/bin/bash
# ncu --set full of the three substep launches of the default kernel (lab harness, 256^3): scripts/gpu_ncu2.sh <tag> [n]
set -u
TAG=${1:-prof}; N=${2:-256}
OUT=gpurun_out/$TAG; mkdir -p $OUT
python scripts/abl.py $N 1 ${3:-0} > $OUT/plain.log 2>&1 || { cat $OUT/plain.log; exit 1; }
cat $OUT/plain.log
ncu --set full --clock-control none --import-source on -k regex:rhs_update -s 6 -c 3 -o $OUT/prof python scripts/abl.py $N 1 ${3:-0} ${3:-0} > $OUT/ncu_full.log 2>&1
tail -2 $OUT/ncu_full.log
