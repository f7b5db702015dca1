#!/bin/bash
# DRAM traffic of the fused kernel at the bench size (512^3): 3 launches (substeps 1-3) -> profiles/traffic.json is
# derived from this CSV by scripts/make_traffic_json.py.  usage: scripts/gpu_traffic.sh <tag>
set -u
OUT=gpurun_out/$1; mkdir -p $OUT
CMD="python bench.py --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline"
timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:rhs_update -s 9 -c 3 --csv --log-file $OUT/traffic_512.csv $CMD > $OUT/traffic.log 2>&1
tail -12 $OUT/traffic_512.csv
