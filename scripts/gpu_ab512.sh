#!/bin/bash
# per-substep kernel times (ncu, no clock control) of two kernel variants at 512^3: scripts/gpu_ab512.sh <tag> "<variants>"
set -u
OUT=gpurun_out/$1; mkdir -p $OUT
for v in $2; do
CMD="python bench.py --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline --kernel-variant $v"
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:rhs_update -s 9 -c 3 --csv --log-file $OUT/t512_v$v.csv $CMD > $OUT/t512_v$v.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("$OUT/t512_v$v.csv")) if len(r)>10 and r[0]!="ID" and r[-3]=="gpu__time_duration.sum"]
print("v$v", [round(float(r[-1].replace(",",""))/1e6,3) for r in rows], "ms")
PY
done
