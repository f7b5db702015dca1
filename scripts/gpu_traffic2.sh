#!/bin/bash
set -u
OUT=gpurun_out/$1; mkdir -p $OUT
CMD="python bench.py --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline"
for pf in 1; do
PCB200_DF_PREFETCH=$pf timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:rhs_update -s 9 -c 3 --csv --log-file $OUT/traffic_pf$pf.csv $CMD > $OUT/traffic_pf$pf.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("$OUT/traffic_pf$pf.csv")) if len(r)>10 and r[0]!="ID"]
for r in rows: print("pf=$pf", r[0], r[-3], r[-2], r[-1])
PY
PCB200_DF_PREFETCH=$pf timeout 120 python bench.py --no-cpu-baseline --steps 5 --e2e-steps 1 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('pf=$pf bench %.3f G/s'%(d['value']/1e9), d['clocks'])"
done
