#!/bin/bash
# tiny experiments: scripts/gpu_try.sh <tag> ; runs bench at 128^3 with env variations
OUT=gpurun_out/$1; mkdir -p $OUT
for pf in 0 1; do
  PCB200_DF_PREFETCH=$pf timeout 120 python bench.py --size 512 --kernel-variant ${V:-2} --no-cpu-baseline --steps 5 > $OUT/b_pf$pf.json 2> $OUT/b_pf$pf.err; echo "pf=$pf exit $?"; python -c "import json; d=json.load(open('$OUT/b_pf$pf.json')); print('%.3f G/s'%(d['value']/1e9))"
done
