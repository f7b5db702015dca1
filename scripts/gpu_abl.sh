#!/bin/bash
# runs scripts/abl.py for every library under build/abl: scripts/gpu_abl.sh [n] [steps] [variant]
OUT=gpurun_out/abl; mkdir -p $OUT
for so in $(ls build/abl/libpcb200_*.so | sort -V); do
  PCB200_LIB=$so timeout 120 python scripts/abl.py ${1:-512} ${2:-3} ${3:-0} 2>&1 | tail -1 | tee -a $OUT/abl.txt
done
