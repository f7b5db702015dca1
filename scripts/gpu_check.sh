#!/bin/bash
# One gpurun call: GPU parity tests, default bench, ncu launch list, ncu --set full of the fused kernel.
# usage: scripts/gpu_check.sh <tag>
set -u
TAG=${1:-run}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
which gfortran mpif90 mpirun > $OUT/fortran_probe.txt 2>&1; nproc >> $OUT/fortran_probe.txt
timeout 300 python -m pytest tests -m gpu -x -q --timeout 60 > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log
tail -3 $OUT/pytest_gpu.log
python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $OUT/smoke.log
python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; cat $OUT/bench.json
python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2>> $OUT/bench.err; cat $OUT/bench_ref.json
CMD="python bench.py --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline"
$CMD > $OUT/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv $CMD > $OUT/ncu_launches.log 2>&1
CMD2="python bench.py --n 256 --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline"
$CMD2 > $OUT/plain256.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rhs_update -s 9 -c 3 -o $OUT/prof256 $CMD2 > $OUT/ncu_full.log 2>&1
ls -la $OUT
