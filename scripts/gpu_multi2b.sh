#!/bin/bash
# N-GPU: decomposition invariance + bench with the in-kernel halo wait (default) and with split launches
set -u
N=${1:-2}
OUT=gpurun_out/multi${N}b; mkdir -p $OUT
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 tests/multi_gpu_check.py > $OUT/check.log 2>&1; echo "check exit $?"
grep -E "case|MULTI_GPU" $OUT/check.log | tail -12
for mode in single split; do
  if [ $mode = split ]; then export PCB200_SPLIT_LAUNCH=1; else unset PCB200_SPLIT_LAUNCH; fi
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N --no-cpu-baseline --steps 10 --e2e-steps 1 > $OUT/bench_${N}_$mode.json 2> $OUT/bench_${N}_$mode.err; echo "bench $N $mode exit $?"
done
unset PCB200_SPLIT_LAUNCH
python - <<PY
import json
for m in ("single", "split"):
    try:
        d=json.loads(open("$OUT/bench_${N}_%s.json"%m).read().strip().splitlines()[-1]); print(m, "%.3f G/s"%(d["value"]/1e9), "frac %.3f"%d["roofline"]["frac"], "share %.3f"%d["roofline"]["kernel_share_of_step"], d["clocks"], d.get("checksum"))
    except Exception as e: print("failed", m, e); print(open("$OUT/bench_${N}_%s.err"%m).read()[-1500:])
PY
