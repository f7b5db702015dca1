#!/bin/bash
# N-GPU correctness (decomposition invariance vs the oracle) and strong-scaling bench lines: scripts/gpu_multi2.sh <N>
set -u
N=${1:-2}
OUT=gpurun_out/multi$N; mkdir -p $OUT
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 tests/multi_gpu_check.py > $OUT/check.log 2>&1; echo "check exit $?"
grep -E "case|MULTI_GPU" $OUT/check.log | tail -12
timeout 300 python bench.py --no-cpu-baseline --steps 10 --e2e-steps 1 > $OUT/bench_1.json 2> $OUT/bench_1.err; echo "bench 1 exit $?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N --no-cpu-baseline --steps 10 --e2e-steps 1 > $OUT/bench_$N.json 2> $OUT/bench_$N.err; echo "bench $N exit $?"
python - <<PY
import json
for n in (1, $N):
    try:
        d=json.loads(open("$OUT/bench_%d.json"%n).read().strip().splitlines()[-1]); print(n, "%.3f G/s"%(d["value"]/1e9), "frac %.3f"%d["roofline"]["frac"], d["clocks"], d.get("checksum"))
    except Exception as e: print("failed", n, e); print(open("$OUT/bench_%d.err"%n).read()[-1500:])
PY
