#!/bin/bash
# 8-GPU box: decomposition invariance, strong scaling 512^3 (N = 1, 8, slabs), config[3] 1024^3 yz-pencils
set -u
OUT=gpurun_out/multi8; mkdir -p $OUT
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 tests/multi_gpu_check.py > $OUT/check.log 2>&1; echo "check exit $?"
grep -E "case|MULTI_GPU" $OUT/check.log | tail -12
timeout 300 python bench.py --no-cpu-baseline --steps 10 --e2e-steps 1 > $OUT/bench_1.json 2> $OUT/bench_1.err; echo "bench 1 exit $?"
for N in 4 8; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$N bench.py --gpus $N --no-cpu-baseline --steps 10 --e2e-steps 1 > $OUT/bench_$N.json 2> $OUT/bench_$N.err; echo "bench $N exit $?"
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 8 --size 1024 --decomp pencil --no-cpu-baseline --steps 10 --e2e-steps 1 > $OUT/bench_pencil_1024.json 2> $OUT/bench_pencil_1024.err; echo "bench pencil exit $?"
python - <<PY
import json
for n in ("1", "4", "8", "pencil_1024"):
    try:
        d=json.loads(open("$OUT/bench_%s.json"%n).read().strip().splitlines()[-1]); print(n, "%.3f G/s"%(d["value"]/1e9), "frac %.3f"%d["roofline"]["frac"], "share %.3f"%d["roofline"]["kernel_share_of_step"], d["clocks"], d.get("checksum"))
    except Exception as e: print("failed", n, e); print(open("$OUT/bench_%s.err"%n).read()[-1500:])
PY
