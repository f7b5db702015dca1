#!/bin/bash
# N-GPU box: decomposition-invariance check (p2p and nccl cases), then the 512^3 bench with both z-halo transports
# usage: scripts/gpu_p2p.sh <tag> <ngpus>
set -u
TAG=$1; N=$2
OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/multi_gpu_check.py > $OUT/multi_check_$N.log 2>&1
echo "multi_gpu_check exit $?"; grep -E "case|MULTI_GPU|rror" $OUT/multi_check_$N.log | head -20
for h in p2p nccl; do
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 5 --warmup 3 --e2e-steps 1 --halo $h > $OUT/bench_${h}_$N.json 2> $OUT/bench_${h}_$N.err
  echo "bench $h exit $?"
  python - $OUT/bench_${h}_$N.json "$h N=$N" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    print(sys.argv[2], "%.3f G/s" % (d["value"] / 1e9), d["config"]["decomposition"], d["config"]["z_halo"], "kernel share %.3f" % d["roofline"]["kernel_share_of_step"], "frac %.3f" % d["roofline"]["frac"])
except Exception as e:
    print(sys.argv[2], "failed", e); print(open(sys.argv[1].replace(".json", ".err")).read()[-1500:])
PY
done
