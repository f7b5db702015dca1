#!/bin/bash
# multi-GPU check + bench: scripts/gpu_multi.sh <tag> <ngpus> [bench n]
set -u
TAG=$1; N=$2; NN=${3:-512}
OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi --query-gpu=index,name --format=csv > $OUT/smi.txt 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/multi_gpu_check.py > $OUT/multi_check.log 2>&1
echo "multi_gpu_check exit $?"; grep -E "case|MULTI_GPU" $OUT/multi_check.log; tail -5 $OUT/multi_check.log | cut -c1-300
for d in slab pencil; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --size $NN --steps 5 --warmup 3 --decomp $d > $OUT/bench_${d}_$N.json 2> $OUT/bench_${d}_$N.err
echo "bench $d exit $?"; python - <<PY
import json
try:
    d=json.loads([l for l in open("$OUT/bench_${d}_$N.json") if l.startswith("{")][-1]); print("$d N=$N", "%.3f G/s"%(d["value"]/1e9), d["config"]["decomposition"], "kernel share %.3f"%d["roofline"]["kernel_share_of_step"], "frac %.3f"%d["roofline"]["frac"])
except Exception as e: print("failed", e); print(open("$OUT/bench_${d}_$N.err").read()[-1500:])
PY
done
timeout 300 python bench.py --gpus 1 --size $NN --steps 5 --warmup 3 --no-cpu-baseline > $OUT/bench_1.json 2> $OUT/bench_1.err; python -c "
import json; d=json.load(open('$OUT/bench_1.json')); print('N=1 %.3f G/s'%(d['value']/1e9))"
