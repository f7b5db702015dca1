#!/bin/bash
# one multi-GPU box: decomposition-invariance check at N ranks, then the strong-scaling series of the 512^3 bench
# (z-slabs) and the 1024^3 yz-pencil case.  usage: scripts/gpu_scale.sh <tag> <ngpus>
set -u
TAG=$1; N=$2
OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi --query-gpu=index,name --format=csv > $OUT/smi.txt 2>&1
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/multi_gpu_check.py > $OUT/multi_check_$N.log 2>&1
echo "multi_gpu_check exit $?"; grep -E "case|MULTI_GPU" $OUT/multi_check_$N.log
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    print(sys.argv[2], "%.3f G/s" % (d["value"] / 1e9), d["config"]["decomposition"], "kernel share %.3f" % d["roofline"]["kernel_share_of_step"], "frac %.3f" % d["roofline"]["frac"], d["clocks"])
except Exception as e:
    print(sys.argv[2], "failed", e)
PY
}
for n in 1 2 4 8; do
  [ $n -gt $N ] && continue
  if [ $n -eq 1 ]; then
    timeout 150 python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $OUT/scale_slab_$n.json 2> $OUT/scale_slab_$n.err
  else
    timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 5 --warmup 3 --e2e-steps 1 > $OUT/scale_slab_$n.json 2> $OUT/scale_slab_$n.err
  fi
  echo "slab N=$n exit $?"; show $OUT/scale_slab_$n.json "slab N=$n"
done
if [ $N -ge 8 ]; then
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --size 1024 --decomp pencil --steps 3 --warmup 3 --e2e-steps 1 > $OUT/pencil_1024_8.json 2> $OUT/pencil_1024_8.err
  echo "pencil 1024 exit $?"; show $OUT/pencil_1024_8.json "pencil 1024^3 N=8"
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --size 1024 --decomp slab --steps 3 --warmup 3 --e2e-steps 1 > $OUT/slab_1024_8.json 2> $OUT/slab_1024_8.err
  echo "slab 1024 exit $?"; show $OUT/slab_1024_8.json "slab 1024^3 N=8"
fi
