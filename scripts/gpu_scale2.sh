#!/bin/bash
# strong-scaling points of the 512^3 bench on one multi-GPU box with the p2p z-halo: scripts/gpu_scale2.sh <tag> "<Ns>" [pencil1024]
set -u
TAG=$1; NS=$2; P=${3:-0}
OUT=gpurun_out/$TAG; mkdir -p $OUT
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    print(sys.argv[2], "%.3f G/s" % (d["value"] / 1e9), d["config"]["decomposition"], d["config"]["z_halo"], "kernel share %.3f" % d["roofline"]["kernel_share_of_step"], "frac %.3f" % d["roofline"]["frac"])
except Exception as e:
    print(sys.argv[2], "failed", e); print(open(sys.argv[1].replace(".json", ".err")).read()[-1200:])
PY
}
for n in $NS; do
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 5 --warmup 3 --e2e-steps 1 > $OUT/scale_p2p_$n.json 2> $OUT/scale_p2p_$n.err
  echo "slab N=$n exit $?"; show $OUT/scale_p2p_$n.json "slab 512^3 N=$n"
done
if [ "$P" = "1" ]; then
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --size 1024 --decomp pencil --steps 3 --warmup 3 --e2e-steps 1 > $OUT/pencil_1024_8.json 2> $OUT/pencil_1024_8.err
  echo "pencil 1024 exit $?"; show $OUT/pencil_1024_8.json "pencil 1024^3 N=8"
fi
