#!/bin/bash
# BASELINE.json configs[1] and configs[4] on one GPU (not the headline; recorded under profiles/)
set -u
OUT=gpurun_out/$1; mkdir -p $OUT
timeout 120 python bench.py --workload iso_mhd --size 128 --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 2 > $OUT/bench_iso_mhd_128.json 2> $OUT/bench_iso_mhd_128.err; echo "iso_mhd exit $?"
timeout 120 python bench.py --workload hydro_hyper3 --size 256 --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 2 > $OUT/bench_hydro_hyper3_256.json 2> $OUT/bench_hydro_hyper3_256.err; echo "hydro_hyper3 exit $?"
timeout 120 python bench.py --size 256 --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 2 > $OUT/bench_mhd_entropy_256.json 2> $OUT/bench_mhd_entropy_256.err; echo "mhd 256 exit $?"
python - $OUT <<'PY'
import json, sys, glob
for f in sorted(glob.glob(sys.argv[1] + "/bench_*.json")):
    try:
        d = json.load(open(f)); print(f.split("/")[-1], "%.3f G/s" % (d["value"] / 1e9), "frac %.3f" % d["roofline"]["frac"], "kernel share %.3f" % d["roofline"]["kernel_share_of_step"], d["config"]["workload"][:50])
    except Exception as e:
        print(f, "failed", e); print(open(f.replace(".json", ".err")).read()[-800:])
PY
