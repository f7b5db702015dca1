#!/bin/bash
# parity subset for the default kernel + lab timings of every build under build/abl
set -u
OUT=gpurun_out/r2c; mkdir -p $OUT
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q --timeout 300 -k "one_substep or tile_edges or three_substeps or 100_steps or loopback or forcing_30" > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log
tail -4 $OUT/pytest.log
bash scripts/gpu_abl.sh 512 3
