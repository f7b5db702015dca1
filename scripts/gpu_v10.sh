#!/bin/bash
# variant 10 bring-up: parity subset (all variants), then lab timing of variants 9 and 10 with the in-tree library
set -u
OUT=gpurun_out/v10; mkdir -p $OUT
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q --timeout 300 -k "one_substep or tile_edges or three_substeps or loopback" > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log
tail -15 $OUT/pytest.log
for kv in 9 10; do timeout 120 python scripts/abl.py ${1:-512} 3 $kv 2>&1 | tail -2 | tee -a $OUT/abl.txt; done
