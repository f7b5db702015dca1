#!/bin/bash
set -u
OUT=gpurun_out/r2b; mkdir -p $OUT
timeout 120 scripts/ubench/ubench > $OUT/ubench.txt 2>&1; echo "ubench exit $?"
cat $OUT/ubench.txt
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 600 > $OUT/pytest.log 2>&1; echo "pytest exit $?" >> $OUT/pytest.log
tail -5 $OUT/pytest.log
