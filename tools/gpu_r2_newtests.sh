#!/bin/bash
TAG=${1:-r2c}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1500 python -m pytest tests/test_rng_bits.py tests/test_closed_loop_reference_runs.py -m gpu -q -s > $OUT/pytest_new.log 2>&1; echo "pytest rc=$?" >> $OUT/rc.txt
cat $OUT/rc.txt; grep -v "generating the HJ" $OUT/pytest_new.log | tail -60 | cut -c1-400
