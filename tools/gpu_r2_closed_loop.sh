#!/bin/bash
# GPU tests + closed-loop report of the reference checkpoints (with the Bullet-model switch sweep)
TAG=${1:-r2b}
OUT=gpurun_out/$TAG
mkdir -p $OUT
export CF_HJ_TABLE_CACHE=/tmp/cf_hj_tables
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/rc.txt
timeout 1500 python tools/closed_loop_report.py $OUT/closed_loop.json --n 8192 --switches > $OUT/closed_loop.log 2>&1; echo "closed loop rc=$?" >> $OUT/rc.txt
cat $OUT/rc.txt; tail -5 $OUT/pytest_gpu.log; grep -v "generating the HJ" $OUT/closed_loop.log | cut -c1-260 | tail -80
