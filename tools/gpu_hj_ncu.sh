#!/bin/bash
TAG=${1:-hjn}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 ncu --set full --clock-control none --cache-control none --import-source on -k regex:"k_update|k_minmax|k_choose" -s 18 -c 6 -o $OUT/hj_full -f python tools/hj_run.py > $OUT/ncu.log 2>&1; echo "ncu rc=$?" >> $OUT/rc.txt
cat $OUT/rc.txt; tail -3 $OUT/ncu.log
