#!/bin/bash
TAG=${1:-s4}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 ncu --set full --cache-control none --clock-control none --import-source on -k regex:k_step -s 12 -c 2 -o $OUT/k_step_flow -f \
  python bench.py --envs 524288 --shards 1 --streams 1 --steps 16 --warmup 5 --no-graph --no-cpu-baseline --e2e-steps 3 > $OUT/ncu_flow.log 2>&1; echo "ncu flow rc=$?" >> $OUT/rc.txt
cat $OUT/rc.txt
