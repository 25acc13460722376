#!/bin/bash
TAG=${1:-r2f}
OUT=gpurun_out/$TAG
mkdir -p $OUT
B="--no-cpu-baseline --e2e-steps 3 --steps 800 --warmup 100"
for v in u1 u3 u1nohist u3nohist; do
  so=disturbance_crazyfile_simulation_b200/csrc/variants/libcf_$v.so
  for opt in "" "--streams 1" "--no-prefetch" "--no-graph" "--streams 1 --no-prefetch"; do
    tagopt=$(echo "$opt" | tr -d ' -')
    CF_LIB_PATH=$so timeout 200 python bench.py $B $opt > $OUT/b_${v}_${tagopt:-default}.json 2>> $OUT/bench.err
    python -c "
import json
d=json.loads(open('$OUT/b_${v}_${tagopt:-default}.json').read().strip().split('\n')[-1]); print('$v [$opt] %.3e'%d['value'], '%.2f us'%(1e3*d['ms_per_step']), 'per-launch %.2f us'%(1e3*d['roofline']['avg_kernel_ms_events_per_launch']))"
  done
done
