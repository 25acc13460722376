#!/bin/bash
TAG=${1:-r2h}
OUT=gpurun_out/$TAG
mkdir -p $OUT
B="--no-cpu-baseline --e2e-steps 3 --steps 1600 --warmup 160"
for opt in "--streams 2" "--streams 3" "--streams 4" "--streams 3 --no-prefetch" "--streams 4 --no-prefetch" "--streams 4 --shards 12" "--streams 3 --shards 9"; do
  tagopt=$(echo "$opt" | tr -d ' -')
  timeout 200 python bench.py $B $opt > $OUT/b_${tagopt}.json 2>> $OUT/bench.err
  python -c "
import json
d=json.loads(open('$OUT/b_${tagopt}.json').read().strip().split('\n')[-1]); print('[$opt] %.3e'%d['value'], '%.2f us'%(1e3*d['ms_per_step']), 'frac %.3f'%d['roofline']['frac'])"
done
