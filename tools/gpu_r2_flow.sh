#!/bin/bash
# the step kernel at full residency: one launch of 524 288 envs (5.5 waves at 5 CTAs/SM), bench + full ncu capture
TAG=${1:-r2o}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 300 python bench.py --no-cpu-baseline --e2e-steps 3 --envs 524288 --shards 1 --streams 1 --steps 200 --warmup 20 > $OUT/bench_flow.json 2> $OUT/bench.err
timeout 300 python bench.py --no-cpu-baseline --e2e-steps 3 --envs 262144 --shards 2 --streams 2 --steps 200 --warmup 20 > $OUT/bench_flow2.json 2>> $OUT/bench.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_step -s 12 -c 1 -o $OUT/k_step_flow -f \
  python bench.py --envs 524288 --shards 1 --streams 1 --steps 16 --warmup 5 --no-graph --no-cpu-baseline --e2e-steps 3 > $OUT/ncu_flow.log 2>&1; echo "ncu flow rc=$?" >> $OUT/rc.txt
cat $OUT/rc.txt
for f in bench_flow bench_flow2; do python -c "
import json
d=json.loads(open('$OUT/$f.json').read().strip().split('\n')[-1]); print('$f %.3e'%d['value'], '%.2f us'%(1e3*d['ms_per_step']), 'frac %.3f'%d['roofline']['frac'])"; done
