#!/bin/bash
TAG=${1:-r2q}
OUT=gpurun_out/$TAG
mkdir -p $OUT
B="--no-cpu-baseline --e2e-steps 3"
for d in 1 0; do
  CF_DEFER_RESET=$d timeout 300 python bench.py $B --envs 524288 --shards 1 --streams 1 --steps 200 --warmup 20 > $OUT/flow_d$d.json 2>> $OUT/bench.err
  CF_DEFER_RESET=$d timeout 300 python bench.py $B --envs 262144 --shards 4 --streams 4 --steps 200 --warmup 20 > $OUT/flow4_d$d.json 2>> $OUT/bench.err
  CF_DEFER_RESET=$d timeout 300 python bench.py $B --shards 16 --streams 16 --steps 1600 --warmup 160 > $OUT/s16_d$d.json 2>> $OUT/bench.err
  CF_DEFER_RESET=$d timeout 300 python tools/kprobe.py --streams 1 2 --graph-reps 4 --traces 1 > $OUT/kprobe_d$d.log 2>&1
  python -c "
import json
L=lambda f: json.loads(open(f).read().strip().split('\n')[-1])
a=L('$OUT/flow_d$d.json'); b=L('$OUT/flow4_d$d.json'); c=L('$OUT/s16_d$d.json')
print('defer=$d flow 524288x1: %.3e %.1f us'%(a['value'],1e3*a['ms_per_step']), '| 262144 x 4 shards x 4 streams: %.3e frac %.3f'%(b['value'], b['roofline']['frac']), '| 16 shards x 16 streams: %.3e frac %.3f'%(c['value'], c['roofline']['frac']))"
  sed -n 1,5p $OUT/kprobe_d$d.log | cut -c1-420
done
