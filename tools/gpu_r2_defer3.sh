#!/bin/bash
TAG=${1:-r2r}
OUT=gpurun_out/$TAG
mkdir -p $OUT
B="--no-cpu-baseline --e2e-steps 3 --steps 1600 --warmup 160"
for cfgs in "8 4" "8 8" "12 12" "16 8" "16 16" "24 24"; do
  set -- $cfgs
  for d in 1 0; do
    CF_DEFER_RESET=$d timeout 300 python bench.py $B --shards $1 --streams $2 > $OUT/b_$1_$2_d$d.json 2>> $OUT/bench.err
    python -c "
import json
c=json.loads(open('$OUT/b_$1_$2_d$d.json').read().strip().split('\n')[-1])
print('shards $1 streams $2 defer=$d: %.3e  %.2f us  frac %.3f  per-launch %.1f us'%(c['value'], 1e3*c['ms_per_step'], c['roofline']['frac'], 1e3*c['roofline']['avg_kernel_ms_events_per_launch']))"
  done
done
CF_DEFER_RESET=1 timeout 300 python bench.py --no-cpu-baseline --e2e-steps 3 --envs 524288 --shards 1 --streams 1 --steps 200 --warmup 20 > $OUT/flow_d1.json 2>> $OUT/bench.err
python -c "
import json
c=json.loads(open('$OUT/flow_d1.json').read().strip().split('\n')[-1]); print('flow defer=1 %.3e %.1f us'%(c['value'],1e3*c['ms_per_step']))"
