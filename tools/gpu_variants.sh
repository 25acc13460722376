#!/bin/bash
# bench + per-CTA timeline for every experimental build under csrc/variants/
TAG=${1:-var}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for v in disturbance_crazyfile_simulation_b200/csrc/variants/libcf_*.so; do
  [ -f "$v" ] || continue
  n=$(basename $v .so)
  CF_LIB_PATH=$v timeout 300 python bench.py --no-cpu-baseline --e2e-steps 3 > $OUT/bench_$n.json 2>> $OUT/bench.err
  CF_LIB_PATH=$v timeout 300 python tools/kprobe.py --streams 1 2 --graph-reps 4 --prefetch --traces 1 > $OUT/kprobe_$n.log 2>&1
  python -c "
import json
d=json.loads(open('$OUT/bench_$n.json').read().strip().split('\n')[-1]); print('$n %.3e'%d['value'], d['ms_per_step'], d['roofline']['frac'])"
  head -5 $OUT/kprobe_$n.log | cut -c1-700
done
