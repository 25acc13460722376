#!/bin/bash
# quick iteration: GPU parity tests, headline bench (no CPU baseline), per-CTA timeline
TAG=${1:-it}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/rc.txt
timeout 300 python bench.py --no-cpu-baseline --e2e-steps 3 > $OUT/bench.json 2>> $OUT/bench.err
timeout 300 python bench.py --no-cpu-baseline --e2e-steps 3 --envs 524288 --shards 1 --streams 1 --steps 400 --warmup 20 > $OUT/bench_flow.json 2>> $OUT/bench.err
timeout 300 python tools/kprobe.py --streams 1 2 --graph-reps 4 --prefetch --traces 1 > $OUT/kprobe.log 2>&1
cat $OUT/rc.txt; tail -3 $OUT/pytest_gpu.log
for f in $OUT/bench.json $OUT/bench_flow.json; do python -c "
import json
d=json.loads(open('$f').read().strip().split('\n')[-1]); print('$f %.3e'%d['value'], d['ms_per_step'], d['roofline']['frac'])"; done
head -8 $OUT/kprobe.log | cut -c1-900
