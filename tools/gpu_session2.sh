#!/bin/bash
TAG=${1:-s2}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -x -q -k "pd_actions" > $OUT/pytest_pd.log 2>&1; echo "pytest rc=$?" >> $OUT/rc.txt
timeout 300 python bench.py --actions pd --no-cpu-baseline --e2e-steps 3 > $OUT/bench_pd.json 2>> $OUT/bench.err
timeout 300 python bench.py --actions uniform --no-cpu-baseline --e2e-steps 3 > $OUT/bench_uniform.json 2>> $OUT/bench.err
timeout 300 python tools/kprobe.py --streams 1 2 --graph-reps 4 --prefetch > $OUT/kprobe.log 2>&1
timeout 900 ncu --set full --cache-control none --clock-control none --import-source on -k regex:k_step -s 40 -c 3 -o $OUT/k_step_warm -f \
  python bench.py --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 3 > $OUT/ncu_warm.log 2>&1; echo "ncu warm rc=$?" >> $OUT/rc.txt
cat $OUT/rc.txt; tail -3 $OUT/pytest_pd.log
