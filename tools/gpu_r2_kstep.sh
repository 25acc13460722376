#!/bin/bash
# k_step iteration: GPU tests, headline bench (driver form + long form), per-CTA timeline
TAG=${1:-r2d}
OUT=gpurun_out/$TAG
mkdir -p $OUT
export CF_HJ_TABLE_CACHE=/tmp/cf_hj_tables
timeout 1500 python -m pytest tests -m gpu -x -q -s ${PYTEST_ARGS} > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/rc.txt
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline > $OUT/bench_driver.json 2>> $OUT/bench.err; echo "bench20 rc=$?" >> $OUT/rc.txt
timeout 300 python bench.py --no-cpu-baseline > $OUT/bench.json 2>> $OUT/bench.err; echo "bench rc=$?" >> $OUT/rc.txt
timeout 300 python bench.py --no-cpu-baseline --e2e-steps 3 --workload initial > $OUT/bench_initial.json 2>> $OUT/bench.err
timeout 300 python tools/kprobe.py --streams 1 2 --graph-reps 4 --prefetch --traces 1 > $OUT/kprobe.log 2>&1
cat $OUT/rc.txt; grep -v "generating the HJ" $OUT/pytest_gpu.log | grep -i "passed\|failed\|error\|worst\|philox 60\|euler conv" | tail -12 | cut -c1-300
for f in $OUT/bench_driver.json $OUT/bench.json $OUT/bench_initial.json; do python -c "
import json
d=json.loads(open('$f').read().strip().split('\n')[-1]); r=d['roofline']; print('$f %.3e'%d['value'], d['ms_per_step'], 'frac %.3f per-launch %.3f issue %.3f'%(r['frac'], r['frac_per_launch'], r['issue_frac']), 'e2e %.3e'%d['e2e']['value'])"; done
head -8 $OUT/kprobe.log | cut -c1-900
