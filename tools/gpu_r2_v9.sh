#!/bin/bash
TAG=${1:-r2g}
OUT=gpurun_out/$TAG
mkdir -p $OUT
export CF_HJ_TABLE_CACHE=/tmp/cf_hj_tables
timeout 1500 python -m pytest tests -m gpu -x -q -s --deselect tests/test_closed_loop_reference_runs.py > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/rc.txt
grep -v "generating the HJ" $OUT/pytest_gpu.log | grep -i "passed\|failed\|error\|worst\|philox 60\|euler conv\|EpLen" | tail -14 | cut -c1-300
B="--no-cpu-baseline --e2e-steps 3"
for v in disturbance_crazyfile_simulation_b200/csrc/variants/libcf_*.so; do
  n=$(basename $v .so)
  CF_LIB_PATH=$v timeout 300 python bench.py $B > $OUT/bench_$n.json 2>> $OUT/bench.err
  CF_LIB_PATH=$v timeout 300 python bench.py $B --no-prefetch --steps 800 --warmup 100 > $OUT/bench_nopf_$n.json 2>> $OUT/bench.err
  CF_LIB_PATH=$v timeout 300 python bench.py $B --workload initial > $OUT/bench_initial_$n.json 2>> $OUT/bench.err
  CF_LIB_PATH=$v timeout 300 python tools/kprobe.py --streams 1 2 --graph-reps 4 --prefetch --traces 1 > $OUT/kprobe_$n.log 2>&1
  python -c "
import json
L=lambda f: json.loads(open(f).read().strip().split('\n')[-1])
d=L('$OUT/bench_$n.json'); p=L('$OUT/bench_nopf_$n.json'); i=L('$OUT/bench_initial_$n.json')
print('$n headline %.3e'%d['value'], '%.2f us'%(1e3*d['ms_per_step']), 'frac %.3f'%d['roofline']['frac'], 'per-launch %.2f us'%(1e3*d['roofline']['avg_kernel_ms_events_per_launch']), '| no-prefetch %.3e'%p['value'], '| initial %.3e'%i['value'], '| k32 %.3e'%d['side']['fused_k32_env_steps_per_s'])"
  sed -n 4,5p $OUT/kprobe_$n.log | cut -c1-420
done
cat $OUT/rc.txt
