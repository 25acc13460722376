#!/bin/bash
# A/B of the experimental builds under csrc/variants/ on one box (headline bench, second workload, per-CTA timeline), then
# a full ncu capture of the product build
TAG=${1:-r2e}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for v in disturbance_crazyfile_simulation_b200/csrc/variants/libcf_*.so; do
  [ -f "$v" ] || continue
  n=$(basename $v .so)
  CF_LIB_PATH=$v timeout 300 python bench.py --no-cpu-baseline --e2e-steps 3 > $OUT/bench_$n.json 2>> $OUT/bench.err
  CF_LIB_PATH=$v timeout 300 python bench.py --no-cpu-baseline --e2e-steps 3 --workload initial > $OUT/bench_initial_$n.json 2>> $OUT/bench.err
  CF_LIB_PATH=$v timeout 300 python tools/kprobe.py --streams 1 2 --graph-reps 4 --prefetch --traces 1 > $OUT/kprobe_$n.log 2>&1
  python -c "
import json
d=json.loads(open('$OUT/bench_$n.json').read().strip().split('\n')[-1]); i=json.loads(open('$OUT/bench_initial_$n.json').read().strip().split('\n')[-1]); print('$n headline %.3e'%d['value'], d['ms_per_step'], 'frac %.3f'%d['roofline']['frac'], 'per-launch %.4f ms'%d['roofline']['avg_kernel_ms_events_per_launch'], '| initial %.3e'%i['value'])"
  sed -n 4,5p $OUT/kprobe_$n.log | cut -c1-420
done
if [ -n "$NCU_VARIANT" ]; then
  CF_LIB_PATH=disturbance_crazyfile_simulation_b200/csrc/variants/libcf_$NCU_VARIANT.so timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step -s 40 -c 2 -o $OUT/k_step_$NCU_VARIANT -f \
    python bench.py --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 3 > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?" >> $OUT/rc.txt
fi
cat $OUT/rc.txt 2>/dev/null
