#!/bin/bash
# warp-specialised step kernel: parity suite with CF_STEP_WS=1, then A/B of the headline and other workloads
TAG=${1:-r2k}
OUT=gpurun_out/$TAG
mkdir -p $OUT
CF_STEP_WS=1 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_rng_bits.py -m gpu -x -q > $OUT/pytest_ws.log 2>&1; echo "pytest ws rc=$?" >> $OUT/rc.txt
tail -3 $OUT/pytest_ws.log | cut -c1-300
B="--no-cpu-baseline --e2e-steps 3"
for ws in 0 1; do
  CF_STEP_WS=$ws timeout 300 python bench.py $B > $OUT/bench_ws$ws.json 2>> $OUT/bench.err
  CF_STEP_WS=$ws timeout 300 python bench.py $B --streams 2 --steps 800 --warmup 100 > $OUT/bench_s2_ws$ws.json 2>> $OUT/bench.err
  CF_STEP_WS=$ws timeout 300 python bench.py $B --workload initial > $OUT/bench_initial_ws$ws.json 2>> $OUT/bench.err
  CF_STEP_WS=$ws timeout 300 python tools/kprobe.py --streams 1 2 --graph-reps 4 --traces 1 > $OUT/kprobe_ws$ws.log 2>&1
  python -c "
import json
L=lambda f: json.loads(open(f).read().strip().split('\n')[-1])
d=L('$OUT/bench_ws$ws.json'); p=L('$OUT/bench_s2_ws$ws.json'); i=L('$OUT/bench_initial_ws$ws.json')
print('ws=$ws headline %.3e'%d['value'], '%.2f us'%(1e3*d['ms_per_step']), 'frac %.3f'%d['roofline']['frac'], 'per-launch %.2f us'%(1e3*d['roofline']['avg_kernel_ms_events_per_launch']), '| 2 streams %.3e'%p['value'], '| initial %.3e'%i['value'], '| k32 %.3e'%d['side']['fused_k32_env_steps_per_s'], '| L2-resident %.3e'%d['side']['one_shard_l2_resident_env_steps_per_s'])"
  sed -n 1,5p $OUT/kprobe_ws$ws.log | cut -c1-420
done
cat $OUT/rc.txt; tail -3 $OUT/bench.err
