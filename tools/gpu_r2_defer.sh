#!/bin/bash
# deferred (list-driven) resets: parity suite, then A/B against the in-kernel reset on one box
TAG=${1:-r2p}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_rng_bits.py tests/test_trainer_gpu.py tests/test_policy_kernel.py -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/rc.txt
tail -3 $OUT/pytest.log | cut -c1-300
B="--no-cpu-baseline --e2e-steps 3"
for d in 1 0; do
  CF_DEFER_RESET=$d timeout 300 python bench.py $B > $OUT/bench_d$d.json 2>> $OUT/bench.err
  CF_DEFER_RESET=$d timeout 300 python bench.py $B --streams 2 --steps 800 --warmup 100 > $OUT/bench_s2_d$d.json 2>> $OUT/bench.err
  CF_DEFER_RESET=$d timeout 300 python bench.py $B --streams 8 --steps 800 --warmup 100 > $OUT/bench_s8_d$d.json 2>> $OUT/bench.err
  CF_DEFER_RESET=$d timeout 300 python bench.py $B --workload initial > $OUT/bench_initial_d$d.json 2>> $OUT/bench.err
  CF_DEFER_RESET=$d timeout 300 python bench.py $B --actions uniform --steps 800 --warmup 100 > $OUT/bench_uniform_d$d.json 2>> $OUT/bench.err
  python -c "
import json
L=lambda f: json.loads(open(f).read().strip().split('\n')[-1])
d=L('$OUT/bench_d$d.json'); p=L('$OUT/bench_s2_d$d.json'); q=L('$OUT/bench_s8_d$d.json'); i=L('$OUT/bench_initial_d$d.json'); u=L('$OUT/bench_uniform_d$d.json')
print('defer=$d headline %.3e'%d['value'], '%.2f us'%(1e3*d['ms_per_step']), 'frac %.3f'%d['roofline']['frac'], 'per-launch %.2f us'%(1e3*d['roofline']['avg_kernel_ms_events_per_launch']), '| 2 streams %.3e'%p['value'], '| 8 streams %.3e'%q['value'], '| initial %.3e'%i['value'], '| uniform %.3e'%u['value'], '| e2e %.3e'%d['e2e']['value'], '| L2-resident %.3e'%d['side']['one_shard_l2_resident_env_steps_per_s'])"
done
cat $OUT/rc.txt; tail -3 $OUT/bench.err
