#!/bin/bash
# Round-2 validation pass on one GPU: all GPU tests, smoke, both bench arms as the driver runs them, other workloads, launch list and
# full ncu capture of the step kernel and of the policy kernel
TAG=${1:-r2t}
OUT=gpurun_out/$TAG
mkdir -p $OUT
export CF_HJ_TABLE_CACHE=/tmp/cf_hj_tables
timeout 1800 python -m pytest tests -m gpu -x -q -s > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/rc.txt
grep -v "generating the HJ" $OUT/pytest_gpu.log | grep -i "passed\|failed\|error\|worst\|philox 60\|euler conv\|EpLen\|consistent" | tail -30 | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" >> $OUT/rc.txt
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?" >> $OUT/rc.txt
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 > $OUT/bench_driver.json 2> $OUT/bench.err; echo "bench20 rc=$?" >> $OUT/rc.txt
timeout 600 python bench.py > $OUT/bench.json 2>> $OUT/bench.err; echo "bench rc=$?" >> $OUT/rc.txt
timeout 300 python bench.py --no-cpu-baseline --e2e-steps 3 --workload initial > $OUT/bench_initial_g0.json 2>> $OUT/bench.err
timeout 300 python bench.py --no-cpu-baseline --e2e-steps 3 --workload initial --ground-effect 1 > $OUT/bench_initial_g1.json 2>> $OUT/bench.err
timeout 300 python bench.py --actions pd --no-cpu-baseline --e2e-steps 3 > $OUT/bench_pd.json 2>> $OUT/bench.err
timeout 300 python bench.py --actions uniform --no-cpu-baseline --e2e-steps 3 > $OUT/bench_uniform.json 2>> $OUT/bench.err
timeout 600 python bench.py --workload ppo --steps 5 --warmup 2 > $OUT/bench_ppo.json 2>> $OUT/bench.err; echo "ppo rc=$?" >> $OUT/rc.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/ncu_launches.csv \
  python bench.py --steps 40 --warmup 5 --no-cpu-baseline --e2e-steps 3 > $OUT/ncu_launches.log 2>&1; echo "ncu list rc=$?" >> $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step -s 40 -c 3 -o $OUT/k_step_full -f \
  python bench.py --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 3 > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?" >> $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_policy -s 10 -c 2 -o $OUT/k_policy_full -f \
  python bench.py --workload ppo --steps 1 --warmup 1 --envs 65536 > $OUT/ncu_policy.log 2>&1; echo "ncu policy rc=$?" >> $OUT/rc.txt
cat $OUT/rc.txt; tail -1 $OUT/smoke.log
for f in bench_driver bench bench_initial_g0 bench_initial_g1 bench_pd bench_uniform; do python -c "
import json
d=json.loads(open('$OUT/$f.json').read().strip().split('\n')[-1]); r=d['roofline']; print('$f %.3e'%d['value'], '%.2f us'%(1e3*d['ms_per_step']), 'frac %.3f per-launch %.3f issue %.3f'%(r['frac'], r['frac_per_launch'], r['issue_frac']), 'e2e %.3e'%d['e2e']['value'], 'gae', d['side'].get('gae',{}).get('achieved_gb_s'))"; done
cut -c1-1500 $OUT/bench_ppo.json; cut -c1-600 $OUT/bench_ref.json; tail -3 $OUT/bench.err
