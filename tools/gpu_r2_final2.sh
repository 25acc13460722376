#!/bin/bash
# round 2, last validation pass: whole GPU suite, smoke, headline bench (driver form and default), reference arm, PPO bench
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1
echo "suite exit $?" | tee -a gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1
echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_driver.json 2> gpurun_out/bench_driver.err
echo "bench driver-form exit $?"
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench exit $?"
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
echo "reference arm exit $?"
timeout 400 python bench.py --workload ppo --steps 3 --warmup 2 > gpurun_out/bench_ppo.json 2> gpurun_out/bench_ppo.err
echo "ppo exit $?"
python - <<'PY'
import json
for f in ('bench_driver','bench','bench_ref','bench_ppo'):
    try:
        d=json.loads([l for l in open(f'gpurun_out/{f}.json') if l.startswith('{')][-1])
        print(f, 'value %.4e'%d['value'], 'frac', (d.get('roofline') or {}).get('frac'), 'e2e %.4e'%(d.get('e2e') or {}).get('value', float('nan')), 'launches', d.get('gpu_launches'), 'clocks', d.get('clocks'))
        if f=='bench': print('   gae', d['side']['gae']['frac_of_hbm_peak'], 'cpu', d['cpu_baseline']['value'], d['cpu_baseline']['cores'])
    except Exception as e: print(f, 'ERR', e)
PY
