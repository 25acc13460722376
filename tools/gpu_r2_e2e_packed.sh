#!/bin/bash
# e2e: one adjacent output block (one download per step) against four separate arrays, depth 8, one box; then the host-API tests
mkdir -p gpurun_out; : > gpurun_out/e2e_packed.log
for rep in 1 2 3; do
for mode in separate packed; do
  if [ "$mode" = separate ]; then flag=--e2e-separate; else flag=; fi
  timeout 200 python bench.py --no-cpu-baseline --steps 20 --warmup 3 $flag > gpurun_out/e2e_p.json 2> gpurun_out/e2e_p.err
  python -c "
import json
d=json.loads([l for l in open('gpurun_out/e2e_p.json') if l.startswith('{')][-1])
print('rep $rep $mode', 'e2e %.4e'%d['e2e']['value'], 'ms/step %.4f'%(65536/d['e2e']['value']*1e3))" >> gpurun_out/e2e_packed.log 2>&1 || { echo "rep $rep $mode FAILED" >> gpurun_out/e2e_packed.log; tail -3 gpurun_out/e2e_p.err >> gpurun_out/e2e_packed.log; }
done; done
cat gpurun_out/e2e_packed.log
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "host" 2>&1 | tail -3
