#!/bin/bash
# e2e: steps in flight (buffer sets) 2..8, DMA upload and kernel upload, one box
mkdir -p gpurun_out; : > gpurun_out/e2e_depth.log
for rep in 1 2; do
for mode in dma kernel; do
  if [ "$mode" = dma ]; then export CF_HOST_UPLOAD=dma; else unset CF_HOST_UPLOAD; fi
  for d in 2 3 4 6 8; do
    timeout 200 python bench.py --no-cpu-baseline --steps 20 --warmup 3 --e2e-depth $d --e2e-steps 400 > gpurun_out/e2e_d.json 2> gpurun_out/e2e_d.err
    python -c "
import json
d=json.loads([l for l in open('gpurun_out/e2e_d.json') if l.startswith('{')][-1])
print('rep $rep upload $mode depth $d', 'e2e %.4e'%d['e2e']['value'], 'ms/step %.4f'%(65536/d['e2e']['value']*1e3))" >> gpurun_out/e2e_depth.log
  done
done
done
cat gpurun_out/e2e_depth.log
