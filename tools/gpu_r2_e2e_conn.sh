#!/bin/bash
# e2e: are the host-pointer streams falsely serialised on the default 8 hardware connections?  Same box: default, 32, default.
mkdir -p gpurun_out
for c in default 32 default; do
  if [ "$c" = default ]; then unset CUDA_DEVICE_MAX_CONNECTIONS; else export CUDA_DEVICE_MAX_CONNECTIONS=$c; fi
  timeout 300 python bench.py --no-cpu-baseline --steps 200 --warmup 20 > gpurun_out/e2e_conn_$c.json 2> gpurun_out/e2e_conn_$c.err
  python -c "
import json
d=json.loads([l for l in open('gpurun_out/e2e_conn_$c.json') if l.startswith('{')][-1])
print('connections $c', 'e2e %.4e'%d['e2e']['value'], 'ms/step %.4f'%(65536/d['e2e']['value']*1e3), 'value %.4e'%d['value'], 'frac %.4f'%d['roofline']['frac'])"
done
export CUDA_DEVICE_MAX_CONNECTIONS=32
CF_PROBE_E2E_TIMELINE=$PWD/gpurun_out/e2e_timeline_conn32.csv timeout 300 python bench.py --no-cpu-baseline --steps 20 --warmup 3 > /dev/null 2>&1
python tools/e2e_timeline_report.py gpurun_out/e2e_timeline_conn32.csv
