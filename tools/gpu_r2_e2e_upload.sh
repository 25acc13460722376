#!/bin/bash
# NOTE: CF_HOST_UPLOAD existed only in the experiment build this script measured (profiles/r2_e2e_timeline.md); not in the tree.
# e2e: action upload by kernel (default now) against the DMA upload, same box; then the host-pointer parity tests
mkdir -p gpurun_out
for mode in dma kernel dma kernel; do
  if [ "$mode" = dma ]; then export CF_HOST_UPLOAD=dma; else unset CF_HOST_UPLOAD; fi
  timeout 300 python bench.py --no-cpu-baseline --steps 200 --warmup 20 > gpurun_out/e2e_up_$mode.json 2> gpurun_out/e2e_up_$mode.err
  python -c "
import json
d=json.loads([l for l in open('gpurun_out/e2e_up_$mode.json') if l.startswith('{')][-1])
print('upload $mode', 'e2e %.4e'%d['e2e']['value'], 'ms/step %.4f'%(65536/d['e2e']['value']*1e3), 'value %.4e'%d['value'])"
done
unset CF_HOST_UPLOAD
CF_PROBE_E2E_TIMELINE=$PWD/gpurun_out/e2e_timeline_upk.csv timeout 300 python bench.py --no-cpu-baseline --steps 20 --warmup 3 > /dev/null 2>&1
python tools/e2e_timeline_report.py gpurun_out/e2e_timeline_upk.csv
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "host" > gpurun_out/host_tests.log 2>&1; tail -3 gpurun_out/host_tests.log
