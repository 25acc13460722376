#!/bin/bash
mkdir -p gpurun_out
CF_PROBE_E2E_TIMELINE=$PWD/gpurun_out/e2e_timeline_cpu.csv timeout 300 python bench.py --no-cpu-baseline --steps 20 --warmup 3 > gpurun_out/e2e_tl_cpu.json 2> gpurun_out/e2e_tl_cpu.err
echo exit $?
python - <<'PY'
import numpy as np
a=np.loadtxt('gpurun_out/e2e_timeline_cpu.csv', delimiter=',', skiprows=1)
b=a[90:104]; g0=b[0,1]; c0=b[0,6]
for r in b:
    print('gpu: start %7.1f kern_done %7.1f obs_done %7.1f all_done %7.1f | cpu: enq_in %7.1f enq_out %7.1f sync_in %7.1f sync_out %7.1f' % (*(1e3*(r[[1,3,4,5]]-g0)), *(1e3*(r[6:10]-c0))))
PY
