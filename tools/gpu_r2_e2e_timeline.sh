#!/bin/bash
# NOTE: the 'skip' leg used CF_PROBE_SKIP_SMALL_D2H, a probe-build switch that is not in the tree; CF_PROBE_E2E_TIMELINE is.
# e2e timeline (CUDA events per host-pointer step): product, then with the small downloads skipped
mkdir -p gpurun_out
for mode in product skip; do
  if [ "$mode" = skip ]; then export CF_PROBE_SKIP_SMALL_D2H=1; else unset CF_PROBE_SKIP_SMALL_D2H; fi
  CF_PROBE_E2E_TIMELINE=$PWD/gpurun_out/e2e_timeline_$mode.csv timeout 300 python bench.py --no-cpu-baseline --steps 20 --warmup 3 \
    > gpurun_out/e2e_tl_$mode.json 2> gpurun_out/e2e_tl_$mode.err
  python -c "
import json
d=json.loads([l for l in open('gpurun_out/e2e_tl_$mode.json') if l.startswith('{')][-1])
print('$mode', 'e2e %.4e'%d['e2e']['value'], 'ms/step %.4f'%(65536/d['e2e']['value']*1e3))"
  python tools/e2e_timeline_report.py gpurun_out/e2e_timeline_$mode.csv
done
