#!/bin/bash
# NOTE: CF_PROBE_SKIP_SMALL_D2H existed only in the probe build this script measured (profiles/r2_e2e_timeline.md); not in the tree.
# e2e: how much of a step's 0.21 ms is the fixed cost of the small D2H copies (rew, done, cost)?  Same box, product then probe then product.
mkdir -p gpurun_out
for mode in product skip product; do
  if [ "$mode" = skip ]; then export CF_PROBE_SKIP_SMALL_D2H=1; else unset CF_PROBE_SKIP_SMALL_D2H; fi
  timeout 300 python bench.py --no-cpu-baseline --steps 200 --warmup 20 > gpurun_out/e2e_$mode.json 2> gpurun_out/e2e_$mode.err
  python -c "
import json
d=json.loads([l for l in open('gpurun_out/e2e_$mode.json') if l.startswith('{')][-1])
print('$mode', 'e2e %.4e'%d['e2e']['value'], 'ms/step %.4f'%(65536/d['e2e']['value']*1e3))"
done
nvidia-smi --query-gpu=pcie.link.gen.current,pcie.link.width.current --format=csv,noheader
