#!/bin/bash
# NOTE: csrc/variants/libcf_t64.so is built ad hoc with build.build(extra_flags=['-DCF_TILE=64'], out=...) and is not kept.
# k_step with 64-env tiles (64-thread CTAs, twice as many): parity tests first, then the headline against the product library
mkdir -p gpurun_out
V=$PWD/disturbance_crazyfile_simulation_b200/csrc/variants/libcf_t64.so
CF_LIB_PATH=$V timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -3
for lib in product t64 product t64; do
  if [ "$lib" = product ]; then unset CF_LIB_PATH; else export CF_LIB_PATH=$V; fi
  timeout 300 python bench.py --no-cpu-baseline --e2e-steps 24 > gpurun_out/t64_$lib.json 2> gpurun_out/t64_$lib.err
  python -c "
import json
d=json.loads([l for l in open('gpurun_out/t64_$lib.json') if l.startswith('{')][-1])
print('$lib', 'value %.4e'%d['value'], 'us/launch %.3f'%(1e3*d['ms_per_step']), 'frac %.4f'%d['roofline']['frac'])" 2>&1 | tail -1
done
