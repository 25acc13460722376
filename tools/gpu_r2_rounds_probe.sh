#!/bin/bash
# NOTE: the variant libraries (csrc/variants/libcf_r7.so, libcf_r4.so) are built ad hoc with build.build(extra_flags=['-DCF_PHILOX_ROUNDS=7'],
# out=...) and are not kept; results: profiles/r2_kstep_ab.md section 8.
# timing-only probe (wrong random numbers): how much of a k_step launch is Philox instruction count?  Product library first, then
# builds with 7 and 4 Philox rounds (tools only; -DCF_PHILOX_ROUNDS), same bench command, one box.
mkdir -p gpurun_out
V=disturbance_crazyfile_simulation_b200/csrc/variants
for name in product r7 r4 product; do
  if [ "$name" = product ]; then unset CF_LIB_PATH; else export CF_LIB_PATH=$PWD/$V/libcf_$name.so; fi
  timeout 300 python bench.py --no-cpu-baseline --e2e-steps 3 > gpurun_out/rounds_$name.json 2> gpurun_out/rounds_$name.err
  echo "$name exit $?"
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/rounds_$name.json') if l.startswith('{')][-1])
print('$name', 'value %.4e'%d['value'], 'us/launch %.3f'%(1e3*d['ms_per_step']), 'frac %.4f'%d['roofline']['frac'])
PY
done
