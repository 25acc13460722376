#!/bin/bash
# NOTE: the variant libraries under csrc/variants/ are built ad hoc with build.build(extra_flags=['-DCF_GAE_U=..', '-DCF_GAE_MIN_CTAS=..',
# '-DCF_GAE_THREADS=..'], out=...) and are not kept; results: profiles/r2_gae_sweep.log.
# GAE kernel: (loads in flight U, min CTAs/SM, threads/CTA) variants, T = 32 x N = 524 288, one box
mkdir -p gpurun_out; : > gpurun_out/gae_sweep.log
V=disturbance_crazyfile_simulation_b200/csrc/variants
for f in $V/libcf_*.so; do
  n=$(basename $f .so)
  echo -n "$n  " >> gpurun_out/gae_sweep.log
  CF_LIB_PATH=$PWD/$f PYTHONPATH=. timeout 120 python tools/gae_probe.py >> gpurun_out/gae_sweep.log 2>&1 || echo "failed" >> gpurun_out/gae_sweep.log
done
echo -n "product  " >> gpurun_out/gae_sweep.log
PYTHONPATH=. timeout 120 python tools/gae_probe.py >> gpurun_out/gae_sweep.log 2>&1
echo -n "product N=2097152  " >> gpurun_out/gae_sweep.log
PYTHONPATH=. timeout 120 python tools/gae_probe.py 2097152 >> gpurun_out/gae_sweep.log 2>&1
cat gpurun_out/gae_sweep.log
