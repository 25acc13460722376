#!/bin/bash
TAG=${1:-var}
OUT=gpurun_out/$TAG
mkdir -p $OUT
bash tools/gpu_variants.sh $TAG
for v in disturbance_crazyfile_simulation_b200/csrc/variants/libcf_*.so; do
  n=$(basename $v .so)
  CF_LIB_PATH=$v timeout 600 python -m pytest tests -m gpu -x -q > $OUT/pytest_$n.log 2>&1; tail -2 $OUT/pytest_$n.log
done
timeout 600 python -m pytest tests -m gpu -x -q -k full_size > $OUT/pytest_fullsize.log 2>&1; tail -2 $OUT/pytest_fullsize.log
