#!/bin/bash
OUT=gpurun_out/${1:-ab}
mkdir -p $OUT
run() { # name, lib
  for w in "hover_hj" "initial"; do
    for rep in 1 2; do
      CF_LIB_PATH=$2 timeout 300 python bench.py --workload $w --no-cpu-baseline --e2e-steps 3 > $OUT/$1_${w}_$rep.json 2>> $OUT/err.txt
      python -c "
import json
d=json.loads([l for l in open('$OUT/$1_${w}_$rep.json').read().strip().split('\n') if l.startswith('{')][-1]); print('$1 $w rep$rep %.4e'%d['value'])"
    done
  done
}
run product disturbance_crazyfile_simulation_b200/csrc/libcf_b200.so
run regg disturbance_crazyfile_simulation_b200/csrc/variants/libcf_regg.so
run legacy disturbance_crazyfile_simulation_b200/csrc/variants/libcf_legacy.so
run product2 disturbance_crazyfile_simulation_b200/csrc/libcf_b200.so
