#!/bin/bash
TAG=${1:-r2s}
OUT=gpurun_out/$TAG
mkdir -p $OUT
B="--no-cpu-baseline --e2e-steps 3 --steps 1600 --warmup 160 --shards 16 --streams 16"
for v in disturbance_crazyfile_simulation_b200/csrc/variants/libcf_*.so; do
  n=$(basename $v .so)
  CF_LIB_PATH=$v timeout 300 python bench.py $B > $OUT/b_$n.json 2>> $OUT/bench.err
  CF_LIB_PATH=$v timeout 300 python bench.py $B --workload initial > $OUT/bi_$n.json 2>> $OUT/bench.err
  python -c "
import json
L=lambda f: json.loads(open(f).read().strip().split('\n')[-1])
c=L('$OUT/b_$n.json'); i=L('$OUT/bi_$n.json')
print('$n: %.3e  %.2f us  frac %.3f | initial %.3e frac %.3f'%(c['value'], 1e3*c['ms_per_step'], c['roofline']['frac'], i['value'], i['roofline']['frac']))"
done
