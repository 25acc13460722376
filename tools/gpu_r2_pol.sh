#!/bin/bash
TAG=${1:-r2l}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 python -m pytest tests/test_policy_kernel.py tests/test_trainer_gpu.py -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/rc.txt
timeout 600 python bench.py --workload ppo --steps 5 --warmup 2 > $OUT/bench_ppo.json 2>> $OUT/bench.err; echo "ppo rc=$?" >> $OUT/rc.txt
cat $OUT/rc.txt; tail -3 $OUT/pytest.log | cut -c1-300; python -c "
import json
d=json.loads(open('$OUT/bench_ppo.json').read().strip().split('\n')[-1]); print('%.3e'%d['value'], json.dumps(d['side']))"
