#!/bin/bash
TAG=${1:-s3}
OUT=gpurun_out/$TAG
mkdir -p $OUT
B="timeout 300 python bench.py --no-cpu-baseline --e2e-steps 3"
$B --envs 524288 --shards 1 --streams 1 --steps 400 --warmup 20 > $OUT/bench_e512k_s1.json 2>> $OUT/bench.err
$B --envs 262144 --shards 2 --steps 800 --warmup 40 > $OUT/bench_e256k_s2.json 2>> $OUT/bench.err
$B --envs 131072 --shards 4 --steps 1200 --warmup 40 > $OUT/bench_e128k_s4.json 2>> $OUT/bench.err
$B --envs 75776 --shards 8 > $OUT/bench_e74k_s8.json 2>> $OUT/bench.err
CF_LIB_PATH=disturbance_crazyfile_simulation_b200/csrc/variants/libcf_tile64.so $B > $OUT/bench_tile64.json 2>> $OUT/bench.err
CF_LIB_PATH=disturbance_crazyfile_simulation_b200/csrc/variants/libcf_tile64.so $B --envs 524288 --shards 1 --streams 1 --steps 400 --warmup 20 > $OUT/bench_tile64_e512k.json 2>> $OUT/bench.err
for f in $OUT/bench_*.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().split('\n')[-1]); print('%.3e'%d['value'], d['ms_per_step'], d['roofline']['frac'])"; done
