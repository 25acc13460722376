#!/bin/bash
# Round-2 check: GPU tests, smoke, both bench arms as the driver runs them (--steps 20 --warmup 3)
TAG=${1:-r2a}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/rc.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" >> $OUT/rc.txt
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?" >> $OUT/rc.txt
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 > $OUT/bench_driver.json 2> $OUT/bench.err; echo "bench20 rc=$?" >> $OUT/rc.txt
timeout 600 python bench.py > $OUT/bench.json 2>> $OUT/bench.err; echo "bench rc=$?" >> $OUT/rc.txt
cat $OUT/rc.txt; tail -3 $OUT/pytest_gpu.log; tail -1 $OUT/smoke.log; cut -c1-600 $OUT/bench_driver.json; cut -c1-400 $OUT/bench.json; tail -5 $OUT/bench.err
