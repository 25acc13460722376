#!/bin/bash
# Round-end style pass on one GPU: GPU tests, both bench arms, launch list, full ncu capture of the step kernel.
TAG=${1:-final}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/rc.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" >> $OUT/rc.txt
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?" >> $OUT/rc.txt
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" >> $OUT/rc.txt
timeout 300 python bench.py --actions pd --no-cpu-baseline --e2e-steps 3 > $OUT/bench_pd.json 2>> $OUT/bench.err
timeout 300 python bench.py --actions uniform --no-cpu-baseline --e2e-steps 3 > $OUT/bench_uniform.json 2>> $OUT/bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/ncu_launches.csv \
  python bench.py --steps 40 --warmup 5 --no-cpu-baseline --e2e-steps 3 > $OUT/ncu_launches.log 2>&1; echo "ncu list rc=$?" >> $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step -s 40 -c 3 -o $OUT/k_step_full -f \
  python bench.py --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 3 > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?" >> $OUT/rc.txt
cat $OUT/rc.txt; tail -2 $OUT/pytest_gpu.log; tail -1 $OUT/smoke.log; cut -c1-300 $OUT/bench.json
