#!/bin/bash
# One gpurun call: GPU tests, both bench arms, ncu launch list + full capture of the step kernel, variant probes.
# Everything it writes goes to gpurun_out/<tag>/.
TAG=${1:-s}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $OUT/rc.txt
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" >> $OUT/rc.txt
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?" >> $OUT/rc.txt
for s in 1 3 4; do
  timeout 300 python bench.py --streams $s --no-cpu-baseline --e2e-steps 3 > $OUT/bench_streams$s.json 2>> $OUT/bench.err
done
for v in disturbance_crazyfile_simulation_b200/csrc/variants/libcf_*.so; do
  [ -f "$v" ] || continue
  n=$(basename $v .so)
  CF_LIB_PATH=$v timeout 300 python bench.py --no-cpu-baseline --e2e-steps 3 > $OUT/bench_$n.json 2>> $OUT/bench.err
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/ncu_launches.csv \
  python bench.py --steps 40 --warmup 5 --no-cpu-baseline --e2e-steps 3 > $OUT/ncu_launches.log 2>&1; echo "ncu list rc=$?" >> $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step -s 40 -c 3 -o $OUT/k_step_full -f \
  python bench.py --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 3 > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?" >> $OUT/rc.txt
cat $OUT/rc.txt
tail -3 $OUT/pytest_gpu.log
cat $OUT/bench.json | cut -c1-400
