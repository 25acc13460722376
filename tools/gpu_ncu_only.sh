#!/bin/bash
# launch list + full capture of the step kernel for the bench command (plain run first)
OUT=gpurun_out/${1:-ncu6}; mkdir -p $OUT
timeout 300 python bench.py --steps 40 --warmup 5 --no-cpu-baseline --e2e-steps 3 > $OUT/plain.json 2> $OUT/plain.err || { echo plain failed; tail -5 $OUT/plain.err; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/ncu_launches.csv \
  python bench.py --steps 40 --warmup 5 --no-cpu-baseline --e2e-steps 3 > $OUT/ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step -s 40 -c 3 -o $OUT/k_step_full -f \
  python bench.py --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 3 > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i $OUT/k_step_full.ncu-rep --page raw --csv > $OUT/k_step_full_raw.csv 2>/dev/null; wc -l $OUT/ncu_launches.csv
