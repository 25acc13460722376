#!/bin/bash
# round 2: allocation-free entry point - its test alone in a fresh process, then the whole GPU suite (it runs last), smoke, bench lines
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/noalloc_alone.log 2>&1
echo "alone exit $?" | tee -a gpurun_out/noalloc_alone.log; tail -5 gpurun_out/noalloc_alone.log
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1
echo "suite exit $?" | tee -a gpurun_out/pytest_gpu.log; tail -6 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1
echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench exit $?"; cut -c1-400 gpurun_out/bench.json
timeout 400 python bench.py --workload ppo --steps 3 --warmup 2 > gpurun_out/bench_ppo.json 2> gpurun_out/bench_ppo.err
echo "ppo exit $?"; cut -c1-200 gpurun_out/bench_ppo.json
