#!/bin/bash
# round 2: fused PPO gradient kernels - parity, trainer tests, bench A/B, ncu capture
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ppo_grad_kernel.py tests/test_trainer_gpu.py -q -m gpu > gpurun_out/ppo_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/ppo_tests.log
tail -8 gpurun_out/ppo_tests.log
timeout 400 python bench.py --workload ppo --steps 3 --warmup 2 --fused-update 1 > gpurun_out/ppo_fused.json 2> gpurun_out/ppo_fused.err
echo "fused exit $?"; cat gpurun_out/ppo_fused.json; tail -3 gpurun_out/ppo_fused.err
PYTHONPATH=. timeout 200 python tools/ppo_grad_probe.py > gpurun_out/ppo_probe.log 2>&1; cat gpurun_out/ppo_probe.log
PYTHONPATH=. timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_ppo_grad -c 1 -f -o gpurun_out/r2_ppo_grad \
  python tools/ppo_grad_probe.py 2097152 2 > gpurun_out/ppo_ncu.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ppo_ncu.log
