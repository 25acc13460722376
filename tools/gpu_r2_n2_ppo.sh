#!/bin/bash
# round 2: two GPUs - the PPO loop with the fused update (NCCL all-reduce of the kernel's gradients inside the captured graph), then the headline
mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus 2 --workload ppo --steps 3 --warmup 2 > gpurun_out/n2_ppo.json 2> gpurun_out/n2_ppo.err
echo "ppo n2 exit $?"; cat gpurun_out/n2_ppo.json | cut -c1-260; tail -3 gpurun_out/n2_ppo.err
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 \
  bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/n2_bench.json 2> gpurun_out/n2_bench.err
echo "bench n2 exit $?"; cat gpurun_out/n2_bench.json | cut -c1-260; tail -3 gpurun_out/n2_bench.err
