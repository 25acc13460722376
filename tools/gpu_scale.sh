#!/bin/bash
# N-GPU run of both bench arms, launched as the driver launches them
N=${1:-8}
OUT=gpurun_out/scale$N
mkdir -p $OUT
nvidia-smi -L > $OUT/smi.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?" >> $OUT/rc.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 2000 --warmup 200 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" >> $OUT/rc.txt
cat $OUT/rc.txt; tail -c 600 $OUT/bench.json; tail -3 $OUT/bench.err
