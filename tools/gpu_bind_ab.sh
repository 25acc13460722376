#!/bin/bash
# A/B of GPU-local CPU binding for the host-pointer (e2e) path at N GPUs, same box, same call
N=${1:-8}
OUT=gpurun_out/bind$N
mkdir -p $OUT
nvidia-smi topo -m > $OUT/topo.txt 2>&1
lscpu | grep -E "Socket|NUMA|^CPU\(s\)" > $OUT/lscpu.txt 2>&1
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $B --master-port 29521 bench.py --gpus $N --steps 400 --warmup 50 --no-cpu-baseline --e2e-steps 200 > $OUT/bound.json 2> $OUT/bound.err; echo "bound rc=$?" >> $OUT/rc.txt
timeout 600 $B --master-port 29522 bench.py --gpus $N --steps 400 --warmup 50 --no-cpu-baseline --e2e-steps 200 --no-bind > $OUT/nobind.json 2> $OUT/nobind.err; echo "nobind rc=$?" >> $OUT/rc.txt
cat $OUT/rc.txt; cat $OUT/lscpu.txt
for f in bound nobind; do python -c "
import json
d=json.loads([l for l in open('$OUT/$f.json').read().strip().split('\n') if l.startswith('{')][-1])
print('$f', 'value %.3e'%d['value'], 'e2e %.3e'%d['e2e']['value'], d['e2e'].get('host_binding_rank0'))"; done
