#!/bin/bash
# N-GPU run of both bench arms (driver form) and the PPO workload, launched as the driver launches them
N=${1:-2}
OUT=gpurun_out/r2_scale$N
mkdir -p $OUT
nvidia-smi -L > $OUT/smi.txt
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $R --master-port 29511 bench.py --impl reference --gpus $N --steps 20 --warmup 3 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?" >> $OUT/rc.txt
timeout 600 $R --master-port 29512 bench.py --gpus $N --steps 20 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" >> $OUT/rc.txt
timeout 600 $R --master-port 29513 bench.py --gpus $N --workload ppo --steps 5 --warmup 2 > $OUT/bench_ppo.json 2>> $OUT/bench.err; echo "ppo rc=$?" >> $OUT/rc.txt
cat $OUT/rc.txt
for f in bench bench_ppo bench_ref; do python -c "
import json
d=json.loads(open('$OUT/$f.json').read().strip().split('\n')[-1]); print('$f n_gpus', d.get('n_gpus'), '%.3e'%d['value'], d.get('ms_per_step'), (d.get('roofline') or {}).get('frac'), 'e2e', (d.get('e2e') or {}).get('value'), json.dumps(d.get('side', {}))[:400] if '$f'=='bench_ppo' else '')"; done
tail -3 $OUT/bench.err
