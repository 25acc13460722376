#!/bin/bash
# launch list of the bench command on the round's last tree (per-launch times are cold-cache and serialised: only the shares count)
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_final3_ncu_launches.csv \
  python bench.py --steps 40 --warmup 5 --no-cpu-baseline --e2e-steps 3 > gpurun_out/ncu_launches.log 2>&1
echo "ncu list exit $?"
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open('gpurun_out/r2_final3_ncu_launches.csv')) if len(r) > 14 and r[0].isdigit()]
t=collections.defaultdict(lambda:[0,0.0])
for r in rows:
    k=r[4].split('(')[0]; t[k][0]+=1; t[k][1]+=float(r[14])
tot=sum(v[1] for v in t.values())
for k,(n,ns) in sorted(t.items(), key=lambda kv:-kv[1][1]):
    print(f'{k:60s} launches {n:4d}  total {ns/1e3:9.1f} us  share {100*ns/tot:5.1f} %  mean {ns/n/1e3:7.2f} us')
PY
