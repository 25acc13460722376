#!/bin/bash
# k_ppo_grad: ncu --set full on one launch (B = 2 097 152), raw metrics exported to CSV on the box (the .ncu-rep is too large to keep)
mkdir -p gpurun_out
PYTHONPATH=. timeout 200 python tools/ppo_grad_probe.py > gpurun_out/ppo_probe.log 2>&1; cat gpurun_out/ppo_probe.log
PYTHONPATH=. timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_ppo_grad -c 1 -f -o /tmp/ppo_grad \
  python tools/ppo_grad_probe.py 2097152 2 > gpurun_out/ppo_ncu.log 2>&1
echo "ncu exit $?"
ncu -i /tmp/ppo_grad.ncu-rep --page raw --csv > /tmp/ppo_raw.csv 2>/dev/null
python - <<'PY'
import csv
rows=list(csv.reader(open('/tmp/ppo_raw.csv')))
hdr, units, vals = rows[0], rows[1], rows[2]
keep=('gpu__time_duration.sum','launch__registers_per_thread','launch__occupancy_limit','launch__grid_size','launch__block_size','smsp__issue_active.avg.pct','sm__inst_executed_pipe_fma.avg.pct','sm__inst_executed_pipe_lsu.avg.pct','sm__inst_executed_pipe_alu.avg.pct','sm__pipe_fmaheavy_cycles_active.avg.pct','smsp__sass_thread_inst_executed_op_ffma_pred_on.sum.per_cycle_elapsed','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','l1tex__data_bank_conflicts_pipe_lsu_mem_shared','dram__bytes_read.sum','dram__bytes_write.sum','smsp__average_warps_issue_stalled','smsp__inst_executed.sum','sm__warps_active.avg','smsp__warps_eligible.avg','Kernel Name')
with open('gpurun_out/r2_ppo_grad_ncu_raw.csv','w') as f:
    f.write('metric,unit,value\n')
    for h,u,v in zip(hdr,units,vals):
        if any(h.startswith(k) or h==k for k in keep):
            f.write(f'"{h}","{u}","{v}"\n')
print(open('gpurun_out/r2_ppo_grad_ncu_raw.csv').read()[:3000])
PY
