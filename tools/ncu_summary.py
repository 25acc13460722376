"""ncu_summary.py - Markdown table of the judged metrics of an .ncu-rep capture (read here, on the CPU box).

    python tools/ncu_summary.py gpurun_out/<tag>/k_step_full.ncu-rep "title" "command line" > profiles/<name>.md

Also prints the warp-stall sampling totals of the first captured launch (source page) when --stalls is given.
"""
import collections
import csv
import io
import subprocess
import sys

METRICS = [
    'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
    'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'launch__registers_per_thread',
    'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'launch__waves_per_multiprocessor',
    'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
    'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
    'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
    'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
    'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio',
    'smsp__warps_eligible.avg.per_cycle_active',
]


def ncu(*args):
    return subprocess.run(['ncu', *args], capture_output=True, text=True).stdout


def main():
    rep, title, cmd = sys.argv[1], sys.argv[2], sys.argv[3]
    rows = list(csv.reader(io.StringIO(ncu('-i', rep, '--page', 'raw', '--csv'))))
    hdr, units, data = rows[0], rows[1], rows[2:]
    print(f'# {title}\n\nCommand: `{cmd}`\n')
    print('| metric | unit | ' + ' | '.join(f'launch {i + 1}' for i in range(len(data))) + ' |')
    print('|---|---|' + '---|' * len(data))
    for m in METRICS:
        if m in hdr:
            i = hdr.index(m)
            print(f'| {m} | {units[i]} | ' + ' | '.join(r[i] for r in data) + ' |')
    if '--stalls' in sys.argv:
        rows = list(csv.reader(io.StringIO(ncu('-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'))))
        hdr, tot, samples, seen = None, collections.Counter(), 0, 0
        for r in rows:
            if r and r[0] == 'Kernel Name':
                seen += 1
                if seen > 1:
                    break
                continue
            if r and r[0] == 'Address':
                hdr = r
                continue
            if hdr and len(r) > 5:
                ix = {h: i for i, h in enumerate(hdr)}
                samples += int(r[ix['# Samples']])
                for h in hdr:
                    if h.startswith('stall_') and 'Not Issued' not in h:
                        tot[h] += int(r[ix[h]])
        print(f'\nWarp-stall sampling, launch 1 ({samples} samples): ' +
              ', '.join(f'{k[6:]} {100 * v / max(1, samples):.1f} %' for k, v in tot.most_common(9)))


main()
