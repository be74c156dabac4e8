#!/usr/bin/env python
"""Print the key metrics of every kernel in an .ncu-rep (read here, no GPU needed): python scripts/ncu_summary.py file.ncu-rep"""
import csv, subprocess, sys, io
WANT = ['gpu__time_duration.sum', 'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'dram__bytes_read.sum',
        'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'smsp__inst_executed_pipe_fma.sum', 'smsp__inst_executed_pipe_alu.sum', 'smsp__inst_executed_pipe_xu.sum',
        'smsp__inst_executed_pipe_lsu.sum', 'smsp__inst_executed_op_shared_atom.sum', 'lts__t_sectors_op_red.sum',
        'lts__t_sectors_op_atom.sum', 'smsp__average_warp_latency_per_inst_issued.ratio']
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = rows[0]
stall = [c for c in h if 'warp_issue_stalled' in c and c.endswith('_per_warp_active.pct')]
for r in rows[2:]:
    print('=====', r[h.index('Kernel Name')][:90])
    for w in WANT:
        if w in h:
            print(f"  {w:72s} {r[h.index(w)]:>18s} {rows[1][h.index(w)]}")
    st = sorted(((float(r[h.index(c)].replace(',', '')), c) for c in stall), reverse=True)[:6]
    for v, c in st:
        print(f"  stall {c.replace('smsp__warp_issue_stalled_', '').replace('_per_warp_active.pct', ''):40s} {v:8.1f} %")
