#!/usr/bin/env python
"""Key metrics of every launch in an ncu report:  ncu -i rep --page raw --csv | python tools/ncu_summary.py [title]"""
import csv, sys
rows = list(csv.reader(sys.stdin))
h, u = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'launch__occupancy_limit_registers',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__inst_executed.sum', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_active',
        'lts__t_bytes.sum.per_second', 'l1tex__t_bytes.sum.per_second', 'dram__bytes.sum.per_second',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'smsp__sass_average_branch_targets_threads_uniform.pct']
if len(sys.argv) > 1:
    print("==", sys.argv[1])
for n, v in enumerate(rows[2:]):
    name = v[h.index('Kernel Name')] if 'Kernel Name' in h else ''
    print("-- launch %d  %s" % (n, name))
    for a, b, c in zip(h, u, v):
        if a in want:
            print('%s [%s] = %s' % (a, b, c))
        if a.startswith('smsp__average_warps_issue_stalled') and a.endswith('_per_issue_active.ratio'):
            try:
                if float(c) >= 0.2:
                    print('  stall %s = %s' % (a[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')], c))
            except ValueError:
                pass
