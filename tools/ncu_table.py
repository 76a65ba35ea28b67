"""Tabulate the per-kernel metrics of an `ncu --page raw --csv` dump (one row per launch)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
H = rows[0]
keys = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio', 'l1tex__throughput.avg.pct_of_peak_sustained_active',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'launch__grid_size',
        'launch__block_size', 'smsp__thread_inst_executed_per_inst_executed.ratio']
short = ['name', 'us', 'rdMB', 'wrMB', 'regs', 'warps%', 'inst', 'issue%', 'longsb', 'shortsb', 'lg', 'mio', 'wait', 'math', 'bar', 'l1%', 'l2%', 'dram%', 'grid', 'blk', 'thr']
idx = [(k, H.index(k)) for k in keys if k in H]
units = rows[1]
print(' '.join(short))
print('units:', ' '.join(units[i] or '-' for _, i in idx))
for r in rows[2:]:
    out = []
    for k, i in idx:
        v = r[i]
        if k == 'Kernel Name':
            v = v.replace('orbfe::<unnamed>::', '')[:22]
        else:
            try: v = f"{float(v.replace(',', '')):.4g}"
            except ValueError: pass
        out.append(v)
    print(' '.join(out))
