#!/usr/bin/env python
"""ncu -i X.ncu-rep --page raw --csv  ->  compact metric,unit,value csv (first captured launch).
usage: ncu_extract.py raw.csv out.csv"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
rows = [r for r in rows if len(r) > 20]
hdr, units, vals = rows[0], rows[1], rows[2]
keep = ('dram__bytes', 'gpu__time_duration', 'gpu__dram_throughput', 'launch__', 'sm__warps_active', 'sm__inst_executed',
        'smsp__issue_active', 'smsp__inst_executed.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared',
        'smsp__average_warps_issue_stalled', 'sm__throughput', 'sm__pipe_tensor', 'smsp__cycles_active.avg',
        'sm__cycles_elapsed.max', 'lts__t_sector_hit_rate', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'lts__t_bytes', 'sm__cycles_active.avg')
with open(sys.argv[2], 'w') as f:
    w = csv.writer(f)
    w.writerow(['metric', 'unit', 'value'])
    w.writerow(['kernel', '', vals[hdr.index('Kernel Name')][:120]])
    for h, u, v in zip(hdr, units, vals):
        if h.startswith(keep) and '.min.' not in h and '.max.pct' not in h and 'sum.pct' not in h:
            w.writerow([h, u, v])
