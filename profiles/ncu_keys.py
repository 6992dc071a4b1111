#!/usr/bin/env python
"""print the metrics of a `ncu --set full` capture that the profile summaries quote: ncu_keys.py file.ncu-rep"""
import csv, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[0]
keys = ['gpu__time_duration.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'dram__bytes_read.sum',
        'dram__bytes_write.sum', 'launch__registers_per_thread', 'smsp__inst_executed.sum', 'launch__grid_size', 'sm__cycles_active.avg',
        'sm__cycles_elapsed.max', 'smsp__warps_eligible.avg.per_cycle_active']
keys += [h for h in hdr if 'issue_stalled' in h and 'per_issue_active' in h]
for vals in rows[2:]:
    d = dict(zip(hdr, vals))
    print("==", d.get("Kernel Name", "")[:60])
    for k in keys:
        if k in d:
            try:
                if 'issue_stalled' in k and float(d[k]) < 0.3: continue
            except ValueError: pass
            print(f"  {k.replace('smsp__average_warps_issue_stalled_','stall_').replace('_per_issue_active.ratio','')}: {d[k]}")
