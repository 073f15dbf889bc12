#!/usr/bin/env python
"""Summarise an .ncu-rep (raw page + SASS hot spots) into text.  usage: ncu_summary.py rep [min_share]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; minshare = float(sys.argv[2]) if len(sys.argv) > 2 else 0.006
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); hdr, units = rows[0], rows[1]
want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__inst_executed.sum', 'launch__grid_size', 'launch__block_size', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'smsp__average_warp_latency_per_inst_issued.ratio',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_bytes.sum', 'l1tex__t_bytes.sum', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem']
for i, h in enumerate(hdr):
    if h in want:
        print(f"{h} [{units[i]}]:", [r[i].split('(')[0][:40] for r in rows[2:]])
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
k = 0
while k < len(rows):
    if rows[k] and rows[k][0] == 'Kernel Name':
        name = rows[k][1].split('(')[0]; hdr = rows[k + 1]; ix = {h: i for i, h in enumerate(hdr)}
        data = []; k += 2
        while k < len(rows) and not (rows[k] and rows[k][0] == 'Kernel Name'):
            if len(rows[k]) >= len(hdr): data.append(rows[k])
            k += 1
        ti = sum(int(r[ix['Instructions Executed']]) for r in data); tt = sum(int(r[ix['Thread Instructions Executed']]) for r in data)
        ts = sum(int(r[ix['# Samples']]) for r in data)
        print(f"\n== {name}: warp inst {ti}, thread inst {tt}, avg threads {tt / max(ti, 1):.2f}, samples {ts}")
        for n, r in enumerate(data):
            ie = int(r[ix['Instructions Executed']]); sm = int(r[ix['# Samples']])
            if ie > ti * minshare or sm > ts * minshare:
                print(f"{n:5d} {r[ix['Source']][:58]:58s} ie={ie:10d} thr={r[ix['Avg. Threads Executed']]:>3s} smp={sm:6d} lsb={r[ix['stall_long_sb']]:>6s}")
        break  # first launch only
    else:
        k += 1
