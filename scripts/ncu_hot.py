#!/usr/bin/env python
"""Per-instruction hot spots of one kernel in an .ncu-rep.  usage: ncu_hot.py rep kernel_regex [min_share]"""
import csv, io, subprocess, sys
rep, kre = sys.argv[1], sys.argv[2]
ms = float(sys.argv[3]) if len(sys.argv) > 3 else 0.012
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name", "regex:" + kre],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
k = max(i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name')
hdr = rows[k + 1]; ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[k + 2:] if len(r) >= len(hdr)]
ts = sum(int(r[ix['# Samples']]) for r in data); ti = sum(int(r[ix['Instructions Executed']]) for r in data)
print(rows[k][1][:60], 'samples', ts, 'warp inst', ti)
cols = ['stall_long_sb', 'stall_membar', 'stall_lg', 'stall_barrier', 'stall_short_sb', 'stall_wait', 'stall_math']
for n, r in enumerate(data):
    sm = int(r[ix['# Samples']]); ie = int(r[ix['Instructions Executed']])
    if sm > ts * ms or ie > ti * ms:
        print(f"{n:5d} {r[ix['Source']][:56]:56s} ie={ie:9d} thr={r[ix['Avg. Threads Executed']]:>3s} smp%={100*sm/ts:5.1f} " +
              ' '.join(f"{c[6:]}={r[ix[c]]}" for c in cols if r[ix[c]] not in ('0', '')))
