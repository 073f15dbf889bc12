#!/usr/bin/env python
"""CUDA-source-line hot spots of one kernel in an .ncu-rep (compiled with -lineinfo, captured with
--import-source on).  usage: ncu_src.py rep kernel_regex [min_share]"""
import csv, io, subprocess, sys
rep, kre = sys.argv[1], sys.argv[2]
ms = float(sys.argv[3]) if len(sys.argv) > 3 else 0.012
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kre],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
agg = {}
hdr = None
for r in rows:
    if r and r[0] == 'Line No':
        hdr = r
        iS, iI, iT = hdr.index('# Samples'), hdr.index('Instructions Executed'), hdr.index('Thread Instructions Executed')
        continue
    if hdr is None or len(r) <= iT or r[0] == '':
        continue
    try:
        key = (int(r[0]), r[1].strip())
        a = agg.setdefault(key, [0, 0, 0])
        a[0] += int(r[iI] or 0); a[1] += int(r[iS] or 0); a[2] += int(r[iT] or 0)
    except ValueError:
        pass
ti = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
print('warp inst', ti, 'samples', ts)
for (ln, src), a in sorted(agg.items()):
    if a[0] > ti * ms or a[1] > ts * ms:
        print('%5d %-88s inst%%=%4.1f smp%%=%4.1f thr=%4.1f' % (ln, src[:88], 100 * a[0] / ti, 100 * a[1] / ts, a[2] / max(a[0], 1)))
