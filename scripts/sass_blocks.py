#!/usr/bin/env python
"""Group the per-instruction SASS profile printed by ncu_summary.py (min_share 0) into basic blocks
(consecutive instructions with the same execution count). usage: sass_blocks.py sass.txt"""
import re, sys
rows = []
for l in open(sys.argv[1]):
    m = re.match(r"\s*(\d+)\s+(.*?)\s+ie=\s*(\d+) thr=\s*(\d+) smp=\s*(\d+) lsb=\s*(\d+)", l)
    if m:
        rows.append((int(m[1]), m[2], int(m[3]), int(m[4]), int(m[5]), int(m[6])))
tot_ie = sum(r[2] for r in rows); tot_s = sum(r[4] for r in rows)
blocks = []
for r in rows:
    if blocks and blocks[-1]["ie"] == r[2] and blocks[-1]["thr"] == r[3]:
        b = blocks[-1]; b["n"] += 1; b["smp"] += r[4]; b["lsb"] += r[5]; b["last"] = r[0]
    else:
        blocks.append(dict(first=r[0], last=r[0], ie=r[2], thr=r[3], n=1, smp=r[4], lsb=r[5], head=r[1][:40]))
print(f"total warp inst {tot_ie}, samples {tot_s}")
for b in blocks:
    share = b["ie"] * b["n"] / tot_ie
    if share > 0.004 or b["smp"] > 0.004 * tot_s:
        print(f"{b['first']:5d}-{b['last']:5d} n={b['n']:4d} exec={b['ie']:9d} thr={b['thr']:2d} inst%={100*share:5.1f} smp%={100*b['smp']/tot_s:5.1f} lsb%={100*b['lsb']/tot_s:5.1f}  {b['head']}")
