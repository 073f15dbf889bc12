#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel. usage: launch_summary.py launches.csv"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'ID')
hdr = rows[hi]; ix = {h: i for i, h in enumerate(hdr)}
agg = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) < len(hdr): continue
    name = r[ix['Kernel Name']].split('(')[0]
    try: v = float(r[ix['Metric Value']].replace(',', ''))
    except ValueError: continue
    unit = r[ix['Metric Unit']]
    v *= {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 's': 1e3}.get(unit, 1e-6)
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
print(f"{'kernel':48s} {'launches':>8s} {'total ms':>10s} {'avg ms':>9s} {'share':>7s}")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:48]:48s} {a[0]:8d} {a[1]:10.3f} {a[1]/a[0]:9.4f} {a[1]/tot*100:6.1f}%")
print(f"{'total':48s} {sum(a[0] for a in agg.values()):8d} {tot:10.3f}")
