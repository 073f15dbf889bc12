#!/usr/bin/env python
"""Print selected metrics of an `ncu --metrics ... --csv` log per launch. usage: ncu_metrics.py file.csv [launch ids...]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'ID')
hdr = rows[hi]
ix = {h: i for i, h in enumerate(hdr)}
d = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) < len(hdr):
        continue
    d.setdefault(r[ix['ID']], collections.OrderedDict())[r[ix['Metric Name']]] = r[ix['Metric Value']]
ids = sys.argv[2:] or list(d)
names = list(next(iter(d.values())))
print('%-66s' % 'metric' + ''.join('%16s' % ('launch ' + i) for i in ids))
for n in names:
    print('%-66s' % n + ''.join('%16s' % d[i].get(n, '') for i in ids))
