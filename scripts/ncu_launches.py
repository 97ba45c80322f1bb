#!/usr/bin/env python3
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time and share per kernel."""
import csv, sys
from collections import defaultdict
rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
H = rows[hdr]
kn, mv, mn, mu = H.index('Kernel Name'), H.index('Metric Value'), H.index('Metric Name'), H.index('Metric Unit')
agg = defaultdict(lambda: [0, 0.0])
for r in rows[hdr + 1:]:
    if len(r) <= mv or r[mn] != 'gpu__time_duration.sum':
        continue
    scale = {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 's': 1e3}.get(r[mu].replace('second', 's').replace('usecond', 'us').replace('nsecond', 'ns').replace('msecond', 'ms'), 1e-6)
    a = agg[r[kn][:90]]; a[0] += 1; a[1] += float(r[mv].replace(',', '')) * scale
tot = sum(v[1] for v in agg.values())
print(f"total {tot:.3f} ms over {sum(v[0] for v in agg.values())} launches")
for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{v[1]:10.3f} ms {v[0]:4d} launches {100 * v[1] / tot:5.1f}%  {k}")
