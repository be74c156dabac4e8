#!/usr/bin/env python
"""Summarise an ncu launch list (gpu__time_duration.sum per launch, --csv): python scripts/launch_summary.py launches.csv"""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
h = rows[hdr]; ki = h.index('Kernel Name'); vi = h.index('Metric Value')
agg = collections.OrderedDict()
for r in rows[hdr + 1:]:
    if len(r) <= vi: continue
    a = agg.setdefault(r[ki][:80], [0, 0.0]); a[0] += 1; a[1] += float(r[vi].replace(',', ''))
tot = sum(a[1] for a in agg.values())
print(f"{'total us':>10s} {'n':>4s} {'avg us':>9s} {'share':>6s}  kernel   (cold-cache, serialised: compare shares)")
for n, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{a[1]/1e3:10.1f} {a[0]:4d} {a[1]/a[0]/1e3:9.1f} {100*a[1]/tot:5.1f}%  {n}")
