#!/usr/bin/env python
"""Aggregate per-line instruction counts of an .ncu-rep into named line ranges:
python scripts/ncu_ranges.py file.ncu-rep "name=file:lo-hi,file:lo-hi" ..."""
import csv, io, subprocess, sys
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'source', '--csv', '--print-source', 'cuda,sass'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
lines = {}
fname = ''
for r in rows:
    if not r: continue
    if r[0] == 'File Path': fname = r[1].split('/')[-1]
    elif r[0] == 'Line No': hdr = r
    elif r[0].isdigit():
        d = dict(zip(hdr, r))
        try: n = int(d['Instructions Executed']); t = int(d['Thread Instructions Executed'])
        except Exception: continue
        k = (fname, int(r[0])); a = lines.setdefault(k, [0, 0]); a[0] += n; a[1] += t
tot = sum(v[0] for v in lines.values()); used = set()
for spec in sys.argv[2:]:
    name, rs = spec.split('=')
    n = t = 0
    for part in rs.split(','):
        f, lh = part.split(':'); lo, hi = map(int, lh.split('-'))
        for (ff, ln), v in lines.items():
            if ff == f and lo <= ln <= hi and (ff, ln) not in used:
                n += v[0]; t += v[1]; used.add((ff, ln))
    print(f"{name:28s} {n/1e6:9.1f} M inst  {100*n/tot:5.1f}%  lanes {t/max(n,1):5.1f}")
rest = sum(v[0] for k, v in lines.items() if k not in used)
print(f"{'(other)':28s} {rest/1e6:9.1f} M inst  {100*rest/tot:5.1f}%   total {tot/1e6:.1f} M")
other = sorted(((v[0], k) for k, v in lines.items() if k not in used), reverse=True)[:8]
for n, k in other: print('     other:', k, f"{n/1e6:.1f}M")
