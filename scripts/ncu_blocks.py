#!/usr/bin/env python
"""Basic-block view of an .ncu-rep: consecutive SASS instructions with the same execution count are merged.
python scripts/ncu_blocks.py file.ncu-rep <kernel substring> [min_pct]"""
import csv, io, subprocess, sys
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'source', '--csv', '--print-source', 'sass'],
                     capture_output=True, text=True).stdout
want = sys.argv[2] if len(sys.argv) > 2 else ''
min_pct = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
cur, hdr, data = None, None, {}
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == 'Kernel Name':
        cur = r[1][:70]; data[cur] = []; continue
    if r[0] == 'Address':
        hdr = r; continue
    d = dict(zip(hdr, r))
    try:
        n = int(d['Instructions Executed'])
    except Exception:
        continue
    data[cur].append((n, d['Source'].strip(), int(d['# Samples'] or 0), d['Avg. Threads Executed']))
for k, v in data.items():
    if want not in k:
        continue
    tot = sum(x[0] for x in v); smp = sum(x[2] for x in v)
    print(k, 'instructions', tot, 'sass lines', len(v), 'samples', smp)
    blocks = []
    for i, x in enumerate(v):
        if blocks and blocks[-1][1] == x[0]:
            blocks[-1][2] += 1; blocks[-1][4] += x[2]
        else:
            blocks.append([i, x[0], 1, x[3], x[2]])
    for b in blocks:
        if b[1] * b[2] > min_pct / 100 * tot:
            print(f"  at {b[0]:4d} count {b[1]:>9d} x {b[2]:3d} = {100 * b[1] * b[2] / tot:5.1f}% inst {100 * b[4] / max(smp, 1):5.1f}% smp  thr {b[3]:>3s}  {v[b[0]][1][:50]}")
