#!/usr/bin/env python
"""SASS-order profile of one kernel of an .ncu-rep, cut into regions at barriers:
python scripts/ncu_sass_regions.py file.ncu-rep <kernel regex>
Prints, per run of SASS instructions between CTA barriers, warp instructions, thread instructions, lanes per
instruction and stall samples -- which phase of a kernel wastes lanes."""
import csv, io, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass', '-k', 'regex:' + pat],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = None
regions, cur = [], dict(n=0, wi=0, ti=0, smp=0, first=None)
seen = 0
for r in rows:
    if r and r[0] == 'Address':
        if hdr is not None:
            break          # first matching launch only
        hdr = r
        continue
    if hdr is None or not r or not r[0].startswith('0x'):
        continue
    d = dict(zip(hdr, r))
    wi, ti, smp = int(d['Instructions Executed']), int(d['Thread Instructions Executed']), int(d['# Samples'])
    src = d['Source'].strip()
    if cur['first'] is None:
        cur['first'] = seen
    cur['n'] += 1; cur['wi'] += wi; cur['ti'] += ti; cur['smp'] += smp
    seen += 1
    if 'BAR.SYNC' in src or 'BAR.RED' in src or src.startswith('EXIT'):
        cur['end'] = src
        regions.append(cur)
        cur = dict(n=0, wi=0, ti=0, smp=0, first=None)
if cur['n']:
    cur['end'] = '(end)'
    regions.append(cur)
tw = sum(x['wi'] for x in regions); ts = sum(x['smp'] for x in regions)
print(f"total warp inst {tw}, thread inst {sum(x['ti'] for x in regions)}, samples {ts}")
for x in regions:
    if x['wi'] == 0:
        continue
    print(f"  sass[{x['first']:5d}+{x['n']:4d}] warp inst {100 * x['wi'] / tw:5.1f}%  lanes/inst {x['ti'] / x['wi']:5.1f}  samples {100 * x['smp'] / max(ts, 1):5.1f}%  ends at {x['end'][:40]}")
