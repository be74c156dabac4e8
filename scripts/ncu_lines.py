#!/usr/bin/env python
"""Per-source-line instruction counts of an .ncu-rep: python scripts/ncu_lines.py file.ncu-rep [top]"""
import csv, io, subprocess, sys
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'source', '--csv', '--print-source', 'cuda,sass'],
                     capture_output=True, text=True).stdout
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(io.StringIO(out)))
sect, cur, fname = [], None, ''
for r in rows:
    if not r:
        continue
    if r[0] == 'Function Name':
        cur = next((x for x in sect if x['kernel'] == r[1][:70]), None)
        if cur is None:
            cur = dict(kernel=r[1][:70], files={})
            sect.append(cur)
    elif r[0] == 'File Path':
        fname = r[1].split('/')[-1]
    elif r[0] == 'Line No':
        hdr = r
    elif cur is not None and r[0].isdigit():
        d = dict(zip(hdr, r))
        try:
            n = int(d['Instructions Executed'])
        except Exception:
            continue
        cur['files'].setdefault(fname, []).append((n, int(r[0]), r[1].strip()[:90], d.get('# Samples', '0'),
                                                   d.get('Avg. Threads Executed', '')))
for s in sect:
    tot = sum(n for f in s['files'].values() for n, *_ in f)
    print('=====', s['kernel'], 'total inst', tot)
    allv = [(n, f, ln, src, smp, thr) for f, v in s['files'].items() for n, ln, src, smp, thr in v]
    for n, f, ln, src, smp, thr in sorted(allv, reverse=True)[:top]:
        print(f"  {100 * n / max(tot, 1):5.1f}% {n:>11d} thr={thr:>3s} smp={smp:>5s} {f}:{ln}: {src}")
