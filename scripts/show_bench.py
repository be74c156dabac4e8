"""Print the key numbers of bench.py JSON lines: python scripts/show_bench.py file.json [...]"""
import json
import sys

for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable:", e)
        continue
    e2e = d.get("e2e") or {}
    print(f"{f}: value {d.get('value', 0):.1f} {d.get('unit')}  {d.get('ms_per_step', 0):.3f} ms/step"
          + (f"  e2e {e2e.get('value', 0):.1f} ({e2e.get('ms_per_step', 0):.3f} ms)" if e2e else ""))
    for k in ("mapping_iteration", "tracking_iteration"):
        if k in d:
            print(f"   {k}: {d[k]['value']:.1f} it/s ({d[k]['ms_per_iteration']:.3f} ms)")
    for k in ("scannetpp_frame", "ba_step"):
        if k in d:
            print(f"   {k}: {d[k]['value']:.1f} {d[k]['unit']} ({d[k]['ms_per_step']:.3f} ms/step)")
    for n, k in (d.get("kernels") or {}).items():
        print(f"   {n:32s} {k['avg_us']:8.1f} us x {k['launches_per_step']:.0f}   hbm {k.get('frac_hbm', 0):.3f}")
