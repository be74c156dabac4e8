import signal
signal.signal(signal.SIGPIPE, signal.SIG_DFL)
import json, sys
for f in sys.argv[1:] or ['gpurun_out/bench_b200.json', 'gpurun_out/bench_ref.json']:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, 'value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value'], 1) if d.get('e2e') else None, 'launches', d.get('gpu_launches'))
        for k, v in (d.get('kernels') or {}).items():
            print(f"    {k:22s} {v['ms_per_step']*1e3:8.1f} us  x{v['launches_per_step']:.0f}  hbm_frac {v.get('frac_hbm', 0):.3f}")
    except Exception as e:
        print(f, 'ERR', e)
