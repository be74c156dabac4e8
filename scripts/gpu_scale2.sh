#!/bin/bash
# bench.py at N = 1 and N = $1 on one box (product arm), short runs: timing statistics, gradient check, collective A/B
N="${1:-2}"
mkdir -p gpurun_out
timeout 300 python bench.py --steps 32 --warmup 4 --no-cpu-baseline --no-loops --no-extra > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err; echo "n1 rc=$?"; tail -2 gpurun_out/scale_n1.err
GS2D_BENCH_WATCHDOG_S=200 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 32 --warmup 4 --no-cpu-baseline --no-loops --no-extra > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err; echo "n$N rc=$?"; tail -3 gpurun_out/scale_n$N.err
python - <<'PY'
import json,sys,glob
for f in sorted(glob.glob("gpurun_out/scale_n*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e: print(f,"unreadable",e); continue
    t=d.get("timing",{}); e=d.get("e2e",{})
    print(f, "value %.1f  ms %.3f  p10/p90 %.3f/%.3f  max %.3f outliers %d unstable %s | e2e %.1f (%.3f ms)" % (d["value"], d["ms_per_step"], t.get("p10_ms",0), t.get("p90_ms",0), t.get("max_ms",0), t.get("n_outliers",-1), d.get("unstable"), e.get("value",0), e.get("ms_per_step",0)))
    print("   ", d.get("collective_ab"), d.get("allreduce_max_rel_err"), d["config"]["parallelism"][:120])
PY
