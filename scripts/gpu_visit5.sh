#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
for i in 1 2; do
timeout 400 python bench.py --steps 64 --warmup 8 --no-cpu-baseline > gpurun_out/bench_b200_$i.json 2> gpurun_out/bench_b200_$i.err; echo "bench rc=$?"
python scripts/show_bench.py gpurun_out/bench_b200_$i.json | head -4
done
