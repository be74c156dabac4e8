#!/bin/bash
# Round-2 first visit: pair statistics, per-kernel baseline, source-level ncu of the kernels of one step.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 300 python scripts/pair_stats.py 500000 0 > gpurun_out/pair_stats_v0.json 2> gpurun_out/pair_stats.err; echo "stats rc=$?"
timeout 300 python scripts/pair_stats.py 500000 5 > gpurun_out/pair_stats_v5.json 2>> gpurun_out/pair_stats.err; echo "stats rc=$?"
timeout 400 python bench.py --steps 48 --warmup 8 --no-cpu-baseline --no-loops > gpurun_out/bench_d1.json 2> gpurun_out/bench_d1.err; echo "bench rc=$?"
python scripts/show_bench.py gpurun_out/bench_d1.json
B="python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-kernel-timing --no-stats"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'scatter|preprocess|tile_sort|scan|composite' -s 18 -c 9 -f -o gpurun_out/prof_d1 $B > gpurun_out/ncu_d1.log 2>&1; echo "ncu rc=$?"
python scripts/ncu_lines.py gpurun_out/prof_d1.ncu-rep 70 > gpurun_out/lines_d1.txt 2>&1
ls -la gpurun_out
