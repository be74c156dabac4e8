#!/bin/bash
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-kernel-timing --no-stats"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'scatter|preprocess|tile_sort|scan|composite' -s 18 -c 9 -f -o gpurun_out/prof_all $B > gpurun_out/ncu_all.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_all.log
