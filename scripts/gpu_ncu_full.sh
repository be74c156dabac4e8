#!/bin/bash
# ncu --set full of the compositing kernels (one capture each) on the bench workload.
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-kernel-timing --no-stats"
PAT="${1:-composite}"
$B > gpurun_out/plain_full.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:$PAT -s ${3:-6} -c ${2:-2} -f -o gpurun_out/prof_$PAT $B > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_full.log
