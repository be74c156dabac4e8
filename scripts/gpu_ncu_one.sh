#!/bin/bash
# ncu --set full + source counters of one kernel family of the bench step: gpu_ncu_one.sh <regex> <tag> [skip] [count]
PAT="$1"; TAG="$2"
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-extra --no-cpu-baseline --no-kernel-timing --no-stats"
$B > gpurun_out/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_$TAG.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$PAT -s ${3:-4} -c ${4:-1} -f -o gpurun_out/prof_$TAG $B > gpurun_out/ncu_$TAG.log 2>&1; echo "ncu rc=$?"
python scripts/ncu_lines.py gpurun_out/prof_$TAG.ncu-rep 60 > gpurun_out/lines_$TAG.txt 2>&1
ncu -i gpurun_out/prof_$TAG.ncu-rep --page raw --csv > gpurun_out/raw_$TAG.csv 2>/dev/null
head -c 600 gpurun_out/lines_$TAG.txt
