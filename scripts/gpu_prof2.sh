#!/bin/bash
# GPU tests + bench + ncu (full set, source counters) of the two compositing kernels.  TAG names the outputs.
TAG="${1:-p}"
bash scripts/gpu_quick.sh $TAG
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-extra --no-cpu-baseline --no-kernel-timing --no-stats"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:composite -s 8 -c 2 -f -o gpurun_out/prof_$TAG $B > gpurun_out/ncu_$TAG.log 2>&1; echo "ncu rc=$?"
python scripts/ncu_lines.py gpurun_out/prof_$TAG.ncu-rep 45 > gpurun_out/lines_$TAG.txt 2>&1
ncu -i gpurun_out/prof_$TAG.ncu-rep --page raw --csv > gpurun_out/raw_$TAG.csv 2>/dev/null
