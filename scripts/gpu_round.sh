#!/bin/bash
# One GPU-box visit: GPU tests, next-row kernel timings, bench (reference arm, then product), ncu of the new kernels.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 300 python scripts/next_rows_bench.py > gpurun_out/next_rows.json 2> gpurun_out/next_rows.err; echo "next_rows rc=$?"; cat gpurun_out/next_rows.json; tail -5 gpurun_out/next_rows.err
if [ -z "$SKIP_REF" ]; then
  timeout 400 python bench.py --impl reference --steps 64 --warmup 8 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"; cat gpurun_out/bench_ref.json; tail -5 gpurun_out/bench_ref.err
fi
timeout 400 python bench.py --steps 64 --warmup 8 > gpurun_out/bench_b200.json 2> gpurun_out/bench_b200.err; echo "bench rc=$?"; cat gpurun_out/bench_b200.json; tail -5 gpurun_out/bench_b200.err
[ -n "$SKIP_NCU" ] && exit 0
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'map_step|pose_transform' -s 20 -c 3 -f -o gpurun_out/prof_next_rows python scripts/next_rows_bench.py --iters 10 > gpurun_out/ncu_next_rows.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_next_rows.log
