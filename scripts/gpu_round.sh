#!/bin/bash
# One GPU-box visit for the round's evidence: GPU tests, smoke, next-row kernel timings, bench (reference arm, then
# product), ncu launch list of the bench, ncu --set full of every kernel of one step and of the next-row kernels.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 300 python scripts/next_rows_bench.py > gpurun_out/next_rows.json 2> gpurun_out/next_rows.err; echo "next_rows rc=$?"
if [ -z "$SKIP_REF" ]; then
  timeout 400 python bench.py --impl reference --steps 64 --warmup 8 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
fi
timeout 400 python bench.py --steps 64 --warmup 8 > gpurun_out/bench_b200.json 2> gpurun_out/bench_b200.err; echo "bench rc=$?"
python scripts/show_bench.py gpurun_out/bench_ref.json gpurun_out/bench_b200.json
[ -n "$SKIP_NCU" ] && exit 0
B="python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-kernel-timing --no-stats"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'scatter|preprocess|tile_sort|scan|composite' -s 18 -c 9 -f -o gpurun_out/prof_all $B > gpurun_out/ncu_all.log 2>&1; echo "ncu all rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'map_step' -s 3 -c 1 -f -o gpurun_out/prof_map_step python scripts/next_rows_bench.py --iters 5 > gpurun_out/ncu_map_step.log 2>&1; echo "ncu map_step rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'pose_transform' -s 4 -c 2 -f -o gpurun_out/prof_pose python scripts/next_rows_bench.py --iters 5 > gpurun_out/ncu_pose.log 2>&1; echo "ncu pose rc=$?"
timeout 300 ncu --set full --clock-control none -k regex:'mapping_loss' -s 4 -c 2 -f -o gpurun_out/prof_loss python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-kernel-timing --no-stats --no-loops > gpurun_out/ncu_loss.log 2>&1; echo "ncu loss rc=$?"
