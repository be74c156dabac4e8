#!/bin/bash
# Round-2 evidence visit on one B200: GPU tests, smoke, both bench arms, ncu launch list of the bench step,
# ncu --set full (+ source counters) of every rasterizer kernel of one step.  TAG names the outputs.
TAG="${1:-r2}"
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi_$TAG.txt 2>&1
if [ -z "$SKIP_TESTS" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$TAG.log
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke_$TAG.log
fi
if [ -z "$SKIP_REF" ]; then
  timeout 500 python bench.py --impl reference > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "bench ref rc=$?"
fi
timeout 600 python bench.py ${BENCH_FLAGS} > gpurun_out/bench_b200_$TAG.json 2> gpurun_out/bench_b200_$TAG.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_b200_$TAG.err
python scripts/show_bench.py gpurun_out/bench_ref_$TAG.json gpurun_out/bench_b200_$TAG.json
[ -n "$SKIP_NCU" ] && exit 0
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-extra --no-cpu-baseline --no-kernel-timing --no-stats --no-loops"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $B > gpurun_out/ncu_launches_$TAG.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'scatter|preprocess|tile_sort|scan|composite' -s 18 -c 9 -f -o gpurun_out/prof_all_$TAG $B > gpurun_out/ncu_all_$TAG.log 2>&1; echo "ncu all rc=$?"
ncu -i gpurun_out/prof_all_$TAG.ncu-rep --page raw --csv > gpurun_out/raw_all_$TAG.csv 2>/dev/null
python scripts/ncu_lines.py gpurun_out/prof_all_$TAG.ncu-rep 50 > gpurun_out/lines_all_$TAG.txt 2>&1
rm -f gpurun_out/prof_all_$TAG.ncu-rep.tmp
ls -la gpurun_out | head -40
