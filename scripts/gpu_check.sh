#!/bin/bash
# One GPU-box visit: golden fixtures from the compiled reference, parity, tests, bench (both arms), ncu launch list.
mkdir -p gpurun_out/golden
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt
if [ "$1" = "golden" ]; then
  python tests/golden/make_golden.py --out tests/golden > gpurun_out/make_golden.log 2>&1; echo "make_golden rc=$?"
  cp tests/golden/*.npz gpurun_out/golden/
fi
[ -n "$SKIP_SMALL" ] || python tests/small_parity.py > gpurun_out/small_parity.log 2>&1; echo "small_parity rc=$?"; tail -30 gpurun_out/small_parity.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
python bench.py --steps 64 --warmup 8 > gpurun_out/bench_b200.json 2> gpurun_out/bench_b200.err; echo "bench rc=$?"; cat gpurun_out/bench_b200.json; tail -5 gpurun_out/bench_b200.err
[ -n "$SKIP_REF" ] || python bench.py --impl reference --steps 64 --warmup 8 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"; cat gpurun_out/bench_ref.json; tail -5 gpurun_out/bench_ref.err
B="python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-kernel-timing --no-stats"
[ -n "$SKIP_NCU" ] && exit 0
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_launches.log 2>&1
echo "ncu rc=$?"
