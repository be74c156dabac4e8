#!/bin/bash
# Quick GPU visit: GPU test suite, then a short bench (value, e2e, per-kernel table).  TAG names the output files.
TAG="${1:-q}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_$TAG.log
timeout 400 python bench.py --steps 48 --warmup 8 --no-cpu-baseline --no-loops --no-extra > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_$TAG.err
python scripts/show_bench.py gpurun_out/bench_$TAG.json
