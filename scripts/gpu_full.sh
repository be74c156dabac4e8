#!/bin/bash
# The driver's two default runs (reference arm, then product) with wall-clock times.  TAG names the outputs.
TAG="${1:-full}"
mkdir -p gpurun_out
t0=$(date +%s); timeout 900 python bench.py --impl reference > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "ref rc=$? $(( $(date +%s) - t0 )) s"; tail -2 gpurun_out/bench_ref_$TAG.err
t0=$(date +%s); timeout 900 python bench.py > gpurun_out/bench_b200_$TAG.json 2> gpurun_out/bench_b200_$TAG.err; echo "b200 rc=$? $(( $(date +%s) - t0 )) s"; tail -2 gpurun_out/bench_b200_$TAG.err
python scripts/show_bench.py gpurun_out/bench_ref_$TAG.json gpurun_out/bench_b200_$TAG.json
