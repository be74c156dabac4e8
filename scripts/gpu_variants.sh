#!/bin/bash
# Compare tuning variants of libgs2d (built with `build.py --variant=<name> -D...`): value + per-kernel times.
mkdir -p gpurun_out
for v in "$@"; do
  if [ "$v" = "default" ]; then unset GS2D_LIBRARY; else export GS2D_LIBRARY=$PWD/gaus-slam_b200/lib/libgs2d_$v.so; fi
  python bench.py --steps 48 --warmup 8 --no-e2e --no-extra --no-cpu-baseline > gpurun_out/variant_$v.json 2> gpurun_out/variant_$v.err
  echo "variant $v rc=$?"; python scripts/show_bench.py gpurun_out/variant_$v.json
done
