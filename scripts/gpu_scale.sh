#!/bin/bash
# Scaling check on one box: bench.py at the given rank counts, back to back.  Every run is bounded twice -- bench.py's
# own watchdog (GS2D_BENCH_WATCHDOG_S: every rank exits with an error line) and an outer `timeout` -- because one hung
# 8-rank run is charged 8x its wall time (round 1 lost its last GPU minutes that way).
# Start with `bash scripts/gpu_scale.sh 2` under `gpurun --gpus 2 --timeout 300` before asking for more ranks.
mkdir -p gpurun_out
export GS2D_BENCH_WATCHDOG_S=${GS2D_BENCH_WATCHDOG_S:-120}
for n in "$@"; do
  if [ "$n" = "1" ]; then
    timeout 180 python bench.py --gpus 1 --steps 64 --warmup 8 --no-cpu-baseline --no-loops > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
  else
    timeout 180 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 64 --warmup 8 --no-cpu-baseline > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
  fi
  rc=$?
  echo "N=$n rc=$rc"; python scripts/show_bench.py gpurun_out/scale_$n.json 2>/dev/null | head -1; tail -2 gpurun_out/scale_$n.err
  if [ $rc -ne 0 ]; then echo "stopping: N=$n did not finish cleanly"; break; fi
done
