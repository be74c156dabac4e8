#!/bin/bash
# Scaling check on one box: bench.py at the given rank counts, back to back.
mkdir -p gpurun_out
for n in "$@"; do
  if [ "$n" = "1" ]; then
    timeout 400 python bench.py --gpus 1 --steps 64 --warmup 8 --no-cpu-baseline --no-loops > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
  else
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 64 --warmup 8 --no-cpu-baseline > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
  fi
  echo "N=$n rc=$?"; python scripts/show_bench.py gpurun_out/scale_$n.json 2>/dev/null | head -1; tail -2 gpurun_out/scale_$n.err
done
