#!/bin/bash
# Multi-GPU visit: gpu_multi.sh N  -> gradient parity + collective timing on N GPUs (scripts/fused_check.py)
N="${1:-2}"
mkdir -p gpurun_out
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  scripts/fused_check.py --out gpurun_out/multi_gpu_check_n$N.json > gpurun_out/fused_check_n$N.log 2>&1
echo "fused_check rc=$?"; tail -4 gpurun_out/fused_check_n$N.log | cut -c1-1500
