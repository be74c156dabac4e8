"""How long does the [P,13] map-gradient all-reduce take on this box?  torchrun --nproc-per-node N scripts/allreduce_probe.py"""
import os, torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
for n in (500_000 * 13, 2_000_000 * 13):
    x = torch.randn(n, device="cuda")
    for _ in range(10):
        dist.all_reduce(x)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        dist.all_reduce(x)
    e1.record(); torch.cuda.synchronize()
    if rank == 0:
        us = e0.elapsed_time(e1) / 50 * 1e3
        print(f"world {world}: all_reduce of {n * 4 / 1e6:.0f} MB: {us:.1f} us  ({2 * (world - 1) / world * n * 4 / us / 1e3:.0f} GB/s bus)")
dist.destroy_process_group()
