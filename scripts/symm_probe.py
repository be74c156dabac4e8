"""Probe torch symmetric memory / NVLS multicast on this box: torchrun --nproc-per-node N scripts/symm_probe.py"""
import os, torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
import torch.distributed._symmetric_memory as symm_mem
n = 500_000 * 13
try:
    t = symm_mem.empty(n, dtype=torch.float32, device=dev)
    hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
    if rank == 0:
        print("handle attrs:", [a for a in dir(hdl) if not a.startswith("_")])
        print("multicast_ptr:", getattr(hdl, "multicast_ptr", None), "world", hdl.world_size, "rank", hdl.rank)
        print("ops:", [o for o in dir(torch.ops.symm_mem) if not o.startswith("_")])
except Exception as e:
    print("rank", rank, "symm_mem failed:", type(e).__name__, e)
    raise

def timeit(name, fn, iters=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    if rank == 0:
        print(f"{name}: {e0.elapsed_time(e1) / iters * 1e3:.1f} us")

g = dist.group.WORLD.group_name
x = torch.randn(n, device=dev)
timeit("nccl all_reduce 26MB", lambda: dist.all_reduce(x))
for name in ("multimem_all_reduce_", "two_shot_all_reduce_", "one_shot_all_reduce"):
    op = getattr(torch.ops.symm_mem, name, None)
    if op is None:
        if rank == 0: print(name, "not available")
        continue
    try:
        t.copy_(x)
        ref = x.clone(); dist.all_reduce(ref)
        out = op(t, "sum", g)
        torch.cuda.synchronize()
        res = out if name == "one_shot_all_reduce" else t
        err = float((res - ref).abs().max())
        if rank == 0: print(name, "max err vs nccl", err)
        timeit(name + " 26MB", lambda: op(t, "sum", g))
    except Exception as e:
        if rank == 0: print(name, "failed:", type(e).__name__, str(e)[:200])
timeit("hdl.barrier", lambda: hdl.barrier(channel=0))
dist.destroy_process_group()
