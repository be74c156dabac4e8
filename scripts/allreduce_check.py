"""Own in-switch all-reduce vs NCCL: correctness and time.  torchrun --nproc-per-node N scripts/allreduce_check.py"""
import os, sys, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gaus-slam_b200"))
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
from gs2d_b200 import mapping, _lib
import ctypes as C
P = 500_000
flat = mapping.symmetric_gradient_buffer(P, dev)
if flat is None:
    print("rank", rank, "no symmetric memory / multicast"); sys.exit(0)
torch.manual_seed(rank)
x = torch.randn(13 * P, device=dev)
ref = x.clone(); dist.all_reduce(ref)
flat.copy_(x)
mapping.allreduce_map_gradient(flat)
torch.cuda.synchronize()
err = float((flat - ref).abs().max()); scale = float(ref.abs().max())
print(f"rank {rank}: max |own - nccl| = {err:.3e} (max |ref| {scale:.2f})")
assert err <= 1e-5 * scale

def timeit(name, fn, iters=50):
    for _ in range(5): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters * 1e3], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0: print(f"{name}: {float(t):.1f} us")

timeit("nccl all_reduce 26 MB", lambda: dist.all_reduce(x))
timeit("own multimem all-reduce 26 MB (incl. 2 barriers)", lambda: mapping.allreduce_map_gradient(flat))
hdl, n, r, w = mapping._symmetric[flat.data_ptr()]
lib = _lib.load()
for blocks in (74, 148, 296, 592, 1184):
    timeit(f"  kernel only, {blocks} blocks", lambda: lib.gs2d_allreduce_multimem(C.c_void_p(hdl.multicast_ptr), n, r, w, blocks, C.c_void_p(torch.cuda.current_stream().cuda_stream)))
dist.destroy_process_group()
