"""Where does an e2e mapping step spend its time?  CPU wall per segment (no extra syncs) + GPU events."""
import os, sys, time, argparse
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
a = argparse.Namespace(surfels=500000, views=8)
dev = torch.device("cuda:0")
wl = bench.build_workload(a, dev)
arm = bench.make_b200_arm(wl, dev) if len(sys.argv) < 2 else bench.make_reference_arm(wl, dev)
s = wl["sets"][0]
params = {k: s[k].detach().clone().requires_grad_(True) for k in bench.PARAMS}
means2D = torch.zeros((wl["P"], 3), device=dev, requires_grad=True)
H, W = wl["H"], wl["W"]
if len(sys.argv) < 2:
    from gs2d_b200 import loss as gs2d_loss
    loss_fn = gs2d_loss.mapping_loss
else:
    loss_fn = torch.cuda.make_graphed_callables(bench.mapping_loss, (
        torch.rand((3, H, W), device=dev, requires_grad=True), torch.rand((7, H, W), device=dev, requires_grad=True),
        s["gt_color"].clone(), s["gt_depth"].clone()))
seg = {k: 0.0 for k in ("raster_ctor", "forward", "loss", "backward", "sync")}
gpu = {k: 0.0 for k in ("forward", "loss", "backward")}
N = 40
for it in range(N + 10):
    if it == 10:
        seg = {k: 0.0 for k in seg}; gpu = {k: 0.0 for k in gpu}
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    for p in params.values(): p.grad = None
    means2D.grad = None
    t0 = time.perf_counter()
    rast = arm.rasterizer(s)
    t1 = time.perf_counter(); ev[0].record()
    color, radii, allmap = rast(means3D=params["means3D"], means2D=means2D, opacities=params["opacities"],
                                colors_precomp=params["colors"], scales=params["scales"], rotations=params["rotations"])
    t2 = time.perf_counter(); ev[1].record()
    loss = loss_fn(color, allmap, s["gt_color"], s["gt_depth"])
    t3 = time.perf_counter(); ev[2].record()
    loss.backward()
    t4 = time.perf_counter(); ev[3].record()
    torch.cuda.synchronize()
    t5 = time.perf_counter()
    seg["raster_ctor"] += t1 - t0; seg["forward"] += t2 - t1; seg["loss"] += t3 - t2; seg["backward"] += t4 - t3; seg["sync"] += t5 - t4
    gpu["forward"] += ev[0].elapsed_time(ev[1]); gpu["loss"] += ev[1].elapsed_time(ev[2]); gpu["backward"] += ev[2].elapsed_time(ev[3])
print("CPU ms/step:", {k: round(1e3 * v / N, 3) for k, v in seg.items()}, "total", round(1e3 * sum(seg.values()) / N, 3))
print("GPU ms/step (event spans):", {k: round(v / N, 3) for k, v in gpu.items()})
