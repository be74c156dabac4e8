"""Where does the HOST spend its time in a mapping iteration through the public API?  cProfile over N steps
(no per-step synchronisation): python scripts/host_profile.py [N]"""
import cProfile, os, pstats, sys, io, argparse
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
a = argparse.Namespace(surfels=500000, views=8)
dev = torch.device("cuda:0")
wl = bench.build_workload(a, dev)
arm = bench.make_b200_arm(wl, dev)
from gs2d_b200 import loss as gs2d_loss
s = wl["sets"][0]
params = {k: s[k].detach().clone().requires_grad_(True) for k in bench.PARAMS}
means2D = torch.zeros((wl["P"], 3), device=dev, requires_grad=True)


def step():
    for p in params.values():
        p.grad = None
    means2D.grad = None
    rast = arm.rasterizer(s)
    color, radii, allmap = rast(means3D=params["means3D"], means2D=means2D, opacities=params["opacities"],
                                colors_precomp=params["colors"], scales=params["scales"], rotations=params["rotations"])
    loss = gs2d_loss.mapping_loss(color, allmap, s["gt_color"], s["gt_depth"])
    loss.backward()


N = int(sys.argv[1]) if len(sys.argv) > 1 else 300
for _ in range(30):
    step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(N):
    step()
pr.disable()
torch.cuda.synchronize()
out = io.StringIO()
st = pstats.Stats(pr, stream=out).sort_stats("tottime")
st.print_stats(28)
txt = out.getvalue().replace(ROOT + "/", "")
print(f"per step (us): total {1e6 * st.total_tt / N:.0f}")
print(txt[txt.index("ncalls"):])
