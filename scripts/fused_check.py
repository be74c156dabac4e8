"""Multi-GPU gradient parity and collective timing on hardware (SURVEY.md 8e "parity under sharding").

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        scripts/fused_check.py [--surfels 500000] [--out gpurun_out/multi_gpu_check.json]

Every rank renders its own keyframe of the bench window (fwd + bwd).  Checked on rank 0 (exit code 1 on failure):
  * NCCL all-reduce of the product's per-view gradients == the sum of the per-view gradients (gathered): the collective;
  * the product's summed gradient vs the sum of the per-view gradients of the UNMODIFIED reference rasterizer
    (oracle/_ref, when present): 1e-3 relative, BASELINE.json's gradient bar;
  * libgs2d's own in-switch all-reduce (multimem.ld_reduce / multimem.st) and the FUSED path (gs2d_backward_reduce: the
    backward-preprocess epilogue multimem.red's into every rank's copy) == the NCCL sum (fp32 re-association only).
Timed (CUDA events, max over ranks): backward + each of the three collectives, and the collectives alone.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gaus-slam_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--surfels", type=int, default=500_000)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    import gaus_2dgs_rasterization as G
    from gaus_2dgs_rasterization import _C
    from gs2d_b200 import mapping, scene
    intr = scene.REPLICA
    W, H, P = intr["W"], intr["H"], a.surfels
    sc = scene.room_scene(P, intr, seed=0)
    cam = scene.view_sweep(8, intr, seed=0)[rank % 8]
    gc, go = scene.image_grads(W, H, seed=rank % 8)
    t = lambda x: torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).to(dev)   # noqa: E731
    params = {k: t(sc[k]).requires_grad_(True) for k in mapping.PARAM_ORDER}
    m2d = torch.zeros((P, 3), device=dev, requires_grad=True)
    dcol, doth = t(gc), t(go)
    st = G.GaussianRasterizationSettings(
        image_height=H, image_width=W, tanfovx=float(cam.tanfovx), tanfovy=float(cam.tanfovy), bg=torch.zeros(3, device=dev),
        scale_modifier=1.0, viewmatrix=t(cam.viewmatrix.reshape(16)), projmatrix=t(cam.projmatrix.reshape(16)), sh_degree=0,
        campos=t(cam.campos), use_sa=True, prefiltered=False, debug=False)
    rast = G.GaussianRasterizer(st)

    def fwd_bwd():
        for p in params.values():
            p.grad = None
        color, radii, allmap = rast(means3D=params["means3D"], means2D=m2d, opacities=params["opacities"],
                                    colors_precomp=params["colors"], scales=params["scales"], rotations=params["rotations"])
        torch.autograd.backward([color, allmap], [dcol, doth])
        return {k: params[k].grad for k in mapping.PARAM_ORDER}

    out = {"world": world, "surfels": P, "checks": {}, "times_us": {}}
    fails = []

    def rel(aa, bb):
        return float((aa - bb).abs().max() / bb.abs().max().clamp_min(1e-30))

    # ---- 1. plain path: local gradients, gathered sum vs NCCL all-reduce
    flat = torch.zeros(13 * P, device=dev)
    local = mapping.pack_grads(fwd_bwd(), flat).clone()
    gathered = [torch.empty_like(local) for _ in range(world)]
    dist.all_gather(gathered, local)
    want = torch.stack(gathered).sum(0)
    nccl = local.clone()
    dist.all_reduce(nccl)
    out["checks"]["nccl_vs_gathered_sum"] = rel(nccl, want)
    if out["checks"]["nccl_vs_gathered_sum"] > 1e-5:
        fails.append("NCCL all-reduce differs from the gathered sum")

    # ---- 2. the sum of the per-view gradients of the unmodified reference rasterizer
    try:
        import refgpu
        have_ref = refgpu.available()
    except Exception:
        have_ref = False
    if have_ref:
        ref = refgpu.RefRasterizer()
        kw = dict(bg=torch.zeros(3, device=dev), means3D=params["means3D"].detach(), viewmatrix=st.viewmatrix,
                  projmatrix=st.projmatrix, campos=st.campos, W=W, H=H, tanfovx=float(cam.tanfovx), tanfovy=float(cam.tanfovy),
                  colors=params["colors"].detach(), scales=params["scales"].detach(), rotations=params["rotations"].detach(),
                  use_sa=True)
        n, color, others, radii = ref.forward(opacities=params["opacities"].detach(), **kw)
        g = ref.backward(radii=radii, dL_dcolor=dcol, dL_dothers=doth, **kw)
        rflat = torch.cat([g[k].reshape(-1) for k in ("means3D", "opacity", "scales", "rotations", "colors")])
        dist.all_reduce(rflat)
        worst = 0.0
        for name, s0, s1 in mapping.flat_sections(P):
            worst = max(worst, rel(nccl[s0:s1], rflat[s0:s1]))
        out["checks"]["product_sum_vs_reference_sum_worst_group"] = worst
        if worst > 1e-3:
            fails.append(f"summed gradients differ from the reference's by {worst:.2e} (bar 1e-3)")
    else:
        out["checks"]["product_sum_vs_reference_sum_worst_group"] = None

    # ---- 3. own collectives: in-switch all-reduce kernel, and the fused backward
    fg = mapping.FusedGradient(P, dev)
    out["symmetric_memory"] = bool(fg.ok)
    if fg.ok:
        sym = fg.flat
        sym.copy_(local)
        mapping.allreduce_map_gradient(sym)
        out["checks"]["multimem_allreduce_vs_nccl"] = rel(sym, nccl)
        fg.begin()
        fwd_bwd()
        fused = fg.finish().clone()
        out["checks"]["fused_backward_vs_nccl"] = rel(fused, nccl)
        for k in ("multimem_allreduce_vs_nccl", "fused_backward_vs_nccl"):
            if out["checks"][k] > 1e-4:
                fails.append(f"{k} = {out['checks'][k]:.2e}")

    # ---- timing (max over ranks)
    def timeit(fn, iters):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        tm = torch.tensor([e0.elapsed_time(e1) / iters * 1e3], device=dev)
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        return float(tm)

    x = local.clone()
    out["times_us"]["nccl_allreduce_alone"] = timeit(lambda: dist.all_reduce(x), a.iters)

    def step_nccl():
        with mapping.gradient_arena(flat):
            mapping.allreduce_map_gradient(mapping.pack_grads(fwd_bwd(), flat))
    out["times_us"]["fwd_bwd_plus_nccl"] = timeit(step_nccl, a.iters)
    out["times_us"]["fwd_bwd_local_only"] = timeit(lambda: fwd_bwd(), a.iters)
    if fg.ok:
        out["times_us"]["multimem_allreduce_alone"] = timeit(lambda: mapping.allreduce_map_gradient(sym), a.iters)

        def step_multimem():
            with mapping.gradient_arena(sym):
                mapping.allreduce_map_gradient(mapping.pack_grads(fwd_bwd(), sym))

        def step_fused():
            fg.begin()
            fwd_bwd()
            fg.finish()
        out["times_us"]["fwd_bwd_plus_multimem"] = timeit(step_multimem, a.iters)
        out["times_us"]["fwd_bwd_fused"] = timeit(step_fused, a.iters)
    out["fails"] = fails
    if rank == 0:
        print(json.dumps(out))
        if a.out:
            os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
            with open(a.out, "w") as f:
                json.dump(out, f, indent=1)
    dist.destroy_process_group()
    return 1 if fails else 0


if __name__ == "__main__":
    sys.exit(main())
