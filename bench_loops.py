"""Whole-iteration loops around the rasterizer, both arms of bench.py (extra keys of its JSON line; the headline
`value` / `e2e` stay the rasterizer's fwd+bwd):

  mapping_iteration  -- Frontend.mapping's body (slam/Frontend.py:116-128): zero_grad, render of one keyframe through
                        get_render_params' activations, mapping loss, backward, map optimiser step.
                        product: gs2d_b200.optim.MapOptimizer (activation chain rule + Adam + next activations = one
                        launch, gradients written by the backward pass straight into its flat buffer) + fused loss;
                        reference: torch activations, the reference rasterizer, the reference's loss as torch ops
                        (CUDA-graphed), torch.optim.Adam(eps=1e-15) -- its stock code path.
  tracking_iteration -- Frontend.tracking's body (:80-92; BASELINE.json configs[1]: 300k-surfel submap, pose only):
                        pose -> camera-frame map (render/__init__.py:35-36) -> render with the identity camera ->
                        tracking loss (slam/Loss.py:35-47, ignore_outliners False, silmask 0.9) -> backward -> Adam on
                        the 7 pose scalars (scene/Frame.py:70-73).
                        product: gs2d_b200.tracking.pose_transform (one kernel each way); reference: the reference's
                        torch expressions with the restated pytorch3d algebra (oracle/pose_ref.py).
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
TRAINING_ARGS = dict(xyz_lr=1e-4, opacity_lr=5e-2, scaling_lr=1e-3, rotation_lr=1e-3, rgb_lr=2.5e-3)   # configs/replica/config.py:125-132
GROUPS = ("xyz", "opacity", "scaling", "rotation", "rgb")


def raw_from_scene(sc, torch, dev):
    """Raw optimiser parameters whose activations are the scene's render parameters."""
    o = np.clip(sc["opacities"].astype(np.float64), 1e-6, 1 - 1e-6)
    raw = dict(xyz=sc["means3D"], opacity=np.log(o / (1 - o)), scaling=np.log(sc["scales"].astype(np.float64)),
               rotation=sc["rotations"] * 1.5, rgb=sc["colors"])
    return {k: torch.tensor(np.ascontiguousarray(v, dtype=np.float32), device=dev) for k, v in raw.items()}


def mapping_iteration(arm, wl, dev, timed, steps, warmup, view_of, loss_fn, impl, multi, allreduce):
    import torch
    raw = raw_from_scene(wl["scene"], torch, dev)

    def render(rp, s):
        means2D = torch.zeros_like(rp["means3D"], requires_grad=True)
        rast = arm.rasterizer(s)
        return rast(means3D=rp["means3D"], means2D=means2D, opacities=rp["opacities"],
                    colors_precomp=rp["colors_precomp"], scales=rp["scales"], rotations=rp["rotations"])

    if impl == "b200":
        from gs2d_b200 import optim
        from gaus_2dgs_rasterization import _C
        opt = optim.MapOptimizer(*[raw[k] for k in GROUPS], TRAINING_ARGS)
        _C.set_gradient_arena(opt.flat_grad)

        def step(i):
            s = view_of(i)
            color, radii, allmap = render(opt.render_params(), s)
            loss = loss_fn(color, allmap, s["gt_color"], s["gt_depth"])
            loss.backward()
            if multi:
                allreduce(opt.flat_grad)
            opt.step()
            return loss
        how = "gs2d_b200.optim.MapOptimizer (fused activation chain rule + Adam + re-activation) + fused loss"
        cleanup = lambda: _C.set_gradient_arena(None)     # noqa: E731
    else:
        p = {k: torch.nn.Parameter(v.clone()) for k, v in raw.items()}
        lrk = dict(xyz="xyz_lr", opacity="opacity_lr", scaling="scaling_lr", rotation="rotation_lr", rgb="rgb_lr")
        optimizer = torch.optim.Adam([{"params": [p[k]], "lr": TRAINING_ARGS[lrk[k]], "name": k} for k in GROUPS],
                                     lr=0.0, eps=1e-15)                                        # Gaussians.py:124-137

        def step(i):
            s = view_of(i)
            optimizer.zero_grad(set_to_none=True)
            rp = dict(means3D=p["xyz"], opacities=torch.sigmoid(p["opacity"]), scales=torch.exp(p["scaling"]),
                      rotations=torch.nn.functional.normalize(p["rotation"]), colors_precomp=p["rgb"])
            color, radii, allmap = render(rp, s)
            loss = loss_fn(color, allmap, s["gt_color"], s["gt_depth"])
            loss.backward()
            with torch.no_grad():
                optimizer.step()
            return loss
        how = "torch activations + reference rasterizer + torch-ops loss (CUDA graphs) + torch.optim.Adam(eps=1e-15)"
        cleanup = lambda: None     # noqa: E731

    try:
        for i in range(warmup):
            step(i)
        last = {}

        def run(i):
            last["loss"] = step(i)
        ms = timed(run, steps, warmup) / steps
        loss = float(last["loss"].detach())
    finally:
        cleanup()
    return {"value": 1000.0 / ms, "unit": "iterations/s", "ms_per_iteration": ms, "final_loss": loss,
            "what": "render -> mapping loss -> backward -> map optimiser step over 5 parameter groups "
                    "(slam/Frontend.py:116-128), inputs resident; " + how}


def tracking_loss(color, allmap, gt_color, gt_depth):
    """slam/Loss.py:22-47 (tracking branch, ignore_outliners False) after render/__init__.py:46-49; masks as
    multiplications (same value, no host synchronisation)."""
    import torch
    depth = allmap[0:1] / (allmap[1:2] + 1e-6)
    depth = torch.where((depth > 1e2) | (depth < 1e-2), torch.zeros_like(depth), depth)
    depth = torch.nan_to_num(depth, 0, 0)
    alpha = torch.nan_to_num(allmap[1:2], 0, 0)
    mask = ((gt_depth > 1e-5) & (depth > 1e-5) & (alpha > 0.9)).to(color.dtype)
    l_color = ((torch.nan_to_num(color, 0, 0) - gt_color).abs() * mask).sum()
    l_depth = ((depth - gt_depth).abs() * mask).sum()
    return 0.5 * l_color + 1.0 * l_depth


def tracking_iteration(arm, dev, timed, steps, warmup, impl, surfels=300_000):
    import torch
    sys.path.insert(0, os.path.join(ROOT, "gaus-slam_b200"))
    from gs2d_b200 import scene
    intr = scene.REPLICA
    W, H = intr["W"], intr["H"]
    sc = scene.room_scene(surfels, intr, seed=1)
    cam = scene.make_camera(np.eye(4, dtype=np.float32), **intr)       # Renderer_tracking: identity camera
    t = lambda x: torch.tensor(np.ascontiguousarray(x, dtype=np.float32), device=dev)     # noqa: E731
    xyz, rot, opa, scl, rgb = (t(sc[k]) for k in ("means3D", "rotations", "opacities", "scales", "colors"))
    rng = np.random.default_rng(9)
    s = dict(viewmatrix=t(cam.viewmatrix.reshape(16)), projmatrix=t(cam.projmatrix.reshape(16)), campos=t(cam.campos),
             tanfovx=float(cam.tanfovx), tanfovy=float(cam.tanfovy),
             gt_color=t(rng.uniform(0, 1, (3, H, W))), gt_depth=t(rng.uniform(1.0, 4.0, (1, H, W))))
    cam_rot = torch.nn.Parameter(t([0.9995, 0.01, -0.02, 0.015]))       # a pose a few cm / degrees off
    cam_trans = torch.nn.Parameter(t([0.03, -0.02, 0.05]))
    optimizer = torch.optim.Adam([{"params": [cam_rot], "lr": 4e-4, "name": "cam_rots"},
                                  {"params": [cam_trans], "lr": 2e-3, "name": "cam_trans"}],
                                 lr=0.0, eps=1e-8, betas=(0.7, 0.99))                     # scene/Frame.py:70-73
    if impl == "b200":
        from gs2d_b200 import tracking

        def to_camera():
            return tracking.pose_transform(xyz, rot, cam_rot=cam_rot, cam_trans=cam_trans)
        from gs2d_b200 import loss as gs2d_loss
        loss_fn = gs2d_loss.tracking_loss
        # the tracker's Adam over the seven pose scalars (scene/Frame.py:70-73) as one launch per step
        optimizer = tracking.PoseOptimizer(cam_rot, cam_trans, lr_rot=4e-4, lr_trans=2e-3, betas=(0.7, 0.99), eps=1e-8)
        how = ("gs2d_b200.tracking.pose_transform (one kernel each way, pose chain rule inside) + fused tracking loss + "
               "gs2d_b200.tracking.PoseOptimizer (torch.optim.Adam's update of the 7 scalars as one launch)")
    else:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import pose_ref

        def to_camera():
            w2c = pose_ref.get_transform_matrix(cam_rot, cam_trans)
            return pose_ref.pose_transform_torch(xyz, rot, w2c)
        loss_fn = tracking_loss
        how = "the reference's torch expressions (render/__init__.py:35-36, scene/Frame.py:87-96; pytorch3d algebra restated) + tracking loss as eager torch ops"
    rast = arm.rasterizer(s)

    def step(i):
        optimizer.zero_grad(set_to_none=True)
        means_c, rots_c = to_camera()
        means2D = torch.zeros_like(xyz, requires_grad=True)
        color, radii, allmap = rast(means3D=means_c, means2D=means2D, opacities=opa, colors_precomp=rgb, scales=scl,
                                    rotations=rots_c)
        loss = loss_fn(color, allmap, s["gt_color"], s["gt_depth"])
        loss.backward()
        with torch.no_grad():
            optimizer.step()
        return loss

    for i in range(warmup):
        step(i)
    last = {}

    def run(i):
        last["loss"] = step(i)
    ms = timed(run, steps, warmup) / steps
    return {"value": 1000.0 / ms, "unit": "iterations/s", "ms_per_iteration": ms, "surfels": surfels,
            "frames_per_s_at_40_iters": 1000.0 / ms / 40.0, "final_loss": float(last["loss"].detach()),
            "what": "pose -> camera-frame map -> render (identity camera) -> tracking loss -> backward -> Adam on 7 pose scalars (slam/Frontend.py:80-92, BASELINE.json configs[1]); " + how}


def scannetpp_frame(arm, dev, timed, steps, warmup, surfels=4_000_000, views=4):
    """BASELINE.json configs[4] in shape: fwd + bwd of one 1752 x 1168 view of a 4 M-surfel map per step (8030 tiles,
    45 sort bits), through the arm's public API, inputs resident, views rotating over `views` per-view input sets."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "gaus-slam_b200"))
    from gs2d_b200 import scene
    intr = scene.SCANNETPP
    W, H = intr["W"], intr["H"]
    sc = scene.room_scene(surfels, intr, seed=4)
    cams = scene.view_sweep(8, intr, seed=4)[:views]
    t = lambda x: torch.tensor(np.ascontiguousarray(x, dtype=np.float32), device=dev)     # noqa: E731
    params = {k: t(sc[k]).requires_grad_(True) for k in ("means3D", "opacities", "scales", "rotations", "colors")}
    means2D = torch.zeros((surfels, 3), device=dev, requires_grad=True)
    sets = []
    for v, cam in enumerate(cams):
        gc, go = scene.image_grads(W, H, seed=v)
        s = dict(viewmatrix=t(cam.viewmatrix.reshape(16)), projmatrix=t(cam.projmatrix.reshape(16)), campos=t(cam.campos),
                 tanfovx=float(cam.tanfovx), tanfovy=float(cam.tanfovy), dL_dcolor=t(gc), dL_dothers=t(go))
        s["rast"] = _rasterizer(arm, s, W, H, dev)
        sets.append(s)

    def step(i):
        s = sets[i % views]
        for p in params.values():
            p.grad = None
        means2D.grad = None
        color, radii, allmap = s["rast"](means3D=params["means3D"], means2D=means2D, opacities=params["opacities"],
                                         colors_precomp=params["colors"], scales=params["scales"],
                                         rotations=params["rotations"])
        torch.autograd.backward([color, allmap], [s["dL_dcolor"], s["dL_dothers"]])
    for i in range(warmup):
        step(i)
    ms = timed(step, steps, warmup) / steps
    return {"value": 1000.0 / ms, "unit": "frames/s", "ms_per_step": ms, "surfels": surfels, "image": f"{W}x{H}",
            "what": f"fwd + bwd of one {W}x{H} view of a {surfels}-surfel map per step (BASELINE.json configs[4] in shape: "
                    f"8030 tiles, 45 sort bits), public API, inputs resident, {views} views rotating"}


def _rasterizer(arm, s, W, H, dev):
    import torch
    G = getattr(arm, "G", None)
    if G is None:                       # the C-shim fallback arm builds its own autograd node
        return arm.rasterizer(s)
    st = G.GaussianRasterizationSettings(image_height=H, image_width=W, tanfovx=s["tanfovx"], tanfovy=s["tanfovy"],
                                         bg=torch.zeros(3, device=dev), scale_modifier=1.0, viewmatrix=s["viewmatrix"],
                                         projmatrix=s["projmatrix"], sh_degree=0, campos=s["campos"], use_sa=True,
                                         prefiltered=False, debug=False)
    return G.GaussianRasterizer(st)


def ba_step(arm, dev, timed, steps, warmup, impl, rank=0, world=1, surfels=2_000_000, keyframes=64):
    """BASELINE.json configs[3] in shape: ONE joint pose + map gradient step of the backend's bundle adjustment over a
    window of 64 keyframes of a 2 M-surfel map (render/__init__.py:84-111 Renderer_BA per keyframe: means3D' = R x + t with
    the map AND the keyframe's pose attached, rotations detached; slam/Backend.py:130-155).  The keyframes are sharded
    over the ranks; every rank accumulates the map gradient of its keyframes IN PLACE in one flat [13 P] buffer
    (param.grad are views of it) and keeps its keyframes' pose gradients; the buffer (104 MB) goes through one all-reduce
    per step.  value = keyframes per second."""
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "gaus-slam_b200"))
    from gs2d_b200 import mapping, scene
    intr = scene.REPLICA
    W, H, P = intr["W"], intr["H"], surfels
    sc = scene.room_scene(P, intr, seed=2)
    cams = scene.view_sweep(keyframes, intr, seed=2)
    mine = [k for k in range(keyframes) if k % world == rank]
    t = lambda x: torch.tensor(np.ascontiguousarray(x, dtype=np.float32), device=dev)     # noqa: E731
    params = {k: t(sc[k]).requires_grad_(True) for k in mapping.PARAM_ORDER}
    flat = torch.zeros(13 * P, device=dev)
    views_of_flat = mapping.unpack_grads(flat, P)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pose_ref
    ident = scene.make_camera(np.eye(4, dtype=np.float32), **intr)       # Renderer_BA renders with the identity camera
    s_id = dict(viewmatrix=t(ident.viewmatrix.reshape(16)), projmatrix=t(ident.projmatrix.reshape(16)), campos=t(ident.campos),
                tanfovx=float(ident.tanfovx), tanfovy=float(ident.tanfovy))
    rast = _rasterizer(arm, s_id, W, H, dev)
    kfs = []
    for k in mine:
        w2c = cams[k].w2c.astype(np.float32)
        q = pose_ref.matrix_to_quaternion(torch.tensor(w2c[None, :3, :3], dtype=torch.float64))[0].numpy()
        gc, go = scene.image_grads(W, H, seed=k % 8)
        kfs.append(dict(cam_rot=t(q).requires_grad_(True), cam_trans=t(w2c[:3, 3]).requires_grad_(True), dL_dcolor=t(gc),
                        dL_dothers=t(go)))
    if impl == "b200":
        from gs2d_b200 import tracking

        def to_camera(kf):
            return tracking.pose_transform(params["means3D"], params["rotations"], cam_rot=kf["cam_rot"], cam_trans=kf["cam_trans"])
        how = "gs2d_b200.tracking.pose_transform + the drop-in rasterizer"
    else:
        def to_camera(kf):
            return pose_ref.pose_transform_torch(params["means3D"], params["rotations"],
                                                 pose_ref.get_transform_matrix(kf["cam_rot"], kf["cam_trans"]))
        how = "the reference's torch expressions (pytorch3d algebra restated) + the reference rasterizer"
    means2D = torch.zeros((P, 3), device=dev, requires_grad=True)

    def step(i):
        flat.zero_()
        for k in mapping.PARAM_ORDER:            # autograd accumulates the keyframes' gradients in place in `flat`
            params[k].grad = views_of_flat[k] if k != "rotations" else None
        means2D.grad = None
        for kf in kfs:
            kf["cam_rot"].grad = None
            kf["cam_trans"].grad = None
            means_c, rots_c = to_camera(kf)
            color, radii, allmap = rast(means3D=means_c, means2D=means2D, opacities=params["opacities"],
                                        colors_precomp=params["colors"], scales=params["scales"], rotations=rots_c)
            torch.autograd.backward([color, allmap], [kf["dL_dcolor"], kf["dL_dothers"]])
        if world > 1:
            dist.all_reduce(flat)
    for i in range(warmup):
        step(i)
    in_place = params["means3D"].grad.data_ptr() == flat.data_ptr()
    ms = timed(step, steps, warmup) / steps
    return {"value": keyframes * 1000.0 / ms, "unit": "keyframes/s", "ms_per_step": ms, "surfels": P, "keyframes": keyframes,
            "keyframes_per_rank": len(mine), "allreduce_bytes": int(flat.numel() * 4) if world > 1 else 0,
            "map_gradient_accumulated_in_place": bool(in_place),
            "what": f"one joint pose + map gradient step over a {keyframes}-keyframe window of a {P}-surfel map (BASELINE.json "
                    f"configs[3] in shape), keyframes sharded over {world} rank(s), one all-reduce of the flat [13 P] buffer per "
                    f"step; " + how}
