"""CPU: one whole mapping iteration of the product, wired the way INTEGRATION.md / bench_loops.py wire it --
MapOptimizer.render_params() -> GaussianRasterizer (backward writing into the optimiser's flat [13 P] buffer through
the gradient arena) -> fused mapping loss -> loss.backward() -> MapOptimizer.step() -- with every native piece replaced
by the host build of its own source text (tests/emu).  Reference pipeline beside it: the reference's parameter handling
and optimiser (oracle/optim_ref.MapTorch = torch activations + torch.optim.Adam), the C restatement of the rasterizer
behind a torch.autograd.Function, the torch transliteration of get_loss.  Checks the WIRING: gradients land in the right
sections of the flat buffer without a copy, the loss gradient reaches the rasterizer, the step consumes what the
backward pass wrote."""
import os
import shutil
import sys

import numpy as np
import pytest
import torch

import cases
import compare
import cpu as O
import loss_ref
import optim_ref as R

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))

pytestmark = pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++ (C++20) for the host build of the kernels")
ARGS = dict(xyz_lr=1e-4, opacity_lr=5e-2, scaling_lr=1e-3, rotation_lr=1e-3, rgb_lr=2.5e-3)


class _OracleRasterizer(torch.autograd.Function):
    """The C restatement of the reference rasterizer as an autograd node (CPU)."""

    @staticmethod
    def forward(ctx, means3D, opacities, scales, rotations, colors, case):
        a = (means3D.numpy(), opacities.numpy(), case["viewmatrix"], case["projmatrix"], case["W"], case["H"], case["bg"])
        kw = dict(scales=scales.numpy(), rotations=rotations.numpy(), campos=case["campos"])
        st = O.forward(*a, colors=colors.numpy(), scale_modifier=1.0, use_sa=True, **kw)
        ctx.st, ctx.case, ctx.kw, ctx.means3D = st, case, kw, means3D.numpy()
        return torch.tensor(np.asarray(st["color"])), torch.tensor(np.asarray(st["others"]))

    @staticmethod
    def backward(ctx, g_color, g_others):
        c = ctx.case
        g = O.backward(ctx.st, ctx.means3D, c["viewmatrix"], c["projmatrix"], c["W"], c["H"], float(c["tanfovx"]),
                       float(c["tanfovy"]), c["bg"], g_color.contiguous().numpy(), g_others.contiguous().numpy(), use_sa=True, **ctx.kw)
        t = lambda k: torch.tensor(np.asarray(g[k]))      # noqa: E731
        return t("means3D"), t("opacity"), t("scales"), t("rotations"), t("colors"), None


def test_one_mapping_iteration_end_to_end():
    import host_only
    G, optim, loss = host_only.load_rasterizer(), host_only.load("optim"), host_only.load("loss")
    case = cases._base(1200, cases._intr(96, 64), seed=1, view=1)          # room_sa's view at a size the emulator walks in seconds
    P, W, H = case["means3D"].shape[0], case["W"], case["H"]
    rng = np.random.default_rng(5)
    raw = dict(xyz=case["means3D"], opacity=np.log(case["opacities"] / (1 - case["opacities"])).astype(np.float32),
               scaling=np.log(case["scales"]).astype(np.float32),
               rotation=(case["rotations"] * rng.uniform(0.5, 2.0, (P, 1))).astype(np.float32), rgb=case["colors"])
    gt_color = rng.uniform(0, 1, (H, W, 3)).astype(np.float32)
    gt_depth = rng.uniform(0.5, 4.0, (H, W, 1)).astype(np.float32)
    t = lambda a: torch.tensor(np.ascontiguousarray(a, np.float32))      # noqa: E731

    # ---- product wiring --------------------------------------------------------------------------------------------
    opt = optim.MapOptimizer(*[t(raw[k]) for k in R.GROUPS], ARGS)
    rs = G.GaussianRasterizationSettings(image_height=H, image_width=W, tanfovx=float(case["tanfovx"]), tanfovy=float(case["tanfovy"]),
                                         bg=t(case["bg"]), scale_modifier=1.0, viewmatrix=t(case["viewmatrix"]).view(4, 4),
                                         projmatrix=t(case["projmatrix"]).view(4, 4), sh_degree=0, campos=t(case["campos"]),
                                         use_sa=True, prefiltered=False, debug=False)
    G._C.set_gradient_arena(opt.flat_grad)
    try:
        rp = opt.render_params()
        color, radii, allmap = G.GaussianRasterizer(rs)(means2D=torch.zeros((P, 3), requires_grad=True), shs=None,
                                                        cov3D_precomp=None, **rp)
        value = loss.mapping_loss(color, allmap, t(gt_color), t(gt_depth))
        value.backward()
    finally:
        G._C.set_gradient_arena(None)
    # zero copy: every leaf's gradient IS its section of the flat buffer
    base, off = opt.flat_grad.data_ptr(), 0
    for name, w in (("means3D", 3), ("opacities", 1), ("scales", 2), ("rotations", 4), ("colors_precomp", 3)):
        assert rp[name].grad.data_ptr() == base + 4 * off, name
        off += w * P
    flat = opt.flat_grad.numpy().copy()
    before = {k: opt.raw[k].clone() for k in R.GROUPS}
    opt.step()

    # ---- reference pipeline ------------------------------------------------------------------------------------------
    ref = R.MapTorch(raw, ARGS)
    rrp = ref.render_params()
    rc, ro = _OracleRasterizer.apply(rrp["means3D"], rrp["opacities"], rrp["scales"], rrp["rotations"], rrp["colors"], case)
    rv = loss_ref.mapping_loss_torch(rc, ro, t(gt_color), t(gt_depth))
    assert abs(float(value.detach()) - float(rv.detach())) <= 1e-4 * abs(float(rv.detach()))
    grads = torch.autograd.grad(rv, [rrp[k] for k, _ in R.FLAT])
    got = R.split_flat(flat, P)
    for (k, _), g in zip(R.FLAT, grads):
        r, e = compare.grad_rel(got[k], g.numpy()), compare.grad_elem_rel(got[k], g.numpy())
        assert r <= compare.ORACLE_SA_GRAD_RTOL and e <= 3 * compare.ORACLE_SA_GRAD_RTOL, (k, r, e)
    # the step took exactly what the backward pass wrote: the reference optimiser fed the same flat gradient agrees
    ref.step(flat)
    st = ref.state()
    for k in R.GROUPS:
        assert float((opt.raw[k] - before[k]).abs().max()) > 0, k
        assert np.abs(opt.raw[k].numpy() - st[k]).max() <= 2e-6 * max(np.abs(st[k]).max(), 1e-30), k


def test_one_tracking_iteration_end_to_end():
    """Frontend.tracking's body (render/__init__.py:22-50, slam/Loss.py:35-47): pose scalars -> fused pose transform ->
    rasterizer with an identity camera -> fused tracking loss -> backward down to dL/dcam_rot, dL/dcam_trans.
    Reference pipeline: the torch transliteration of the pose path, the C restatement of the rasterizer, the
    transliterated tracking loss."""
    import host_only
    import pose_ref
    G, tracking, loss = host_only.load_rasterizer(), host_only.load("tracking"), host_only.load("loss")
    case = cases._base(1200, cases._intr(96, 64), seed=1, view=1)          # room_sa's view at a size the emulator walks in seconds
    P, W, H = case["means3D"].shape[0], case["W"], case["H"]
    rng = np.random.default_rng(11)
    gt_color = rng.uniform(0, 1, (H, W, 3)).astype(np.float32)
    gt_depth = rng.uniform(0.5, 4.0, (H, W, 1)).astype(np.float32)
    t = lambda a, rg=False: torch.tensor(np.ascontiguousarray(a, np.float32), requires_grad=rg)      # noqa: E731
    # the golden view as the pose: w2c = viewmatrix^T (column-major storage), camera at identity afterwards
    w2c = np.ascontiguousarray(case["viewmatrix"].reshape(4, 4).T)
    q0 = pose_ref.matrix_to_quaternion(torch.tensor(w2c[None, :3, :3], dtype=torch.float64))[0].numpy() * 1.3   # unnormalised
    t0 = w2c[:3, 3]
    ident = dict(case, viewmatrix=np.eye(4, dtype=np.float32).reshape(16), campos=np.zeros(3, np.float32))
    proj = case["projmatrix"].reshape(4, 4).T @ np.linalg.inv(w2c.astype(np.float64))              # P V -> P (identity view)
    ident["projmatrix"] = np.ascontiguousarray(proj.T.astype(np.float32)).reshape(16)
    xyz, rot = t(case["means3D"]), t(case["rotations"])

    q, tr = t(q0, True), t(t0, True)
    means, rots = tracking.pose_transform(xyz, rot, cam_rot=q, cam_trans=tr)
    rs = G.GaussianRasterizationSettings(image_height=H, image_width=W, tanfovx=float(case["tanfovx"]), tanfovy=float(case["tanfovy"]),
                                         bg=t(case["bg"]), scale_modifier=1.0, viewmatrix=t(ident["viewmatrix"]).view(4, 4),
                                         projmatrix=t(ident["projmatrix"]).view(4, 4), sh_degree=0, campos=t(ident["campos"]),
                                         use_sa=True, prefiltered=False, debug=False)
    color, radii, allmap = G.GaussianRasterizer(rs)(means3D=means, means2D=torch.zeros((P, 3), requires_grad=True),
                                                    opacities=t(case["opacities"]), colors_precomp=t(case["colors"]),
                                                    scales=t(case["scales"]), rotations=rots)
    value = loss.tracking_loss(color, allmap, t(gt_color), t(gt_depth), silmask_th=0.5)
    value.backward()

    rq, rt = t(q0, True), t(t0, True)
    rm, rr = pose_ref.pose_transform_torch(xyz, rot, pose_ref.get_transform_matrix(rq, rt))
    rc, ro = _OracleRasterizer.apply(rm, t(case["opacities"]), t(case["scales"]), rr, t(case["colors"]), ident)
    rv = loss_ref.tracking_loss_torch(rc, ro, t(gt_color), t(gt_depth), silmask_th=0.5)
    rv.backward()
    assert int((radii > 0).sum()) > P // 2                                  # the scene is in view under the pose
    assert abs(float(value.detach()) - float(rv.detach())) <= 1e-3 * abs(float(rv.detach()))
    for got, want in ((q.grad, rq.grad), (tr.grad, rt.grad)):
        scale = float(want.abs().max())
        assert scale > 0 and float((got - want).abs().max()) <= 3 * compare.ORACLE_SA_GRAD_RTOL * scale


def test_one_bundle_adjustment_step_end_to_end():
    """Renderer_BA's shape (render/__init__.py:84-111): map AND pose get gradients through the fused pose transform,
    rotations are detached, so the optimiser step skips the rotation group as torch.optim.Adam skips .grad = None."""
    import host_only
    import pose_ref
    G, tracking, loss, optim = (host_only.load_rasterizer(), host_only.load("tracking"), host_only.load("loss"),
                                host_only.load("optim"))
    case = cases._base(1200, cases._intr(96, 64), seed=1, view=1)
    P, W, H = case["means3D"].shape[0], case["W"], case["H"]
    rng = np.random.default_rng(23)
    gt_color = rng.uniform(0, 1, (H, W, 3)).astype(np.float32)
    gt_depth = rng.uniform(0.5, 4.0, (H, W, 1)).astype(np.float32)
    t = lambda a, rg=False: torch.tensor(np.ascontiguousarray(a, np.float32), requires_grad=rg)      # noqa: E731
    raw = dict(xyz=case["means3D"], opacity=np.log(case["opacities"] / (1 - case["opacities"])).astype(np.float32),
               scaling=np.log(case["scales"]).astype(np.float32), rotation=case["rotations"], rgb=case["colors"])
    w2c = np.ascontiguousarray(case["viewmatrix"].reshape(4, 4).T)
    q0 = pose_ref.matrix_to_quaternion(torch.tensor(w2c[None, :3, :3], dtype=torch.float64))[0].numpy()
    t0 = w2c[:3, 3]
    ident = dict(case, viewmatrix=np.eye(4, dtype=np.float32).reshape(16), campos=np.zeros(3, np.float32))
    proj = case["projmatrix"].reshape(4, 4).T @ np.linalg.inv(w2c.astype(np.float64))
    ident["projmatrix"] = np.ascontiguousarray(proj.T.astype(np.float32)).reshape(16)

    opt = optim.MapOptimizer(*[t(raw[k]) for k in R.GROUPS], ARGS)
    rp = opt.render_params()
    q, tr = t(q0, True), t(t0, True)
    means, rots = tracking.pose_transform(rp["means3D"], rp["rotations"], cam_rot=q, cam_trans=tr)
    rs = G.GaussianRasterizationSettings(image_height=H, image_width=W, tanfovx=float(case["tanfovx"]), tanfovy=float(case["tanfovy"]),
                                         bg=t(case["bg"]), scale_modifier=1.0, viewmatrix=t(ident["viewmatrix"]).view(4, 4),
                                         projmatrix=t(ident["projmatrix"]).view(4, 4), sh_degree=0, campos=t(ident["campos"]),
                                         use_sa=True, prefiltered=False, debug=False)
    color, radii, allmap = G.GaussianRasterizer(rs)(means3D=means, means2D=torch.zeros((P, 3), requires_grad=True),
                                                    opacities=rp["opacities"], colors_precomp=rp["colors_precomp"],
                                                    scales=rp["scales"], rotations=rots)
    loss.mapping_loss(color, allmap, t(gt_color), t(gt_depth)).backward()
    assert rp["rotations"].grad is None                                   # detached inside pose_transform, like the reference
    flat = torch.cat([rp["means3D"].grad.reshape(-1), rp["opacities"].grad.reshape(-1), rp["scales"].grad.reshape(-1),
                      torch.zeros(4 * P), rp["colors_precomp"].grad.reshape(-1)])
    rot_before = opt.raw["rotation"].clone()
    opt.step(flat, skip=("rotation",))
    assert torch.equal(opt.raw["rotation"], rot_before) and not opt.exp_avg["rotation"].any()

    ref = R.MapTorch(raw, ARGS)
    rrp = ref.render_params()
    rq, rt = t(q0, True), t(t0, True)
    rm, rr = pose_ref.pose_transform_torch(rrp["means3D"], rrp["rotations"], pose_ref.get_transform_matrix(rq, rt))
    rc, ro = _OracleRasterizer.apply(rm, rrp["opacities"], rrp["scales"], rr, rrp["colors"], ident)
    loss_ref.mapping_loss_torch(rc, ro, t(gt_color), t(gt_depth)).backward()
    assert ref.p["rotation"].grad is None
    got = R.split_flat(flat.numpy(), P)
    # gradients w.r.t. the RAW parameters on the reference side -> compare where raw = render (xyz, rgb) and the poses
    for k, leaf in (("means3D", ref.p["xyz"]), ("colors", ref.p["rgb"])):
        r = compare.grad_rel(got[k], leaf.grad.numpy())
        assert r <= compare.ORACLE_SA_GRAD_RTOL, (k, r)
    for a, b in ((q.grad, rq.grad), (tr.grad, rt.grad)):
        assert float((a - b).abs().max()) <= 3 * compare.ORACLE_SA_GRAD_RTOL * float(b.abs().max())
    ref.step(flat.numpy(), detached=("rotations",))
    st = ref.state()
    for k in R.GROUPS:
        assert np.abs(opt.raw[k].numpy() - st[k]).max() <= 2e-6 * max(np.abs(st[k]).max(), 1e-30), k
