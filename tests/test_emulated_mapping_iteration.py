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
    case = cases.room_sa()
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
