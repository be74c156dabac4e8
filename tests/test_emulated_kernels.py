"""CPU: rows f1-f3 of SURVEY.md 8(f) END TO END without a GPU -- the product's Python wrappers (gs2d_b200.optim / loss /
tracking, loaded host-only by tests/emu/host_only.py) over the SOURCE TEXT of their CUDA files (csrc/optim.cu,
compact.cu, loss.cu, pose.cu) built for the host under a CUDA-thread emulator (tests/emu/build_emu.py, cuda_emu.h),
held to the golden vectors of the reference's own code (tests/golden/optim_*.npz, loss_*.npz) and to the torch
transliteration of the pose path.  Same asserts and tolerances as the GPU tests of the same rows.

This checks the kernels' arithmetic, indexing, reductions and the Python around them; it is a pre-check for kernel
edits in a container with no GPU, not a replacement for the GPU tests (host libm vs device math, no memory model)."""
import os
import shutil
import sys

import numpy as np
import pytest
import torch

import optim_ref as R
import test_loss as TL
import test_optim as TO
import test_pose as TP
import test_zz_map_compaction as TS

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))
import host_only  # noqa: E402

pytestmark = pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++ (C++20) for the host build of the kernels")


@pytest.fixture(scope="module")
def optim():
    return host_only.load("optim")


@pytest.fixture(scope="module")
def loss():
    return host_only.load("loss")


@pytest.fixture(scope="module")
def tracking():
    return host_only.load("tracking")


def test_emulated_library_exports_the_next_row_entry_points():
    lib = host_only.emulated_library()
    for name in ("gs2d_map_activate", "gs2d_map_adam_step", "gs2d_map_compact_scratch_bytes", "gs2d_map_compact_plan",
                 "gs2d_map_compact_move", "gs2d_mapping_loss_scratch_bytes", "gs2d_mapping_loss_forward",
                 "gs2d_mapping_loss_backward", "gs2d_pose_transform_forward", "gs2d_pose_transform_scratch_bytes",
                 "gs2d_pose_transform_backward", "gs2d_pose_adam_step"):
        assert hasattr(lib, name), name


# ---- f3: optimiser step and densification surgery ---------------------------------------------------------------------
@pytest.mark.parametrize("name", TO.GOLDEN_CASES)
def test_map_step_kernel_text_matches_the_reference_gaussians_class(optim, name):
    g, raw, args = TO._golden(name)
    opt = optim.MapOptimizer(*[torch.tensor(raw[k]) for k in R.GROUPS], args)
    for step in range(1, g["in_grad"].shape[0] + 1):
        opt.flat_grad.copy_(torch.from_numpy(g["in_grad"][step - 1]))
        opt.step()
        for k in R.GROUPS:
            assert TO._rel(opt.raw[k].numpy(), g[f"ref_{k}_{step}"].astype(np.float64)) <= 2e-6, (step, k)
            assert TO._rel(opt.exp_avg[k].numpy(), g[f"ref_m_{k}_{step}"].astype(np.float64)) <= 1e-5, (step, k)
            assert TO._rel(opt.exp_avg_sq[k].numpy(), g[f"ref_v_{k}_{step}"].astype(np.float64)) <= 1e-5, (step, k)
        for k in ("opacities", "scales", "rotations"):
            assert TO._rel(getattr(opt, k).numpy(), g[f"ref_{k}_{step}"].astype(np.float64)) <= 2e-6, (step, k)
    last = g["in_grad"].shape[0]
    moved, want = opt.raw["opacity"].numpy() - raw["opacity"], g[f"ref_opacity_{last}"] - raw["opacity"]
    assert np.abs(moved - want).max() <= 1e-4 * np.abs(want).max() + 1e-7


def test_surgery_through_wrappers_and_kernel_text(optim):
    g, raw, args = TS._golden()
    opt = optim.MapOptimizer(*[torch.tensor(raw[k]) for k in R.GROUPS], args)

    def step(tag):
        opt.step(torch.tensor(g[f"in_grad_{tag}"]))

    step("1"); TS._check(opt, g, "1", False)
    step("2"); TS._check(opt, g, "2", False)
    keep = torch.tensor(~g["in_remove_mask"])
    before = {k: (opt.raw[k].clone(), opt.exp_avg[k].clone(), opt.exp_avg_sq[k].clone()) for k in R.GROUPS}
    n = opt.prune(keep)
    assert n == int(keep.sum()) == opt.P and opt.flat_grad.numel() == 13 * n and opt.step_count == 2
    for k in R.GROUPS:
        for got, old in zip((opt.raw[k], opt.exp_avg[k], opt.exp_avg_sq[k]), before[k]):
            assert torch.equal(got, old[keep]), k                      # tensor[mask] semantics, bit for bit
    TS._check(opt, g, "pruned", False)
    step("3"); TS._check(opt, g, "3", False)
    P0 = opt.P
    opt.extend(*[torch.tensor(g["in_new_" + k]) for k in R.GROUPS])
    assert opt.P == P0 + g["in_new_xyz"].shape[0] and opt.step_count == 3
    TS._check(opt, g, "extended", False)
    step("4"); TS._check(opt, g, "4", False)
    opt.replace("opacity", torch.tensor(g["ref_opacity_reset"]))
    TS._check(opt, g, "reset", False)
    step("5"); TS._check(opt, g, "5", False)
    rp = opt.render_params()
    assert rp["means3D"].data_ptr() == opt.raw["xyz"].data_ptr() and rp["scales"].shape == (opt.P, 2)
    n = opt.prune(torch.zeros(opt.P, dtype=torch.bool))                   # nothing kept
    assert n == 0 and opt.P == 0 and opt.scales.shape == (0, 2) and opt.flat_grad.numel() == 0


# ---- f2: fused loss ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", TL.GOLDEN_LOSS)
@pytest.mark.parametrize("mode", ["mapping", "tracking"])
def test_loss_kernel_text_matches_the_reference_get_loss(loss, name, mode):
    g = TL._golden(name)
    c = torch.tensor(g["in_color"], requires_grad=True)
    a = torch.tensor(g["in_allmap"], requires_grad=True)
    gc, gd = torch.tensor(g["in_gt_color"]), torch.tensor(g["in_gt_depth"])       # [H,W,3], [H,W,1]: the reference's layout
    for rep in range(2):                                                          # twice: scratch left ready
        c.grad = a.grad = None
        out = loss.mapping_loss(c, a, gc, gd) if mode == "mapping" else loss.tracking_loss(c, a, gc, gd)
        out.backward()
        want = float(g[f"ref_loss_{mode}"])
        assert abs(float(out.detach()) - want) <= 2e-5 * abs(want)
        assert TL._close(c.grad.numpy(), g[f"ref_dcolor_{mode}"], 5e-6)
        assert TL._close(a.grad.numpy(), g[f"ref_dallmap_{mode}"], 1e-5)


@pytest.mark.parametrize("name", TL.GOLDEN_EDGE)
def test_loss_kernel_text_with_edge_growth_matches_the_reference_get_loss(loss, name):
    """densify.use_edge_growth = True: colour and distortion over alpha > edge_thres (slam/Loss.py:53)."""
    g = TL._golden(name)
    c = torch.tensor(g["in_color"], requires_grad=True)
    a = torch.tensor(g["in_allmap"], requires_grad=True)
    out = loss.mapping_loss(c, a, torch.tensor(g["in_gt_color"]), torch.tensor(g["in_gt_depth"]), use_edge_growth=True,
                            edge_thres=float(g["edge_thres"]))
    out.backward()
    want = float(g["ref_loss_mapping"])
    assert abs(float(out.detach()) - want) <= 2e-5 * abs(want)
    assert TL._close(c.grad.numpy(), g["ref_dcolor_mapping"], 5e-6)
    assert TL._close(a.grad.numpy(), g["ref_dallmap_mapping"], 1e-5)


def test_loss_kernel_text_on_nasty_inputs_and_upstream_scale(loss):
    """NaN / inf in every input channel, CHW ground truth, upstream gradient 2: against the float64 restatement."""
    import loss_ref
    x = TL._inputs(24, 40, 0)
    color, allmap, gt_color, gt_depth = x
    l_np, dc_np, da_np = loss_ref.mapping_loss_numpy(*x)
    c = torch.tensor(color, requires_grad=True)
    a = torch.tensor(allmap, requires_grad=True)
    out = loss.mapping_loss(c, a, torch.tensor(gt_color), torch.tensor(gt_depth))
    (2.0 * out).backward()
    assert abs(float(out.detach()) - l_np) <= 1e-5 * abs(l_np)
    assert TL._close(c.grad.numpy(), 2.0 * dc_np, 2e-6) and TL._close(a.grad.numpy(), 2.0 * da_np, 2e-6)
    assert float(a.grad[2:6].abs().max()) == 0.0


# ---- f1: pose transform + pose-gradient reduction -------------------------------------------------------------------------
def _ours(tracking, xyz, rot, g, q=None, t=None, w2c=None, xyz_grad=True):
    x = torch.tensor(xyz, requires_grad=xyz_grad)
    r = torch.tensor(rot)
    leaves = {}
    if w2c is None:
        leaves["q"] = torch.tensor(q, dtype=torch.float32, requires_grad=True)
        leaves["t"] = torch.tensor(t, dtype=torch.float32, requires_grad=True)
        kw = dict(cam_rot=leaves["q"], cam_trans=leaves["t"])
    else:
        leaves["w2c"] = torch.tensor(w2c, dtype=torch.float32, requires_grad=True)
        kw = dict(w2c=leaves["w2c"])
    means, rots = tracking.pose_transform(x, r, **kw)
    assert not rots.requires_grad
    means.backward(torch.tensor(g))
    out = dict(means=means.detach().numpy(), rots=rots.numpy())
    out.update({"d" + k: v.grad.numpy() for k, v in leaves.items()})
    if xyz_grad:
        out["dxyz"] = x.grad.numpy()
    return out


@pytest.mark.parametrize("P,pose", [(300, 0), (300, 1), (1500, 2), (1, 3), (300, 4)])
def test_pose_kernel_text_equals_the_reference_path(tracking, P, pose):
    xyz, rot, g = TP._scene(P, 10 + pose)
    q, t = TP.POSES[pose], np.array([0.3, -0.2, 0.5])
    ref = TP._torch_path(xyz, rot, g, q=q, t=t)                    # float64 torch
    for rep in range(2):                                            # twice: the ticket must be left ready
        got = _ours(tracking, xyz, rot, g, q=q, t=t, xyz_grad=(rep == 0))
        assert TP._rel(got["means"], ref["means"]) <= 1e-6
        assert TP._rot_err(got["rots"], ref["rots"]) <= 2e-6
        assert TP._rel(got["dt"], ref["dt"]) <= 5e-4 and TP._rel(got["dq"], ref["dq"]) <= 5e-4
        if rep == 0:
            assert TP._rel(got["dxyz"], ref["dxyz"]) <= 1e-6
    got2 = _ours(tracking, xyz, rot, g, w2c=ref["w2c"])
    ref2 = TP._torch_path(xyz, rot, g, w2c=ref["w2c"])
    assert TP._rel(got2["means"], ref2["means"]) <= 1e-6 and TP._rot_err(got2["rots"], ref2["rots"]) <= 2e-6
    assert TP._rel(got2["dw2c"], ref2["dw2c"]) <= 5e-4 and np.all(got2["dw2c"][3] == 0)


def test_groups_without_gradient_are_left_untouched_like_torch_adam(optim):
    """Renderer_BA / Renderer_tracking detach the rotations: `_rotation.grad` is None and torch.optim.Adam skips the
    parameter (moments and step included).  MapOptimizer.step(skip=("rotation",)) must do the same, step after step."""
    g, raw, args = TO._golden("aniso")
    opt = optim.MapOptimizer(*[torch.tensor(raw[k]) for k in R.GROUPS], args)
    ref = R.MapTorch(raw, args)
    for step in (1, 2, 3):
        grad = g["in_grad"][step - 1]
        skip = ("rotation",) if step != 2 else ()                     # with, without, with again
        ref.step(grad, detached=("rotations",) if skip else ())
        opt.step(torch.from_numpy(grad.copy()), skip=skip)
        st = ref.state()
        for k in R.GROUPS:
            assert TO._rel(opt.raw[k].numpy(), st[k].astype(np.float64)) <= 2e-6, (step, k)
            assert TO._rel(opt.exp_avg[k].numpy(), st["m_" + k].astype(np.float64)) <= 1e-5 or not st["m_" + k].any(), (step, k)
        assert TO._rel(opt.rotations.numpy(), st["rotations"].astype(np.float64)) <= 2e-6
        if step == 1:
            assert np.array_equal(opt.raw["rotation"].numpy(), raw["rotation"]) and not opt.exp_avg["rotation"].any()
    with pytest.raises(KeyError):
        opt.step(torch.zeros(13 * opt.P), skip=("f_dc",))


def test_pose_adam_kernel_text_equals_torch_adam(tracking):
    """The fused optimiser of the seven pose scalars (host build of its kernel) against torch.optim.Adam with the
    tracker's settings (scene/Frame.py:70-73), including a step without a gradient for one parameter."""
    want = TP.pose_adam_sequence(TP.torch_pose_adam, "cpu")
    got = TP.pose_adam_sequence(lambda q, t: tracking.PoseOptimizer(q, t, lr_rot=4e-4, lr_trans=2e-3), "cpu")
    TP.check_pose_adam(got, want)
