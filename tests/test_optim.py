"""Fused map optimiser step.  CPU: the float64 restatement with hand-written activation
gradients agrees with the reference's own machinery (torch activations + torch.optim.Adam, run on the CPU).  GPU:
gs2d_map_activate / gs2d_map_adam_step through the C ABI (gs2d_b200.optim.MapOptimizer) agree with both over several
steps -- fp32 tolerances written at each assert.  Golden: tests/golden/optim_*.npz hold the states the reference's own
`Gaussians.training_setup()` / `get_render_params` / `optimizer.step()` produced (scene/Gaussians.py imported
unmodified, tests/golden/make_optim_golden.py); the restatements (CPU) and the fused kernel (GPU) are held to them."""
import os

import numpy as np
import pytest
import torch

import optim_ref as R

ARGS = dict(xyz_lr=1e-4, opacity_lr=5e-2, scaling_lr=5e-3, rotation_lr=1e-3, rgb_lr=2.5e-3)


def _raw(P, scale_cols, seed):
    rng = np.random.default_rng(seed)
    return dict(xyz=rng.normal(0, 2, (P, 3)).astype(np.float32), opacity=rng.normal(0, 1.5, (P, 1)).astype(np.float32),
                scaling=rng.normal(-4, 0.5, (P, scale_cols)).astype(np.float32),
                rotation=(rng.normal(0, 1, (P, 4)) * rng.uniform(0.5, 2.0, (P, 1))).astype(np.float32),
                rgb=rng.uniform(0, 1, (P, 3)).astype(np.float32))


def _grad(P, seed, step):
    rng = np.random.default_rng(1000 * seed + step)
    g = (rng.normal(0, 1, 13 * P) * 10.0 ** rng.uniform(-6, -2, 13 * P)).astype(np.float32)
    g[rng.uniform(size=13 * P) < 0.3] = 0.0        # surfels a view does not see get exactly zero gradient
    return g


GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLDEN_CASES = ("aniso", "iso", "aligned")


def _golden(name):
    path = os.path.join(GOLDEN, f"optim_{name}.npz")
    assert os.path.exists(path), f"{path} missing (tests/golden/make_optim_golden.py)"
    g = dict(np.load(path))
    raw = {k: g["in_" + k] for k in R.GROUPS}
    args = dict(zip(R.LR_KEYS, (float(x) for x in g["in_lr"])))
    return g, raw, args


def _rel(a, b):
    return float(np.abs(a.astype(np.float64) - b).max() / max(np.abs(b).max(), 1e-30))


@pytest.mark.parametrize("P,scale_cols,seed", [(257, 2, 0), (64, 1, 1)])
def test_numpy_restatement_equals_torch_adam(P, scale_cols, seed):
    raw = _raw(P, scale_cols, seed)
    ref = R.MapTorch(raw, ARGS)
    x = {k: raw[k].astype(np.float64) for k in R.GROUPS}
    m = {k: np.zeros_like(x[k]) for k in R.GROUPS}
    v = {k: np.zeros_like(x[k]) for k in R.GROUPS}
    lr = [ARGS[k] for k in R.LR_KEYS]
    for step in range(1, 5):
        g = _grad(P, seed, step)
        ref.step(g)
        x, m, v, act = R.map_step_numpy(x, m, v, g, lr, step)
        st = ref.state()
        for k in R.GROUPS:
            assert _rel(st[k], x[k]) <= 2e-6, (step, k)
            assert _rel(st["m_" + k], m[k]) <= 1e-5, (step, k)
            assert _rel(st["v_" + k], v[k]) <= 1e-5, (step, k)
        for k in ("opacities", "scales", "rotations"):
            assert _rel(st[k], act[k]) <= 1e-6, (step, k)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_restatements_match_the_reference_gaussians_class(name):
    """MapTorch (torch ops, fp32) must reproduce the reference's states BIT FOR BIT -- same torch, same ops, same
    order; the float64 numpy restatement to fp32 rounding."""
    g, raw, args = _golden(name)
    ref = R.MapTorch(raw, args)
    x = {k: raw[k].astype(np.float64) for k in R.GROUPS}
    m = {k: np.zeros_like(x[k]) for k in R.GROUPS}
    v = {k: np.zeros_like(x[k]) for k in R.GROUPS}
    lr = [args[k] for k in R.LR_KEYS]
    for step in range(1, g["in_grad"].shape[0] + 1):
        grad = g["in_grad"][step - 1]
        ref.step(grad)
        x, m, v, act = R.map_step_numpy(x, m, v, grad, lr, step)
        st = ref.state()
        for k in R.GROUPS:
            assert np.array_equal(st[k], g[f"ref_{k}_{step}"]), (step, k)
            assert np.array_equal(st["m_" + k], g[f"ref_m_{k}_{step}"]), (step, k)
            assert np.array_equal(st["v_" + k], g[f"ref_v_{k}_{step}"]), (step, k)
            assert _rel(g[f"ref_{k}_{step}"], x[k]) <= 2e-6, (step, k)
            assert _rel(g[f"ref_m_{k}_{step}"], m[k]) <= 1e-5, (step, k)
            assert _rel(g[f"ref_v_{k}_{step}"], v[k]) <= 1e-5, (step, k)
        for k in ("opacities", "scales", "rotations"):
            assert np.array_equal(st[k], g[f"ref_{k}_{step}"]), (step, k)
            assert _rel(g[f"ref_{k}_{step}"], act[k]) <= 1e-6, (step, k)


@pytest.mark.gpu
@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_fused_step_matches_the_reference_gaussians_class(name):
    from gs2d_b200 import optim
    dev = torch.device("cuda:0")
    g, raw, args = _golden(name)
    opt = optim.MapOptimizer(*[torch.tensor(raw[k], device=dev) for k in R.GROUPS], args)
    for step in range(1, g["in_grad"].shape[0] + 1):
        opt.flat_grad.copy_(torch.from_numpy(g["in_grad"][step - 1]))
        opt.step()
        for k in R.GROUPS:
            assert _rel(opt.raw[k].cpu().numpy(), g[f"ref_{k}_{step}"].astype(np.float64)) <= 2e-6, (step, k)
            assert _rel(opt.exp_avg[k].cpu().numpy(), g[f"ref_m_{k}_{step}"].astype(np.float64)) <= 1e-5, (step, k)
            assert _rel(opt.exp_avg_sq[k].cpu().numpy(), g[f"ref_v_{k}_{step}"].astype(np.float64)) <= 1e-5, (step, k)
        for k in ("opacities", "scales", "rotations"):
            assert _rel(getattr(opt, k).cpu().numpy(), g[f"ref_{k}_{step}"].astype(np.float64)) <= 2e-6, (step, k)
    last = g["in_grad"].shape[0]
    moved = opt.raw["opacity"].cpu().numpy() - raw["opacity"]
    want = g[f"ref_opacity_{last}"] - raw["opacity"]
    assert np.abs(moved - want).max() <= 1e-4 * np.abs(want).max() + 1e-7


@pytest.mark.gpu
@pytest.mark.parametrize("P,scale_cols,seed", [(257, 2, 0), (64, 1, 1), (4096, 2, 2), (1, 2, 3), (100003, 2, 4)])
def test_fused_step_equals_torch_adam(P, scale_cols, seed):
    """Odd P: the flat-buffer sections are not 16-byte aligned (scalar path); P % 4 == 0: 128-bit path."""
    from gs2d_b200 import optim
    dev = torch.device("cuda:0")
    raw = _raw(P, scale_cols, seed)
    ref = R.MapTorch(raw, ARGS)                       # the reference's optimiser on the CPU
    opt = optim.MapOptimizer(*[torch.tensor(raw[k], device=dev) for k in R.GROUPS], ARGS)
    act0 = R.activate_numpy(raw)
    for k in ("opacities", "scales", "rotations"):
        assert _rel(getattr(opt, k).cpu().numpy(), act0[k]) <= 1e-6
    for step in range(1, 5):
        g = _grad(P, seed, step)
        ref.step(g)
        opt.flat_grad.copy_(torch.from_numpy(g))
        opt.step()
        st = ref.state()
        for k in R.GROUPS:
            # parameters move by ~lr per step; compare the values (2e-6 rel) AND the moments
            assert _rel(opt.raw[k].cpu().numpy(), st[k].astype(np.float64)) <= 2e-6, (step, k)
            assert _rel(opt.exp_avg[k].cpu().numpy(), st["m_" + k].astype(np.float64)) <= 1e-5, (step, k)
            assert _rel(opt.exp_avg_sq[k].cpu().numpy(), st["v_" + k].astype(np.float64)) <= 1e-5, (step, k)
        for k in ("opacities", "scales", "rotations"):
            assert _rel(getattr(opt, k).cpu().numpy(), st[k].astype(np.float64)) <= 2e-6, (step, k)
    # the step the parameters took is the reference's step (not just "close to the start value")
    moved = opt.raw["opacity"].cpu().numpy() - raw["opacity"]
    want = ref.state()["opacity"] - raw["opacity"]
    assert np.abs(moved - want).max() <= 1e-4 * np.abs(want).max() + 1e-7


@pytest.mark.gpu
def test_render_params_are_views_of_the_optimiser_state():
    from gs2d_b200 import optim
    dev = torch.device("cuda:0")
    raw = _raw(128, 2, 7)
    opt = optim.MapOptimizer(*[torch.tensor(raw[k], device=dev) for k in R.GROUPS], ARGS)
    rp = opt.render_params()
    assert rp["means3D"].data_ptr() == opt.raw["xyz"].data_ptr() and rp["opacities"].requires_grad
    assert rp["scales"].shape == (128, 2) and rp["rotations"].shape == (128, 4)
    with pytest.raises(RuntimeError):
        optim.MapOptimizer(*[torch.tensor(raw[k]) for k in R.GROUPS], ARGS)       # CPU tensors: no CPU path
    with pytest.raises(RuntimeError):
        opt.step(torch.zeros(5, device=dev))
