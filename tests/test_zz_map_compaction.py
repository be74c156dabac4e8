"""Densification surgery on the map optimiser (second half): prune / extend / replace.

Golden: tests/golden/optim_surgery.npz is the reference's OWN code (scene/Gaussians.py imported unmodified by
tests/golden/make_optim_golden.py) running  step, step, remove_gaussians_from_mask, step, densification_postfix,
step, reset_opacity, step.  CPU: the restatement (oracle/optim_ref.py MapTorch.prune / extend / replace) reproduces
every state bit for bit.  GPU: gs2d_b200.optim.MapOptimizer (gs2d_map_compact_plan / _move through the C ABI) is
held to the same states -- moved rows bit-exact, optimiser steps at the fp32 tolerances of tests/test_optim.py.

The file sorts last on purpose: the compaction kernels were written after this round's GPU budget was spent, so
their first run on a B200 is the driver's round-end run.
"""
import os

import numpy as np
import pytest
import torch

import optim_ref as R

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "optim_surgery.npz")


def _golden():
    assert os.path.exists(GOLDEN), "tests/golden/optim_surgery.npz missing (tests/golden/make_optim_golden.py)"
    g = dict(np.load(GOLDEN))
    raw = {k: g["in_" + k] for k in R.GROUPS}
    args = dict(zip(R.LR_KEYS, (float(x) for x in g["in_lr"])))
    return g, raw, args


def _rel(a, b):
    return float(np.abs(a.astype(np.float64) - b).max() / max(np.abs(b).max(), 1e-30))


def test_restatement_reproduces_the_reference_surgery_bit_for_bit():
    g, raw, args = _golden()
    ref = R.MapTorch(raw, args)

    def same(tag):
        st = ref.state()
        for k in R.GROUPS:
            for pre in ("", "m_", "v_"):
                assert np.array_equal(st[pre + k], g[f"ref_{pre}{k}_{tag}"]), (tag, pre + k)
        for k in ("opacities", "scales", "rotations"):
            assert np.array_equal(st[k], g[f"ref_{k}_{tag}"]), (tag, k)

    for tag in ("1", "2"):
        ref.step(g[f"in_grad_{tag}"]); same(tag)
    ref.prune(~g["in_remove_mask"]); same("pruned")
    ref.step(g["in_grad_3"]); same("3")
    ref.extend({k: g["in_new_" + k] for k in R.GROUPS}); same("extended")
    ref.step(g["in_grad_4"]); same("4")
    ref.replace("opacity", ref.reset_opacity_values()); same("reset")
    ref.step(g["in_grad_5"]); same("5")


def _check(opt, g, tag, exact):
    tol = dict(p=0.0, m=0.0, a=2e-6) if exact else dict(p=2e-6, m=1e-5, a=2e-6)
    for k in R.GROUPS:
        want = g[f"ref_{k}_{tag}"]
        assert tuple(opt.raw[k].shape) == want.shape, (tag, k)
        assert _rel(opt.raw[k].cpu().numpy(), want.astype(np.float64)) <= tol["p"], (tag, k)
        assert _rel(opt.exp_avg[k].cpu().numpy(), g[f"ref_m_{k}_{tag}"].astype(np.float64)) <= tol["m"], (tag, "m", k)
        assert _rel(opt.exp_avg_sq[k].cpu().numpy(), g[f"ref_v_{k}_{tag}"].astype(np.float64)) <= tol["m"], (tag, "v", k)
    for k in ("opacities", "scales", "rotations"):
        assert _rel(getattr(opt, k).cpu().numpy(), g[f"ref_{k}_{tag}"].astype(np.float64)) <= tol["a"], (tag, k)


@pytest.mark.gpu
def test_map_optimizer_surgery_matches_the_reference():
    from gs2d_b200 import optim
    dev = torch.device("cuda:0")
    g, raw, args = _golden()
    opt = optim.MapOptimizer(*[torch.tensor(raw[k], device=dev) for k in R.GROUPS], args)

    def step(tag):
        opt.step(torch.tensor(g[f"in_grad_{tag}"], device=dev))

    # the states the kernel is compared with drift by fp32 rounding (2e-6 / 1e-5, tests/test_optim.py); surgery
    # itself must move what the optimiser holds without changing a bit -- checked against its own state below
    step("1"); _check(opt, g, "1", False)
    step("2"); _check(opt, g, "2", False)
    keep = torch.tensor(~g["in_remove_mask"], device=dev)
    before = {k: (opt.raw[k].clone(), opt.exp_avg[k].clone(), opt.exp_avg_sq[k].clone()) for k in R.GROUPS}
    n = opt.prune(keep)
    assert n == int(keep.sum()) == opt.P and opt.flat_grad.numel() == 13 * n
    for k in R.GROUPS:
        for got, old in zip((opt.raw[k], opt.exp_avg[k], opt.exp_avg_sq[k]), before[k]):
            assert torch.equal(got, old[keep]), k                      # tensor[mask] semantics, bit for bit
    _check(opt, g, "pruned", False)
    step("3"); _check(opt, g, "3", False)
    P0 = opt.P
    opt.extend(*[torch.tensor(g["in_new_" + k], device=dev) for k in R.GROUPS])
    assert opt.P == P0 + g["in_new_xyz"].shape[0]
    for k in R.GROUPS:
        assert torch.equal(opt.raw[k][P0:].cpu(), torch.tensor(g["in_new_" + k])) and float(opt.exp_avg[k][P0:].abs().max()) == 0.0
    _check(opt, g, "extended", False)
    step("4"); _check(opt, g, "4", False)
    opt.replace("opacity", torch.tensor(g["ref_opacity_reset"], device=dev))
    assert float(opt.exp_avg["opacity"].abs().max()) == 0.0 and float(opt.exp_avg_sq["opacity"].abs().max()) == 0.0
    _check(opt, g, "reset", False)
    step("5"); _check(opt, g, "5", False)


@pytest.mark.gpu
@pytest.mark.parametrize("P,frac,seed", [(1, 1.0, 0), (1, 0.0, 1), (255, 0.5, 2), (256, 0.5, 3), (257, 0.99, 4),
                                         (100003, 0.7, 5), (300000, 0.01, 6), (2000, 0.0, 7), (2000, 1.0, 8)])
def test_prune_equals_boolean_indexing(P, frac, seed):
    """Sizes around the 256-surfel block and the 1024-block scan pass; nothing / everything kept; isotropic map."""
    from gs2d_b200 import optim
    dev = torch.device("cuda:0")
    gen = torch.Generator(device="cpu").manual_seed(seed)
    cols = 1 if seed % 2 else 2
    raw = [torch.randn((P, w), generator=gen).to(dev) for w in (3, 1, cols, 4, 3)]
    opt = optim.MapOptimizer(*raw, dict(zip(R.LR_KEYS, (1e-4, 5e-2, 1e-3, 1e-3, 2.5e-3))))
    opt.step(torch.randn(13 * P, generator=gen).to(dev) * 1e-3)          # non-trivial moments
    keep = (torch.rand(P, generator=gen) < frac).to(dev)
    before = {k: (opt.raw[k].clone(), opt.exp_avg[k].clone(), opt.exp_avg_sq[k].clone()) for k in R.GROUPS}
    n = opt.prune(keep)
    assert n == int(keep.sum())
    for k in R.GROUPS:
        for got, old in zip((opt.raw[k], opt.exp_avg[k], opt.exp_avg_sq[k]), before[k]):
            assert got.shape == old[keep].shape and torch.equal(got, old[keep]), k
    assert opt.scales.shape == (n, 2) and opt.step_count == 1
    if n:
        opt.step(torch.zeros(13 * n, device=dev))                         # the compacted state is usable
        assert torch.isfinite(opt.raw["xyz"]).all()


@pytest.mark.gpu
def test_surgery_argument_checks():
    from gs2d_b200 import optim
    dev = torch.device("cuda:0")
    raw = [torch.randn((8, w), device=dev) for w in (3, 1, 2, 4, 3)]
    opt = optim.MapOptimizer(*raw, dict(zip(R.LR_KEYS, (1e-4,) * 5)))
    with pytest.raises(RuntimeError):
        opt.prune(torch.ones(7, dtype=torch.bool, device=dev))
    with pytest.raises(RuntimeError):
        opt.prune(torch.ones(8, dtype=torch.uint8, device=dev))
    with pytest.raises(RuntimeError):
        opt.extend(*[torch.randn((2, w)) for w in (3, 1, 2, 4, 3)])      # CPU tensors
    with pytest.raises(RuntimeError):
        opt.extend(*[torch.randn((2, w), device=dev) for w in (3, 1, 1, 4, 3)])
    with pytest.raises(KeyError):
        opt.replace("f_dc", raw[0])
