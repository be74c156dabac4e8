"""Fused mapping loss.  CPU: the two restatements of the reference's loss (torch
transliteration with autograd, numpy with hand-written gradients) agree.  GPU: the CUDA kernels agree with both,
through the C ABI (gs2d_b200.loss) -- fp32 tolerance 1e-5 relative on the loss, 1e-6 abs-to-max on gradients."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
import loss_ref  # noqa: E402


def _inputs(H, W, seed, nasty=True):
    rng = np.random.default_rng(seed)
    color = rng.uniform(0, 1, (3, H, W)).astype(np.float32)
    allmap = rng.normal(0, 1, (7, H, W)).astype(np.float32)
    alpha = rng.uniform(0.0, 1.0, (H, W)).astype(np.float32)
    depth = rng.uniform(0.5, 5.0, (H, W)).astype(np.float32)
    allmap[1] = alpha
    allmap[0] = depth * alpha
    allmap[6] = rng.uniform(0, 0.05, (H, W)).astype(np.float32)
    gt_color = rng.uniform(0, 1, (3, H, W)).astype(np.float32)
    gt_depth = rng.uniform(0.5, 5.0, (H, W)).astype(np.float32)
    if nasty:
        gt_depth[rng.uniform(size=(H, W)) < 0.1] = 0.0           # holes in the sensor depth
        allmap[1][rng.uniform(size=(H, W)) < 0.05] = 0.0         # nothing rendered: depth = D / eps -> outlier
        allmap[0][rng.uniform(size=(H, W)) < 0.02] = np.nan
        allmap[0][rng.uniform(size=(H, W)) < 0.02] = np.inf
        color[:, rng.uniform(size=(H, W)) < 0.02] = np.nan
        color[0][rng.uniform(size=(H, W)) < 0.01] = np.inf
        allmap[6][rng.uniform(size=(H, W)) < 0.02] = np.inf
        allmap[0][0, :4] = 1e6                                    # beyond depth_far
        color[1][1, :4] = gt_color[1][1, :4]                      # exact zeros of |c - gt|: sign(0) = 0
    return color, allmap, gt_color, gt_depth


def _close(got, want_torch, rel):
    """Compare with torch autograd.  One documented divergence: where the rendered depth sum allmap[0] is not
    finite, autograd back-propagates 0 * inf = NaN into dL/dallmap[1]; the restatements (and the kernels) write 0."""
    bad = ~np.isfinite(want_torch)
    assert np.all(got[bad] == 0)
    w = np.where(bad, 0.0, want_torch)
    return np.abs(got - w).max() <= rel * np.abs(w).max()


def _torch_ref(color, allmap, gt_color, gt_depth, dev="cpu", **kw):
    c = torch.tensor(color, device=dev, requires_grad=True)
    a = torch.tensor(allmap, device=dev, requires_grad=True)
    gc = torch.tensor(gt_color, device=dev).permute(1, 2, 0).contiguous()
    gd = torch.tensor(gt_depth, device=dev)[..., None]
    loss = loss_ref.mapping_loss_torch(c, a, gc, gd, **kw)
    loss.backward()
    return float(loss.detach()), c.grad.cpu().numpy(), a.grad.cpu().numpy()


@pytest.mark.parametrize("H,W,seed", [(24, 40, 0), (17, 33, 1), (64, 96, 2)])
def test_numpy_restatement_equals_torch_transliteration(H, W, seed):
    x = _inputs(H, W, seed)
    l0, dc0, da0 = _torch_ref(*x)
    l1, dc1, da1 = loss_ref.mapping_loss_numpy(*x)
    assert abs(l0 - l1) <= 1e-5 * abs(l1)
    assert _close(dc1, dc0, 1e-6) and _close(da1, da0, 1e-5)
    assert np.isnan(da0[1]).sum() > 0 and np.all(da1[2:6] == 0)


def test_weights_and_upstream_gradient_scale():
    x = _inputs(20, 28, 5, nasty=False)
    kw = dict(weights=(0.3, 2.0, 0.7), eps=1e-5, depth_near=0.6, depth_far=4.0)
    l0, dc0, da0 = _torch_ref(*x, **kw)
    l1, dc1, da1 = loss_ref.mapping_loss_numpy(*x, **kw)
    assert abs(l0 - l1) <= 1e-5 * abs(l1)
    assert np.abs(da0 - da1).max() <= 1e-5 * np.abs(da1).max()
    _, dc2, _ = loss_ref.mapping_loss_numpy(*x, grad_out=3.0, **kw)
    assert np.allclose(dc2, 3.0 * dc1)


@pytest.mark.gpu
@pytest.mark.parametrize("H,W,seed,hwc", [(24, 40, 0, False), (17, 33, 1, True), (680, 1200, 2, True)])
def test_fused_kernels_equal_the_reference_loss(H, W, seed, hwc):
    from gs2d_b200 import loss as gl
    dev = torch.device("cuda:0")
    x = _inputs(H, W, seed)
    color, allmap, gt_color, gt_depth = x
    l_np, dc_np, da_np = loss_ref.mapping_loss_numpy(*x)
    l_t, dc_t, da_t = _torch_ref(*x, dev=dev)
    c = torch.tensor(color, device=dev, requires_grad=True)
    a = torch.tensor(allmap, device=dev, requires_grad=True)
    gc = torch.tensor(gt_color, device=dev)
    gc = gc.permute(1, 2, 0).contiguous() if hwc else gc
    gd = torch.tensor(gt_depth, device=dev)
    for rep in range(2):                      # twice: the scratch must be left ready for the next call
        c.grad = a.grad = None
        loss = gl.mapping_loss(c, a, gc, gd)
        (2.0 * loss).backward()
        assert abs(float(loss) - l_np) <= 1e-5 * abs(l_np)
        assert abs(float(loss) - l_t) <= 1e-5 * abs(l_t)
        for got, want in ((c.grad, 2.0 * dc_np), (a.grad, 2.0 * da_np), (c.grad, 2.0 * dc_t), (a.grad, 2.0 * da_t)):
            assert _close(got.cpu().numpy(), want, 2e-6)
        assert float(a.grad[2:6].abs().max()) == 0.0


def _torch_tracking(color, allmap, gt_color, gt_depth, dev="cpu"):
    c = torch.tensor(color, device=dev, requires_grad=True)
    a = torch.tensor(allmap, device=dev, requires_grad=True)
    gc = torch.tensor(gt_color, device=dev).permute(1, 2, 0).contiguous()
    gd = torch.tensor(gt_depth, device=dev)[..., None]
    loss = loss_ref.tracking_loss_torch(c, a, gc, gd)
    loss.backward()
    return float(loss.detach()), c.grad.cpu().numpy(), a.grad.cpu().numpy()


@pytest.mark.parametrize("H,W,seed", [(24, 40, 0), (33, 17, 4)])
def test_tracking_restatement_equals_torch_transliteration(H, W, seed):
    x = _inputs(H, W, seed)
    x[1][1][np.random.default_rng(seed).uniform(size=(H, W)) < 0.5] = 0.95      # half the pixels pass the silhouette mask
    l0, dc0, da0 = _torch_tracking(*x)
    l1, dc1, da1 = loss_ref.mapping_loss_numpy(*x, weights=(0.5, 1.0, 0.0), tracking=True)
    assert l1 > 0 and abs(l0 - l1) <= 1e-5 * abs(l1)
    assert _close(dc1, dc0, 1e-6) and _close(da1, da0, 1e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("H,W,seed,hwc", [(24, 40, 0, False), (680, 1200, 3, True)])
def test_fused_tracking_loss_equals_the_reference_loss(H, W, seed, hwc):
    from gs2d_b200 import loss as gl
    dev = torch.device("cuda:0")
    x = _inputs(H, W, seed)
    x[1][1][np.random.default_rng(seed).uniform(size=(H, W)) < 0.5] = 0.95
    color, allmap, gt_color, gt_depth = x
    l_np, dc_np, da_np = loss_ref.mapping_loss_numpy(*x, weights=(0.5, 1.0, 0.0), tracking=True)
    l_t, dc_t, da_t = _torch_tracking(*x, dev=dev)
    c = torch.tensor(color, device=dev, requires_grad=True)
    a = torch.tensor(allmap, device=dev, requires_grad=True)
    gc = torch.tensor(gt_color, device=dev)
    gc = gc.permute(1, 2, 0).contiguous() if hwc else gc
    loss = gl.tracking_loss(c, a, gc, torch.tensor(gt_depth, device=dev))
    loss.backward()
    assert abs(float(loss.detach()) - l_np) <= 1e-5 * abs(l_np) and abs(float(loss.detach()) - l_t) <= 2e-5 * abs(l_t)
    for got, want in ((c.grad, dc_np), (a.grad, da_np), (c.grad, dc_t), (a.grad, da_t)):
        assert _close(got.cpu().numpy(), want, 2e-6)
    with pytest.raises(NotImplementedError):
        gl.tracking_loss(c, a, gc, torch.tensor(gt_depth, device=dev), use_normal_loss=True)


@pytest.mark.gpu
def test_fused_loss_rejects_what_is_not_built():
    from gs2d_b200 import loss as gl
    t = torch.zeros((3, 8, 8), device="cuda:0")
    with pytest.raises(NotImplementedError):
        gl.mapping_loss(t, torch.zeros((7, 8, 8), device="cuda:0"), t, torch.zeros((8, 8), device="cuda:0"),
                        use_normal_loss=True)
    with pytest.raises(RuntimeError):
        gl.mapping_loss(t.cpu(), torch.zeros((7, 8, 8)), t.cpu(), torch.zeros((8, 8)))


# ---- golden vectors from the reference's OWN get_loss (tests/golden/make_loss_golden.py: slam/Loss.py imported
# unmodified in the authoring container, run on the CPU) --------------------------------------------------------
GOLDEN_LOSS = ["small", "ragged", "clean"]


def _golden(name):
    p = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", f"loss_{name}.npz")
    with np.load(p) as z:
        return {k: z[k] for k in z.files}


@pytest.mark.parametrize("name", GOLDEN_LOSS)
@pytest.mark.parametrize("mode", ["mapping", "tracking"])
def test_restatements_match_the_reference_get_loss(name, mode):
    """Pins oracle/loss_ref.py: both restatements against what the reference's code computed."""
    g = _golden(name)
    color, allmap = g["in_color"], g["in_allmap"]
    gt_chw = np.ascontiguousarray(g["in_gt_color"].transpose(2, 0, 1))
    gt_depth = g["in_gt_depth"][..., 0]
    want = float(g[f"ref_loss_{mode}"]), g[f"ref_dcolor_{mode}"], g[f"ref_dallmap_{mode}"]
    if mode == "mapping":
        got_np = loss_ref.mapping_loss_numpy(color, allmap, gt_chw, gt_depth)
        got_t = _torch_ref(color, allmap, gt_chw, gt_depth)
    else:
        got_np = loss_ref.mapping_loss_numpy(color, allmap, gt_chw, gt_depth, weights=(0.5, 1.0, 0.0), tracking=True)
        got_t = _torch_tracking(color, allmap, gt_chw, gt_depth)
    for got in (got_np, got_t):
        assert abs(got[0] - want[0]) <= 1e-5 * abs(want[0])
        assert _close(np.asarray(got[1]), want[1], 2e-6) and _close(np.asarray(got[2]), want[2], 1e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("name", GOLDEN_LOSS)
@pytest.mark.parametrize("mode", ["mapping", "tracking"])
def test_fused_kernels_match_the_reference_get_loss(name, mode):
    from gs2d_b200 import loss as gl
    dev = torch.device("cuda:0")
    g = _golden(name)
    c = torch.tensor(g["in_color"], device=dev, requires_grad=True)
    a = torch.tensor(g["in_allmap"], device=dev, requires_grad=True)
    gc = torch.tensor(g["in_gt_color"], device=dev)               # [H,W,3], the reference's layout
    gd = torch.tensor(g["in_gt_depth"], device=dev)               # [H,W,1]
    loss = gl.mapping_loss(c, a, gc, gd) if mode == "mapping" else gl.tracking_loss(c, a, gc, gd)
    loss.backward()
    want = float(g[f"ref_loss_{mode}"])
    assert abs(float(loss.detach()) - want) <= 2e-5 * abs(want)
    assert _close(c.grad.cpu().numpy(), g[f"ref_dcolor_{mode}"], 5e-6)
    assert _close(a.grad.cpu().numpy(), g[f"ref_dallmap_{mode}"], 1e-5)


GOLDEN_EDGE = ["edge", "edge_ragged"]


@pytest.mark.parametrize("name", GOLDEN_EDGE)
def test_edge_growth_restatement_matches_the_reference_get_loss(name):
    """CPU: the transliteration with densify.use_edge_growth = True (colour and distortion averaged over
    alpha > edge_thres, slam/Loss.py:53) against what the reference's own get_loss computed."""
    g = _golden(name)
    c = torch.tensor(g["in_color"], requires_grad=True)
    a = torch.tensor(g["in_allmap"], requires_grad=True)
    loss = loss_ref.mapping_loss_torch(c, a, torch.tensor(g["in_gt_color"]), torch.tensor(g["in_gt_depth"]),
                                       use_edge_growth=True, edge_thres=float(g["edge_thres"]))
    loss.backward()
    want = float(g["ref_loss_mapping"])
    assert abs(float(loss.detach()) - want) <= 1e-5 * abs(want)
    assert _close(c.grad.numpy(), g["ref_dcolor_mapping"], 2e-6) and _close(a.grad.numpy(), g["ref_dallmap_mapping"], 1e-5)
    # the branch is really exercised: the two masks differ
    plain = loss_ref.mapping_loss_torch(torch.tensor(g["in_color"]), torch.tensor(g["in_allmap"]),
                                        torch.tensor(g["in_gt_color"]), torch.tensor(g["in_gt_depth"]))
    assert abs(float(plain) - want) > 1e-3 * abs(want)


@pytest.mark.gpu
@pytest.mark.parametrize("name", GOLDEN_EDGE)
def test_fused_mapping_loss_with_edge_growth_matches_the_reference_get_loss(name):
    from gs2d_b200 import loss as gl
    dev = torch.device("cuda:0")
    g = _golden(name)
    c = torch.tensor(g["in_color"], device=dev, requires_grad=True)
    a = torch.tensor(g["in_allmap"], device=dev, requires_grad=True)
    gc = torch.tensor(g["in_gt_color"], device=dev)
    gd = torch.tensor(g["in_gt_depth"], device=dev)
    loss = gl.mapping_loss(c, a, gc, gd, use_edge_growth=True, edge_thres=float(g["edge_thres"]))
    loss.backward()
    want = float(g["ref_loss_mapping"])
    assert abs(float(loss.detach()) - want) <= 2e-5 * abs(want)
    assert _close(c.grad.cpu().numpy(), g["ref_dcolor_mapping"], 5e-6)
    assert _close(a.grad.cpu().numpy(), g["ref_dallmap_mapping"], 1e-5)


def test_inlier_mask_restatement_matches_the_reference_get_loss():
    """CPU: the torch ops that build the `ignore_outliners` mask (gs2d_b200.loss._inlier_gt_depth) + the transliterated
    tracking loss reproduce what the reference's get_loss computed with ignore_outliners = True, silmask_th = 0.99
    (configs/replica/splatam.py)."""
    from gs2d_b200 import loss as gl
    g = _golden("outliers")
    a = torch.tensor(g["in_allmap"])
    gd = gl._inlier_gt_depth(a, torch.tensor(g["in_gt_depth"]), 1e-6, 1e-2, 1e2)
    gt_chw = np.ascontiguousarray(g["in_gt_color"].transpose(2, 0, 1))
    got = loss_ref.mapping_loss_numpy(g["in_color"], g["in_allmap"], gt_chw, gd.numpy()[..., 0], weights=(0.5, 1.0, 0.0),
                                      tracking=True, silmask_th=float(g["silmask_th"]))
    want = float(g["ref_loss_tracking_inliers"])
    assert abs(got[0] - want) <= 1e-5 * abs(want)
    assert abs(float(g["ref_loss_tracking_all"]) - want) > 0.05 * abs(want)       # the mask matters in this case
    assert _close(np.asarray(got[1]), g["ref_dcolor_tracking_inliers"], 2e-6)
    assert _close(np.asarray(got[2]), g["ref_dallmap_tracking_inliers"], 1e-5)


@pytest.mark.gpu
def test_fused_tracking_loss_with_the_outlier_mask_matches_the_reference_get_loss():
    from gs2d_b200 import loss as gl
    dev = torch.device("cuda:0")
    g = _golden("outliers")
    c = torch.tensor(g["in_color"], device=dev, requires_grad=True)
    a = torch.tensor(g["in_allmap"], device=dev, requires_grad=True)
    gc = torch.tensor(g["in_gt_color"], device=dev)
    gd = torch.tensor(g["in_gt_depth"], device=dev)
    loss = gl.tracking_loss(c, a, gc, gd, silmask_th=float(g["silmask_th"]), ignore_outliners=True)
    loss.backward()
    want = float(g["ref_loss_tracking_inliers"])
    assert abs(float(loss.detach()) - want) <= 2e-5 * abs(want)
    assert _close(c.grad.cpu().numpy(), g["ref_dcolor_tracking_inliers"], 5e-6)
    assert _close(a.grad.cpu().numpy(), g["ref_dallmap_tracking_inliers"], 1e-5)


@pytest.mark.gpu
def test_two_losses_before_one_backward():
    """(loss(view 1) + loss(view 2)).backward(): every backward must divide by the mask count of ITS forward
    (round 1 kept the count in one scratch per stream)."""
    from gs2d_b200 import loss as gl
    dev = torch.device("cuda:0")
    g1, g2 = _golden("small"), _golden("clean")

    def leaves(g):
        return (torch.tensor(g["in_color"], device=dev, requires_grad=True),
                torch.tensor(g["in_allmap"], device=dev, requires_grad=True),
                torch.tensor(g["in_gt_color"], device=dev), torch.tensor(g["in_gt_depth"], device=dev))
    x1, x2 = leaves(g1), leaves(g2)
    (gl.mapping_loss(*x1) + gl.mapping_loss(*x2)).backward()
    for x, g in ((x1, g1), (x2, g2)):
        assert _close(x[0].grad.cpu().numpy(), g["ref_dcolor_mapping"], 5e-6)
        assert _close(x[1].grad.cpu().numpy(), g["ref_dallmap_mapping"], 1e-5)
