"""TEST INFRASTRUCTURE ONLY (never imported by the product): the reference's mapping loss, restated.

`mapping_loss_torch` follows slam/Loss.py:22-25, :36, :51-63 and render/__init__.py:46-49 line by line with the
same torch calls (so torch autograd gives the gradients the reference would get); `mapping_loss_numpy` is a
loop-free float64 restatement with hand-written gradients, used to check the torch transliteration itself.
Parity pin: tests/golden/loss_*.npz hold what the reference's own get_loss (slam/Loss.py imported unmodified, CPU)
computes for both branches (tests/golden/make_loss_golden.py); tests/test_loss.py checks both restatements against
them.
"""
import numpy as np


def mapping_loss_torch(color, allmap, gt_color_hwc, gt_depth_hw1, weights=(0.5, 1.0, 0.1), eps=1e-6,
                       depth_near=1e-2, depth_far=1e2, use_edge_growth=False, edge_thres=0.4):
    import torch
    # render/__init__.py:46-49 (use_weight_norm)
    render_depth = allmap[0:1] / (allmap[1:2] + eps)
    outliner_mask = torch.logical_or(render_depth > depth_far, render_depth < depth_near)
    render_depth = render_depth.clone()
    render_depth[outliner_mask] = 0
    # slam/Loss.py:22-25
    render_depth = torch.nan_to_num(render_depth, 0, 0).permute(1, 2, 0)
    render_color = torch.nan_to_num(color, 0, 0).permute(1, 2, 0)
    render_dist = torch.nan_to_num(allmap[6:7], 0, 0).permute(1, 2, 0)
    # :36
    depth_mask = (gt_depth_hw1 > 1e-5).view(-1) & (render_depth > 1e-5).view(-1)
    # :52-58 (use_edge_growth False: colour mask = depth mask; True: alpha > edge_thres, :53)
    color_mask = depth_mask
    if use_edge_growth:
        render_alpha = torch.nan_to_num(allmap[1:2], 0, 0).permute(1, 2, 0)        # :22
        color_mask = (render_alpha > edge_thres).reshape(-1)
    l_color = (torch.abs(render_color - gt_color_hwc).view(-1, 3))[color_mask].mean()
    l_depth = (torch.abs(render_depth - gt_depth_hw1).view(-1, 1))[depth_mask].mean()
    l_dist = (render_dist.reshape(-1, 1)[color_mask]).mean()
    return weights[0] * l_color + weights[1] * l_depth + weights[2] * l_dist


def tracking_loss_torch(color, allmap, gt_color_hwc, gt_depth_hw1, weights=(0.5, 1.0), eps=1e-6, depth_near=1e-2,
                        depth_far=1e2, silmask_th=0.9):
    """slam/Loss.py:22-47, tracking=True, ignore_outliners False, use_normal_loss False."""
    import torch
    render_depth = allmap[0:1] / (allmap[1:2] + eps)
    outliner_mask = torch.logical_or(render_depth > depth_far, render_depth < depth_near)
    render_depth = render_depth.clone()
    render_depth[outliner_mask] = 0
    render_alpha = torch.nan_to_num(allmap[1:2], 0, 0).permute(1, 2, 0)
    render_depth = torch.nan_to_num(render_depth, 0, 0).permute(1, 2, 0)
    render_color = torch.nan_to_num(color, 0, 0).permute(1, 2, 0)
    depth_mask = (gt_depth_hw1 > 1e-5).view(-1) & (render_depth > 1e-5).view(-1)
    alpha_mask = (render_alpha > silmask_th).view(-1)
    m = depth_mask & alpha_mask
    l_color = torch.abs(render_color - gt_color_hwc).view(-1, 3)[m].sum()
    l_depth = torch.abs(render_depth - gt_depth_hw1).view(-1, 1)[m].sum()
    return weights[0] * l_color + weights[1] * l_depth


def mapping_loss_numpy(color, allmap, gt_color_chw, gt_depth, weights=(0.5, 1.0, 0.1), eps=1e-6, depth_near=1e-2,
                       depth_far=1e2, grad_out=1.0, tracking=False, silmask_th=0.9):
    """float64; returns (loss, dL_dcolor [3,H,W], dL_dallmap [7,H,W])."""
    c = color.astype(np.float64); a = allmap.astype(np.float64)
    gc = gt_color_chw.astype(np.float64); gd = gt_depth.astype(np.float64).reshape(a.shape[1:])
    lowest = float(np.finfo(np.float32).min)

    def n2n(x):
        ok = np.isfinite(x)
        return np.where(ok, x, np.where(np.isneginf(x), lowest, 0.0)), ok

    with np.errstate(all="ignore"):
        inv = 1.0 / (a[1].astype(np.float32) + np.float32(eps)).astype(np.float64)
        d = a[0] * inv
        d = np.where((d > depth_far) | (d < depth_near), 0.0, d)
    d, _ = n2n(d)
    cc, c_ok = n2n(c)
    r, r_ok = n2n(a[6])
    m = (gd > 1e-5) & (d > 1e-5)
    if tracking:         # sums over (depth mask & alpha mask), no distortion term
        with np.errstate(all="ignore"):
            m = m & np.isfinite(a[1]) & (a[1] > silmask_th)
        loss = weights[0] * (np.abs(cc - gc) * m).sum() + weights[1] * (np.abs(d - gd) * m).sum()
        dc = grad_out * weights[0] * np.sign(cc - gc) * m * c_ok
        dd = grad_out * weights[1] * np.sign(d - gd) * m
        da = np.zeros_like(a)
        da[0] = dd * inv
        da[1] = -dd * d * inv
        return float(loss), dc, da
    n = float(m.sum())
    if n == 0:
        return float("nan"), np.zeros_like(c), np.zeros_like(a)
    loss = weights[0] * (np.abs(cc - gc) * m).sum() / (3 * n) + weights[1] * (np.abs(d - gd) * m).sum() / n + \
        weights[2] * (r * m).sum() / n
    dc = grad_out * weights[0] / (3 * n) * np.sign(cc - gc) * m * c_ok
    dd = grad_out * weights[1] / n * np.sign(d - gd) * m
    da = np.zeros_like(a)
    da[0] = dd * inv
    da[1] = -dd * d * inv
    da[6] = grad_out * weights[2] / n * m * r_ok
    return float(loss), dc, da
