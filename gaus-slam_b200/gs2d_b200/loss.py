"""Fused mapping loss: the mapping branch of the reference's `get_loss`
(slam/Loss.py:6-64) together with the weight normalisation of the rendered depth that `Renderer_*` applies
before it (render/__init__.py:46-49), as ONE forward and ONE backward kernel of libgs2d.so instead of ~25 torch
launches each way.

    loss = mapping_loss(color, allmap, gt_color, gt_depth)          # color [3,H,W], allmap [7,H,W] from the rasterizer
    loss.backward()

Argument meaning follows the reference's config keys: `weights` = config['loss']['mapping'] (color, depth, dist),
`eps`, `depth_near`, `depth_far` = config['render'][...].  `gt_color` may be [3,H,W] or the reference's [H,W,3].
Built for every branch of get_loss but the normal loss (`use_normal_loss`, off in every shipped config: it raises):
`use_edge_growth`, the tracking branch, `ignore_outliners`.  `enable_exposure` is not part of get_loss: the caller
applies `exposure[0] * color + exposure[1]` (render/__init__.py:105-107) to the rendered colour before the loss, as a
torch op, exactly like the reference.  CUDA only: there is no CPU path.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib as L

class _Scratch:
    """Partials, ticket and the forward's mask counts: ~38 KB that belong to ONE forward call and travel to its backward
    in ctx -- several losses may be formed before any backward (the sum over the keyframes of a bundle-adjustment
    window; round 1 shared one scratch per stream, so every backward then used the last count).  The kernel leaves the
    ticket at zero when it is done, so a buffer whose loss object has died goes back to a free list and is handed out
    again as it is: no allocation and no fill launch per step once the pool is warm."""
    _free = {}

    def __init__(self, dev, W, H):
        # (per stream: a buffer is reused in stream order, behind the kernels of its previous owner)
        self.key = (str(dev), L.raw_stream(torch, dev), L.load().gs2d_mapping_loss_scratch_bytes(W, H))
        pool = _Scratch._free.get(self.key)
        self.buf = pool.pop() if pool else torch.zeros(self.key[2], dtype=torch.uint8, device=dev)

    def __del__(self):
        try:
            _Scratch._free.setdefault(self.key, []).append(self.buf)
        except Exception:       # interpreter shutdown
            pass

    def data_ptr(self):
        return self.buf.data_ptr()


def _new_scratch(dev, W, H):
    return _Scratch(dev, W, H)


def _check(t, name, shape):
    if not t.is_cuda or t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be a float32 CUDA tensor")
    if tuple(t.shape) not in shape:
        raise RuntimeError(f"{name} has shape {tuple(t.shape)}, expected one of {shape}")
    return t.contiguous()


class _MappingLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, color, allmap, gt_color, gt_depth, params):
        lib = L.load()
        H, W = int(color.shape[1]), int(color.shape[2])
        dev = color.device
        color = _check(color, "color", [(3, H, W)])
        allmap = _check(allmap, "allmap", [(7, H, W)])
        gt_color = _check(gt_color, "gt_color", [(3, H, W), (H, W, 3)])
        gt_depth = _check(gt_depth, "gt_depth", [(H, W), (1, H, W), (H, W, 1)])
        hwc = int(tuple(gt_color.shape) == (H, W, 3) and (H, W, 3) != (3, H, W))
        lp = L.LossParams(*params)
        with torch.cuda.device(dev):
            scratch = _new_scratch(dev, W, H)
            loss = torch.empty((), dtype=torch.float32, device=dev)
            st = C.c_void_p(L.raw_stream(torch, dev))
            L.check(lib.gs2d_mapping_loss_forward(W, H, color.data_ptr(), allmap.data_ptr(), gt_color.data_ptr(),
                                                  gt_depth.data_ptr(), hwc, C.byref(lp), scratch.data_ptr(),
                                                  loss.data_ptr(), st), "gs2d_mapping_loss_forward")
        ctx.save_for_backward(color, allmap, gt_color, gt_depth)
        ctx.lp, ctx.hwc, ctx.scratch = lp, hwc, scratch
        return loss

    @staticmethod
    def backward(ctx, grad_loss):
        lib = L.load()
        color, allmap, gt_color, gt_depth = ctx.saved_tensors
        H, W = int(color.shape[1]), int(color.shape[2])
        dev = color.device
        g = grad_loss.to(torch.float32).contiguous()
        d_color = torch.empty_like(color)
        d_allmap = torch.empty_like(allmap)
        with torch.cuda.device(dev):
            st = C.c_void_p(L.raw_stream(torch, dev))
            L.check(lib.gs2d_mapping_loss_backward(W, H, color.data_ptr(), allmap.data_ptr(), gt_color.data_ptr(),
                                                   gt_depth.data_ptr(), ctx.hwc, C.byref(ctx.lp),
                                                   ctx.scratch.data_ptr(), g.data_ptr(), d_color.data_ptr(),
                                                   d_allmap.data_ptr(), st), "gs2d_mapping_loss_backward")
        return d_color, d_allmap, None, None, None


def mapping_loss(color, allmap, gt_color, gt_depth, weights=(0.5, 1.0, 0.1), eps=1e-6, depth_near=1e-2,
                 depth_far=1e2, use_edge_growth=False, edge_thres=0.4, use_normal_loss=False):
    """-> scalar loss tensor (differentiable w.r.t. color and allmap).  `use_edge_growth` / `edge_thres` =
    config['densify'][...] (slam/Loss.py:53): colour and distortion are then averaged over alpha > edge_thres instead of
    over the depth mask."""
    if use_normal_loss:
        raise NotImplementedError("the normal loss is not built (off in every shipped config)")
    if use_edge_growth and not float(edge_thres) >= 0.0:
        raise ValueError("edge_thres must be >= 0")
    params = (float(weights[0]), float(weights[1]), float(weights[2]), float(eps), float(depth_near), float(depth_far),
              0, 0.0, float(edge_thres) if use_edge_growth else -1.0)
    return _MappingLoss.apply(color, allmap, gt_color, gt_depth, params)


def _inlier_gt_depth(allmap, gt_depth, eps, depth_near, depth_far):
    """The median-based outlier mask of the tracking loss (slam/Loss.py:37-40, `ignore_outliners`), folded into the
    sensor depth: depth_error = |depth - gt| * depth_mask over ALL pixels, inliers = depth_error < 10 * median.  A pixel
    whose gt depth is set to 0 fails the depth mask of the fused kernels, which removes it from the colour and the
    depth sums exactly like the reference's `inliner_mask & tracking_depth_mask`.  Masks carry no gradient."""
    with torch.no_grad():
        d = allmap[0:1] / (allmap[1:2] + eps)                    # render/__init__.py:46-49
        d = torch.where((d > depth_far) | (d < depth_near), torch.zeros_like(d), d)
        d = torch.nan_to_num(d, 0, 0)
        gd = gt_depth.reshape(d.shape)
        err = (d - gd).abs() * ((gd > 1e-5) & (d > 1e-5))
        return torch.where(err < 10 * err.median(), gd, torch.zeros_like(gd)).reshape(gt_depth.shape)


def tracking_loss(color, allmap, gt_color, gt_depth, weights=(0.5, 1.0), eps=1e-6, depth_near=1e-2, depth_far=1e2,
                  silmask_th=0.9, ignore_outliners=False, use_normal_loss=False):
    """The tracking branch of get_loss (slam/Loss.py:35-47): mask = depth mask [& inliers] & nan_to_num(alpha) >
    silmask_th, loss = weights[0] * SUM |color - gt_color| + weights[1] * SUM |depth - gt_depth| over the mask.  Same two
    kernels as mapping_loss; `ignore_outliners` (configs/replica/splatam.py) adds a handful of torch ops that build the
    inlier mask (a median over the image) in front of them."""
    if use_normal_loss:
        raise NotImplementedError("the normal loss is not built (off in every shipped config)")
    if ignore_outliners:
        gt_depth = _inlier_gt_depth(allmap, gt_depth, float(eps), float(depth_near), float(depth_far))
    params = (float(weights[0]), float(weights[1]), 0.0, float(eps), float(depth_near), float(depth_far), 1,
              float(silmask_th), -1.0)
    return _MappingLoss.apply(color, allmap, gt_color, gt_depth, params)
