"""B200-native drop-in for `gaus_2dgs_rasterization` (the package gaus-slam imports in
render/render_2dgs.py:3-4).

Public surface, identical to RAST/gaus_2dgs_rasterization/__init__.py:
  * GaussianRasterizationSettings  -- NamedTuple, 13 fields in the reference's order (:163-176)
  * GaussianRasterizer(nn.Module)  -- forward(means3D, means2D, opacities, shs=, colors_precomp=,
                                      scales=, rotations=, cov3D_precomp=) -> (color, radii, allmap)
                                      and markVisible(positions) (:178-227)
  * rasterize_gaussians(...)       -- functional form (:21-42)
Gradients come back in the reference's order (means3D, means2D, sh, colors_precomp, opacities,
scales, rotations, cov3Ds_precomp, None) (:146-156); `means2D.grad` receives the densification
proxy of CR/backward.cu:660-663.

The compute lives in libgs2d.so (hand-written CUDA for sm_100a, C ABI in include/gs2d.h).
"""
from __future__ import annotations

import os
import sys
from typing import NamedTuple

import torch
import torch.nn as nn

_PKG_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _PKG_ROOT not in sys.path:
    sys.path.insert(0, _PKG_ROOT)

from . import _C  # noqa: E402

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    use_sa: bool
    prefiltered: bool
    debug: bool


def cpu_deep_copy_tuple(input_tuple):
    """Host-side snapshot of a call's arguments (used for the debug dumps)."""
    return tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in input_tuple)


def _call_with_snapshot(fn, args, debug: bool, dump_name: str, what: str):
    # debug=True reproduces the reference's behaviour: keep a CPU copy of the arguments and write
    # it next to the process if the native call throws (__init__.py:84-91, :135-142).
    if not debug:
        return fn(*args)
    snapshot = cpu_deep_copy_tuple(args)
    try:
        return fn(*args)
    except Exception:
        torch.save(snapshot, dump_name)
        print(f"\nAn error occured in {what}. Please forward {dump_name} for debugging.")
        raise


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        rs = raster_settings
        args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height,
                rs.image_width, sh, rs.sh_degree, rs.campos, rs.use_sa, rs.prefiltered, rs.debug)
        num_rendered, color, others, radii, geom, binning, img = _call_with_snapshot(
            _C.rasterize_gaussians, args, rs.debug, "snapshot_fw.dump", "forward")
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              geom, binning, img)
        ctx.mark_non_differentiable(radii)
        # no zero-filled stand-ins for the gradients nobody supplied (radii always, colour / allmap when only one of
        # them feeds the loss): backward() gets None and handles it, which saves a fill launch per step
        ctx.set_materialize_grads(False)
        return color, radii, others

    @staticmethod
    def backward(ctx, grad_out_color, grad_radii, grad_others):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img) = ctx.saved_tensors
        if grad_out_color is None:
            grad_out_color = torch.zeros((3, rs.image_height, rs.image_width), dtype=torch.float32,
                                         device=means3D.device)
        if grad_others is None:
            grad_others = torch.zeros((7, rs.image_height, rs.image_width), dtype=torch.float32,
                                      device=means3D.device)
        args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color, grad_others, sh,
                rs.sh_degree, rs.campos, geom, ctx.num_rendered, binning, img, rs.use_sa, rs.debug)
        (g_means2D, g_colors, g_opacities, g_means3D, g_cov, g_sh, g_scales, g_rotations) = _call_with_snapshot(
            _C.rasterize_gaussians_backward, args, rs.debug, "snapshot_bw.dump", "backward")
        return (g_means3D, g_means2D, g_sh, g_colors, g_opacities, g_scales, g_rotations, g_cov, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings: GaussianRasterizationSettings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        rs = self.raster_settings
        if (shs is None) == (colors_precomp is None):
            raise Exception("Please provide excatly one of either SHs or precomputed colors!")
        have_sr = scales is not None or rotations is not None
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (have_sr and cov3D_precomp is not None):
            raise Exception("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")

        def _or_empty(t):
            return t if t is not None else torch.empty(0, dtype=torch.float32, device=means3D.device)

        return rasterize_gaussians(means3D, means2D, _or_empty(shs), _or_empty(colors_precomp), opacities,
                                   _or_empty(scales), _or_empty(rotations), _or_empty(cov3D_precomp), rs)
