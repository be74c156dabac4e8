"""oracle/refgpu.py -- ctypes access to oracle/_ref/libgs2d_ref.so.  TEST INFRASTRUCTURE ONLY.

libgs2d_ref.so is the UNMODIFIED reference CUDA rasterizer + simple-knn compiled for sm_100a by
oracle/build.py, behind the C shim oracle/ref_shim.cu.  It is the primary parity oracle on the GPU
box and the timed "reference" arm of bench.py.  Nothing in the product path may import this.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libgs2d_ref.so")


class RefStatePtrs(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "depths", "internal_radii", "means2D", "transMat", "normal_opacity", "tiles_touched",
        "point_offsets", "point_list_keys_unsorted", "point_list_keys", "point_list_unsorted",
        "point_list", "ranges", "n_contrib", "accum_alpha", "median_depth", "depth_std")]


_lib = None


def available() -> bool:
    return os.path.exists(REF_SO)


def load():
    global _lib
    if _lib is None:
        lib = C.CDLL(REF_SO)
        lib.ref_create.restype = C.c_void_p
        lib.ref_destroy.argtypes = [C.c_void_p]
        lib.ref_forward.restype = C.c_int
        lib.ref_forward.argtypes = ([C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int] +
                                    [C.c_void_p] * 5 + [C.c_float] + [C.c_void_p] * 5 +
                                    [C.c_float, C.c_float, C.c_int] + [C.c_void_p] * 3 + [C.c_int, C.c_int])
        lib.ref_state.restype = C.c_int
        lib.ref_state.argtypes = [C.c_void_p, C.POINTER(RefStatePtrs)]
        lib.ref_backward.restype = None
        lib.ref_backward.argtypes = ([C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int] +
                                     [C.c_void_p] * 4 + [C.c_float] + [C.c_void_p] * 5 +
                                     [C.c_float, C.c_float] + [C.c_void_p] * 11 + [C.c_int, C.c_int])
        lib.ref_mark_visible.argtypes = [C.c_int] + [C.c_void_p] * 4
        lib.ref_dist_cuda2.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        lib.ref_last_error.restype = C.c_int
        lib.ref_copy_d2d.restype = C.c_int
        lib.ref_copy_d2d.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        _lib = lib
    return _lib


def _p(t):
    return None if t is None or t.numel() == 0 else t.data_ptr()


def _view(ptr, shape, dtype, device):
    """Clone a raw device array into a torch tensor (copy, so it survives buffer reuse)."""
    n = 1
    for s in shape:
        n *= s
    out = torch.empty(shape, dtype=dtype, device=device)
    if n == 0 or not ptr:
        return out
    nbytes = n * out.element_size()
    torch.cuda.synchronize()
    rc = load().ref_copy_d2d(C.c_void_p(out.data_ptr()), C.c_void_p(ptr), C.c_size_t(nbytes))
    assert rc == 0, f"cudaMemcpy failed: {rc}"
    return out


class RefRasterizer:
    """One reference context (its three byte buffers are reused between calls, like a cached allocator)."""

    def __init__(self):
        self.lib = load()
        self.h = self.lib.ref_create()
        self.last = None

    def __del__(self):
        try:
            self.lib.ref_destroy(self.h)
        except Exception:
            pass

    def forward(self, *, bg, means3D, opacities, viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy,
                colors=None, shs=None, scales=None, rotations=None, transMat=None, scale_modifier=1.0,
                sh_degree=0, use_sa=True, prefiltered=False, debug=False, out=None):
        # raw pointers go to the reference kernels: every tensor must be dense (RAST/rasterize_points.cu calls
        # .contiguous() on its inputs too)
        c = lambda x: x.contiguous() if isinstance(x, torch.Tensor) else x   # noqa: E731
        bg, means3D, opacities, viewmatrix, projmatrix, campos = (c(x) for x in (bg, means3D, opacities, viewmatrix, projmatrix, campos))
        colors, shs, scales, rotations, transMat = (c(x) for x in (colors, shs, scales, rotations, transMat))
        dev = means3D.device
        P = means3D.shape[0]
        M = shs.shape[1] if shs is not None and shs.numel() else 0
        if out is None:
            color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
            others = torch.empty((7, H, W), dtype=torch.float32, device=dev)
            radii = torch.empty((P,), dtype=torch.int32, device=dev)
        else:
            color, others, radii = out
        n = self.lib.ref_forward(self.h, P, sh_degree, M, _p(bg), W, H, _p(means3D), _p(shs), _p(colors),
                                 _p(opacities), _p(scales), scale_modifier, _p(rotations), _p(transMat),
                                 _p(viewmatrix), _p(projmatrix), _p(campos), tanfovx, tanfovy, int(prefiltered),
                                 _p(color), _p(others), _p(radii), int(use_sa), int(debug))
        self.last = dict(P=P, W=W, H=H, M=M, n=n, dev=dev)
        return n, color, others, radii

    def state(self):
        """Copies of the reference's internal arrays (binning + per-pixel state)."""
        L = self.last
        P, W, H, n, dev = L["P"], L["W"], L["H"], L["n"], L["dev"]
        sp = RefStatePtrs()
        assert self.lib.ref_state(self.h, C.byref(sp)) == 0
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        i32, u8 = torch.int32, torch.uint8
        st = dict(
            depths=_view(sp.depths, (P,), torch.float32, dev),
            means2D=_view(sp.means2D, (P, 2), torch.float32, dev),
            transMat=_view(sp.transMat, (P, 9), torch.float32, dev),
            normal_opacity=_view(sp.normal_opacity, (P, 4), torch.float32, dev),
            tiles_touched=_view(sp.tiles_touched, (P,), i32, dev),
            point_offsets=_view(sp.point_offsets, (P,), i32, dev),
            keys_unsorted=_view(sp.point_list_keys_unsorted, (n,), torch.int64, dev),
            keys_sorted=_view(sp.point_list_keys, (n,), torch.int64, dev),
            values_unsorted=_view(sp.point_list_unsorted, (n,), i32, dev),
            point_list=_view(sp.point_list, (n,), i32, dev),
            ranges=_view(sp.ranges, (tiles, 2), i32, dev),
            n_contrib=_view(sp.n_contrib, (2, H, W), i32, dev),
            final_T=_view(sp.accum_alpha, (3, H, W), torch.float32, dev),
            median_depth=_view(sp.median_depth, (H, W), torch.float32, dev),
            depth_std=_view(sp.depth_std, (H, W), torch.float32, dev),
        )
        return st

    def backward(self, *, bg, means3D, radii, viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy,
                 dL_dcolor, dL_dothers, colors=None, shs=None, scales=None, rotations=None, transMat=None,
                 scale_modifier=1.0, sh_degree=0, use_sa=True, debug=False, out=None):
        c = lambda x: x.contiguous() if isinstance(x, torch.Tensor) else x   # noqa: E731
        bg, means3D, radii, viewmatrix, projmatrix, campos = (c(x) for x in (bg, means3D, radii, viewmatrix, projmatrix, campos))
        dL_dcolor, dL_dothers, colors, shs, scales, rotations, transMat = (c(x) for x in (dL_dcolor, dL_dothers, colors, shs, scales, rotations, transMat))
        dev = means3D.device
        P = means3D.shape[0]
        M = shs.shape[1] if shs is not None and shs.numel() else 0
        f32 = dict(dtype=torch.float32, device=dev)
        if out is None:
            g = dict(means2D=torch.empty((P, 3), **f32), colors=torch.empty((P, 3), **f32),
                     opacity=torch.empty((P, 1), **f32), means3D=torch.empty((P, 3), **f32),
                     transMat=torch.empty((P, 9), **f32), sh=torch.empty((P, M, 3), **f32),
                     scales=torch.empty((P, 2), **f32), rotations=torch.empty((P, 4), **f32))
        else:
            g = out
        self.lib.ref_backward(self.h, P, sh_degree, M, _p(bg), W, H, _p(means3D), _p(shs), _p(colors),
                              _p(scales), scale_modifier, _p(rotations), _p(transMat), _p(viewmatrix),
                              _p(projmatrix), _p(campos), tanfovx, tanfovy, _p(radii), _p(dL_dcolor),
                              _p(dL_dothers), _p(g["means2D"]), _p(g["colors"]), _p(g["opacity"]),
                              _p(g["means3D"]), _p(g["transMat"]), _p(g["sh"]), _p(g["scales"]),
                              _p(g["rotations"]), int(use_sa), int(debug))
        return g


def mark_visible(means3D, viewmatrix, projmatrix):
    lib = load()
    P = means3D.shape[0]
    present = torch.zeros((P,), dtype=torch.bool, device=means3D.device)
    lib.ref_mark_visible(P, _p(means3D), _p(viewmatrix), _p(projmatrix), present.data_ptr())
    return present


def dist_cuda2(points):
    lib = load()
    P = points.shape[0]
    out = torch.empty((P,), dtype=torch.float32, device=points.device)
    lib.ref_dist_cuda2(P, _p(points), out.data_ptr())
    return out
