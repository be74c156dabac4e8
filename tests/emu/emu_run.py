"""TEST INFRASTRUCTURE: drive the C ABI of include/gs2d.h on HOST memory against libgs2d_emu_full.so (the product's
sources built for the host by tests/emu/build_emu.py).  The call sequence is the one gaus_2dgs_rasterization/_C.py
makes on a GPU (sizes -> gs2d_forward_preprocess -> read num_rendered -> gs2d_forward_render -> gs2d_backward, state
through gs2d_state_view_get); buffers are numpy arrays filled with garbage first, as torch.empty leaves them.
Returns the same dict of arrays as tests/gpu_run.py: run_ours."""
from __future__ import annotations

import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import build_emu  # noqa: E402
from gs2d_b200 import _lib as L  # noqa: E402

GRAD_NAMES = ("means2D", "colors", "opacity", "means3D", "transMat", "sh", "scales", "rotations")
_lib = None


def library():
    global _lib
    if _lib is None:
        lib = C.CDLL(build_emu.build_full())
        for name, (restype, argtypes) in L.PROTOTYPES.items():
            if hasattr(lib, name):
                fn = getattr(lib, name)
                fn.restype, fn.argtypes = restype, argtypes
        _lib = lib
    return _lib


def _ptr(a):
    return None if a is None or a.size == 0 else a.ctypes.data_as(C.c_void_p)


def _garbage(nbytes):
    """256-byte aligned uint8 view over memory filled with 0xCD."""
    raw = np.full(int(nbytes) + 512, 0xCD, np.uint8)
    off = (-raw.ctypes.data) % 256
    return raw[off:off + int(nbytes)]


def _view(ptr, shape, dtype):
    n = int(np.prod(shape))
    if not n:
        return np.zeros(shape, dtype)
    buf = (C.c_byte * (n * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).reshape(shape).copy()


def _check(rc, what):
    if rc != L.GS2D_OK:
        raise RuntimeError(f"{what} -> {rc}")


def run_ours(case, debug=True, backward=True):
    lib = library()
    f32 = lambda a: None if a is None else np.ascontiguousarray(a, np.float32)   # noqa: E731
    P, W, H = case["means3D"].shape[0], case["W"], case["H"]
    M = 0 if case["shs"] is None else case["shs"].shape[1]
    s = L.Settings(P=P, image_width=W, image_height=H, sh_degree=int(case["sh_degree"]), sh_coeffs=M,
                   use_sa=int(bool(case["use_sa"])), prefiltered=0, debug=int(debug), tanfovx=float(case["tanfovx"]),
                   tanfovy=float(case["tanfovy"]), scale_modifier=float(case["scale_modifier"]), reserved_=0.0)
    k = {n: f32(case[n]) for n in ("bg", "means3D", "shs", "colors", "opacities", "scales", "rotations", "transMat",
                                   "viewmatrix", "projmatrix", "campos", "dL_dcolor", "dL_dothers")}
    inp = L.Inputs(background=_ptr(k["bg"]), means3D=_ptr(k["means3D"]), shs=_ptr(k["shs"]), colors_precomp=_ptr(k["colors"]),
                   opacities=_ptr(k["opacities"]), scales=_ptr(k["scales"]), rotations=_ptr(k["rotations"]),
                   transMat_precomp=_ptr(k["transMat"]), viewmatrix=_ptr(k["viewmatrix"]), projmatrix=_ptr(k["projmatrix"]),
                   campos=_ptr(k["campos"]))
    geom = _garbage(lib.gs2d_geom_bytes(P, M) if P else 0)
    img = _garbage(lib.gs2d_image_bytes(W, H))
    bufs = L.Buffers(geom=_ptr(geom), geom_bytes=geom.size, binning=None, binning_bytes=0, image=_ptr(img), image_bytes=img.size)
    radii = np.full(P, -7, np.int32)
    n_host = np.zeros(1, np.int64)
    _check(lib.gs2d_forward_preprocess(C.byref(s), C.byref(inp), C.byref(bufs), _ptr(radii), _ptr(n_host), None), "forward_preprocess")
    n = int(n_host[0]) if P else 0
    binning = _garbage(lib.gs2d_binning_bytes(n, W, H) if n else 0)
    bufs.binning, bufs.binning_bytes = _ptr(binning), binning.size
    color, others = np.full((3, H, W), np.nan, np.float32), np.full((7, H, W), np.nan, np.float32)
    pairs = np.full(1, -1, np.int64)
    _check(lib.gs2d_forward_render(C.byref(s), C.byref(inp), C.byref(bufs), n, _ptr(color), _ptr(others), _ptr(pairs), None),
           "forward_render")
    out = dict(num_rendered=np.int64(n), color=color, others=others, radii=radii, num_pairs=int(pairs[0]))
    sv = L.StateView()
    _check(lib.gs2d_state_view_get(C.byref(s), C.byref(bufs), n, C.byref(sv)), "state_view")
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    counts = (C.c_uint64 * 3)()
    _check(lib.gs2d_footprint_audit(C.byref(s), C.byref(bufs), n, counts, None), "footprint_audit")
    out["footprint_audit"] = np.array(list(counts), dtype=np.int64)
    if P:
        splat = _view(sv.splat, (P, 20), np.float32)
        out.update(tiles_touched=_view(sv.tiles_touched, (P,), np.uint32), point_offsets=_view(sv.point_offsets, (P,), np.uint32),
                   depths=_view(sv.depths, (P,), np.float32),
                   transMat=np.concatenate([splat[:, 0:3], splat[:, 4:7], splat[:, 8:11]], axis=1),
                   means2D=np.stack([splat[:, 3], splat[:, 7]], axis=1),
                   normal_opacity=np.concatenate([splat[:, 12:15], splat[:, 11:12]], axis=1), rgb=splat[:, 16:19].copy())
    if n:
        out.update(keys_sorted=_view(sv.keys_sorted, (n,), np.uint64), point_list=_view(sv.point_list, (n,), np.uint32))
        if debug:
            out.update(keys_unsorted=_view(sv.keys_unsorted, (n,), np.uint64), values_unsorted=_view(sv.values_unsorted, (n,), np.uint32))
    out.update(ranges=_view(sv.ranges, (tiles, 2), np.uint32), n_contrib=_view(sv.n_contrib, (2, H, W), np.uint32),
               final_T=_view(sv.final_T, (3, H, W), np.float32), median_depth=_view(sv.median_depth, (H, W), np.float32),
               depth_std=_view(sv.depth_std, (H, W), np.float32))
    if backward:
        g = dict(means2D=np.full((P, 3), np.nan, np.float32), colors=np.full((P, 3), np.nan, np.float32),
                 opacity=np.full((P, 1), np.nan, np.float32), means3D=np.full((P, 3), np.nan, np.float32),
                 transMat=np.full((P, 9), np.nan, np.float32), sh=np.full((P, M, 3), np.nan, np.float32),
                 scales=np.full((P, 2), np.nan, np.float32), rotations=np.full((P, 4), np.nan, np.float32))
        if P:
            num_pairs = max(int(pairs[0]), 0) if n else 0
            scratch = _garbage(lib.gs2d_backward_scratch_bytes(P, n, num_pairs, W, H))
            grads = L.Grads(dL_dmeans2D=_ptr(g["means2D"]), dL_dcolors=_ptr(g["colors"]), dL_dopacity=_ptr(g["opacity"]),
                            dL_dmeans3D=_ptr(g["means3D"]), dL_dtransMat=_ptr(g["transMat"]), dL_dsh=_ptr(g["sh"]),
                            dL_dscales=_ptr(g["scales"]), dL_drotations=_ptr(g["rotations"]))
            _check(lib.gs2d_backward(C.byref(s), C.byref(inp), C.byref(bufs), _ptr(radii), n, num_pairs, _ptr(k["dL_dcolor"]),
                                     _ptr(k["dL_dothers"]), C.byref(grads), _ptr(scratch), None), "backward")
        for nm in GRAD_NAMES:
            out["grad_" + nm] = g[nm]
    return out
