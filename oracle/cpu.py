"""oracle/cpu.py -- numpy/ctypes front end of the CPU restatement (oracle/gs2d_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
reference legs, never by the product package.  `forward()` / `backward()` chain the stages in the
order of CudaRasterizer::Rasterizer::forward / backward (CR/rasterizer_impl.cu:201-350, :354-460)
and return every intermediate array, so tests can compare stage by stage.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "_build", "libgs2d_oracle.so")
_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(SO):
            import importlib.util
            spec = importlib.util.spec_from_file_location("oracle_build", os.path.join(HERE, "build.py"))
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            mod.build_oracle()
        _lib = C.CDLL(SO)
        _lib.oracle_scan.restype = C.c_int64
        _lib.oracle_higher_msb.restype = C.c_uint32
        _lib.oracle_num_threads.restype = C.c_int
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def num_threads() -> int:
    return int(load().oracle_num_threads())


def set_num_threads(n: int) -> None:
    load().oracle_set_num_threads(C.c_int(n))


def higher_msb(n: int) -> int:
    return int(load().oracle_higher_msb(C.c_uint32(n)))


def tiles(W, H):
    return ((W + 15) // 16) * ((H + 15) // 16)


def preprocess(means3D, opacities, viewmatrix, projmatrix, W, H, scales=None, rotations=None, transMat=None,
               colors=None, shs=None, sh_degree=0, campos=None, scale_modifier=1.0):
    lib = load()
    means3D = _f32(means3D); P = means3D.shape[0]
    scales, rotations, transMat, colors, shs = map(_f32, (scales, rotations, transMat, colors, shs))
    opacities = _f32(opacities); viewmatrix = _f32(viewmatrix); projmatrix = _f32(projmatrix)
    campos = _f32(campos if campos is not None else np.zeros(3))
    M = 0 if shs is None else shs.shape[1]
    st = dict(
        radii=np.zeros(P, np.int32), means2D=np.zeros((P, 2), np.float32), depths=np.zeros(P, np.float32),
        transMat=np.zeros((P, 9), np.float32), rgb=np.zeros((P, 3), np.float32),
        normal_opacity=np.zeros((P, 4), np.float32), tiles_touched=np.zeros(P, np.uint32),
        clamped=np.zeros((P, 3), np.uint8))
    lib.oracle_preprocess(C.c_int(P), C.c_int(sh_degree), C.c_int(M), _p(means3D), _p(scales),
                          C.c_float(scale_modifier), _p(rotations), _p(opacities), _p(shs), _p(st["clamped"]),
                          _p(transMat), _p(colors), _p(viewmatrix), _p(projmatrix), _p(campos), C.c_int(W), C.c_int(H),
                          _p(st["radii"]), _p(st["means2D"]), _p(st["depths"]), _p(st["transMat"]), _p(st["rgb"]),
                          _p(st["normal_opacity"]), _p(st["tiles_touched"]))
    return st


def binning(st, W, H):
    """scan + duplicate + sort + ranges from the preprocess state (dict with radii, means2D, depths, tiles_touched)."""
    lib = load()
    P = st["radii"].shape[0]
    tt = np.ascontiguousarray(st["tiles_touched"], dtype=np.uint32)
    offs = np.zeros(P, np.uint32)
    n = int(lib.oracle_scan(C.c_int(P), _p(tt), _p(offs)))
    ku = np.zeros(n, np.uint64); vu = np.zeros(n, np.uint32)
    ks = np.zeros(n, np.uint64); pl = np.zeros(n, np.uint32)
    m2d = _f32(st["means2D"]); dep = _f32(st["depths"]); radii = np.ascontiguousarray(st["radii"], dtype=np.int32)
    lib.oracle_duplicate(C.c_int(P), C.c_int(W), C.c_int(H), _p(m2d), _p(dep), _p(offs), _p(radii), _p(ku), _p(vu))
    end_bit = 32 + higher_msb(tiles(W, H))
    lib.oracle_sort_pairs(C.c_int64(n), _p(ku), _p(vu), _p(ks), _p(pl), C.c_int(end_bit))
    ranges = np.zeros((tiles(W, H), 2), np.uint32)
    lib.oracle_tile_ranges(C.c_int64(n), _p(ks), C.c_int(W), C.c_int(H), _p(ranges))
    return dict(num_rendered=n, point_offsets=offs, keys_unsorted=ku, values_unsorted=vu, keys_sorted=ks,
                point_list=pl, ranges=ranges, end_bit=end_bit)


def sort_pairs(keys, vals, end_bit):
    lib = load()
    keys = np.ascontiguousarray(keys, dtype=np.uint64); vals = np.ascontiguousarray(vals, dtype=np.uint32)
    ko = np.zeros_like(keys); vo = np.zeros_like(vals)
    lib.oracle_sort_pairs(C.c_int64(keys.shape[0]), _p(keys), _p(vals), _p(ko), _p(vo), C.c_int(end_bit))
    return ko, vo


def render_forward(W, H, ranges, point_list, means2D, features, transMat, normal_opacity, bg, use_sa=True):
    lib = load()
    out = dict(final_T=np.zeros((3, H, W), np.float32), n_contrib=np.zeros((2, H, W), np.uint32),
               color=np.zeros((3, H, W), np.float32), others=np.zeros((7, H, W), np.float32),
               median_depth=np.zeros((H, W), np.float32), depth_std=np.zeros((H, W), np.float32))
    ranges = np.ascontiguousarray(ranges, dtype=np.uint32); point_list = np.ascontiguousarray(point_list, dtype=np.uint32)
    a = [_f32(x) for x in (means2D, features, transMat, normal_opacity, bg)]
    lib.oracle_render_forward(C.c_int(W), C.c_int(H), _p(ranges), _p(point_list), _p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]),
                              _p(a[4]), C.c_int(int(use_sa)), _p(out["final_T"]), _p(out["n_contrib"]), _p(out["color"]),
                              _p(out["others"]), _p(out["median_depth"]), _p(out["depth_std"]))
    return out


def forward(means3D, opacities, viewmatrix, projmatrix, W, H, bg, scales=None, rotations=None, transMat=None,
            colors=None, shs=None, sh_degree=0, campos=None, scale_modifier=1.0, use_sa=True):
    """Whole forward pass; returns a dict with every stage's arrays."""
    st = preprocess(means3D, opacities, viewmatrix, projmatrix, W, H, scales, rotations, transMat, colors, shs,
                    sh_degree, campos, scale_modifier)
    st.update(binning(st, W, H))
    feats = _f32(colors) if colors is not None else st["rgb"]
    tm = _f32(transMat) if transMat is not None else st["transMat"]
    st.update(render_forward(W, H, st["ranges"], st["point_list"], st["means2D"], feats, tm, st["normal_opacity"],
                             bg, use_sa))
    st["features"] = feats
    st["transMat_used"] = tm
    return st


def backward(st, means3D, viewmatrix, projmatrix, W, H, tanfovx, tanfovy, bg, dL_dcolor, dL_dothers, scales=None,
             rotations=None, transMat=None, shs=None, sh_degree=0, campos=None, use_sa=True):
    """Whole backward pass from a forward() state; returns the reference's gradient tuple as a dict."""
    lib = load()
    means3D = _f32(means3D); P = means3D.shape[0]
    scales, rotations, transMat, shs = map(_f32, (scales, rotations, transMat, shs))
    viewmatrix = _f32(viewmatrix); projmatrix = _f32(projmatrix); bg = _f32(bg)
    campos = _f32(campos if campos is not None else np.zeros(3))
    dL_dcolor = _f32(dL_dcolor); dL_dothers = _f32(dL_dothers)
    M = 0 if shs is None else shs.shape[1]
    g = dict(means2D=np.zeros((P, 3), np.float32), colors=np.zeros((P, 3), np.float32),
             opacity=np.zeros((P, 1), np.float32), means3D=np.zeros((P, 3), np.float32),
             transMat=np.zeros((P, 9), np.float32), sh=np.zeros((P, M, 3), np.float32),
             scales=np.zeros((P, 2), np.float32), rotations=np.zeros((P, 4), np.float32))
    dnormal = np.zeros((P, 3), np.float32)
    lib.oracle_render_backward(C.c_int(W), C.c_int(H), _p(st["ranges"]), _p(st["point_list"]), _p(bg), _p(st["means2D"]),
                               _p(st["normal_opacity"]), _p(st["transMat_used"]), _p(st["features"]), _p(st["final_T"]),
                               _p(st["n_contrib"]), _p(dL_dcolor), _p(dL_dothers), _p(st["median_depth"]),
                               _p(st["depth_std"]), C.c_int(int(use_sa)), _p(g["transMat"]), _p(g["means2D"]), _p(dnormal),
                               _p(g["opacity"]), _p(g["colors"]))
    g["normal3D"] = dnormal
    focal_y = np.float32(H) / (np.float32(2.0) * np.float32(tanfovy))
    focal_x = np.float32(W) / (np.float32(2.0) * np.float32(tanfovx))
    lib.oracle_preprocess_backward(C.c_int(P), C.c_int(sh_degree), C.c_int(M), _p(means3D), _p(st["transMat_used"]),
                                   _p(st["radii"]), _p(shs), _p(st["clamped"]), _p(scales), _p(rotations), _p(viewmatrix),
                                   _p(projmatrix), C.c_float(focal_x), C.c_float(focal_y), C.c_float(tanfovx),
                                   C.c_float(tanfovy), _p(campos), _p(g["transMat"]), _p(dnormal), _p(g["colors"]),
                                   _p(g["sh"]), _p(g["means2D"]), _p(g["means3D"]), _p(g["scales"]), _p(g["rotations"]))
    return g


def mark_visible(means3D, viewmatrix):
    lib = load()
    means3D = _f32(means3D); viewmatrix = _f32(viewmatrix)
    out = np.zeros(means3D.shape[0], np.uint8)
    lib.oracle_mark_visible(C.c_int(means3D.shape[0]), _p(means3D), _p(viewmatrix), _p(out))
    return out.astype(bool)


def dist2(points):
    lib = load()
    points = _f32(points)
    out = np.zeros(points.shape[0], np.float32)
    lib.oracle_dist2(C.c_int(points.shape[0]), _p(points), _p(out))
    return out
