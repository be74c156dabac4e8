"""Run one parity case through the product (libgs2d.so via the drop-in `_C` module) or through the compiled
reference (oracle/_ref) on the GPU and return every stage as numpy arrays with the SAME key names, so
that tests/ and tests/golden/make_golden.py can compare them field by field.  GPU only.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

import cases  # noqa: F401  (sets sys.path)
from gs2d_b200 import _lib as L
import gaus_2dgs_rasterization._C as GC

GRAD_NAMES = ["means2D", "colors", "opacity", "means3D", "transMat", "sh", "scales", "rotations"]


def _t(a, dev):
    if a is None:
        return torch.empty(0, device=dev)
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)


def _raw(ptr, shape, dtype, dev):
    """Copy a raw device array (pointer from gs2d_state_view) into a torch tensor."""
    n = int(np.prod(shape)) if len(shape) else 1
    out = torch.empty(shape, dtype=dtype, device=dev)
    if n == 0 or not ptr:
        return out
    torch.cuda.synchronize()
    rt = C.CDLL("libcudart.so.12") if not hasattr(_raw, "rt") else _raw.rt
    _raw.rt = rt
    rt.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
    rc = rt.cudaMemcpy(C.c_void_p(out.data_ptr()), C.c_void_p(ptr), C.c_size_t(n * out.element_size()), 3)
    assert rc == 0, f"cudaMemcpy failed: {rc}"
    return out


def to_device(case, dev="cuda:0"):
    dev = torch.device(dev)
    return {k: _t(case[k], dev) for k in ("bg", "means3D", "colors", "opacities", "scales", "rotations", "transMat",
                                          "viewmatrix", "projmatrix", "shs", "campos", "dL_dcolor", "dL_dothers")}


def run_ours(case, dev="cuda:0", debug=True, backward=True):
    """-> dict of numpy arrays: every stage of forward (+ gradients)."""
    dev = torch.device(dev)
    lib = L.load()
    d = to_device(case, dev)
    W, H, P = case["W"], case["H"], case["means3D"].shape[0]
    use_sa = bool(case["use_sa"])
    n, color, others, radii, geom, binning, img = GC.rasterize_gaussians(
        d["bg"], d["means3D"], d["colors"], d["opacities"], d["scales"], d["rotations"], float(case["scale_modifier"]),
        d["transMat"], d["viewmatrix"], d["projmatrix"], float(case["tanfovx"]), float(case["tanfovy"]), H, W,
        d["shs"], int(case["sh_degree"]), d["campos"], use_sa, False, debug)
    torch.cuda.synchronize()
    out = dict(num_rendered=np.int64(n), color=color.cpu().numpy(), others=others.cpu().numpy(),
               radii=radii.cpu().numpy())
    M = 0 if case["shs"] is None else case["shs"].shape[1]
    s = GC._settings(P, W, H, int(case["sh_degree"]), M, use_sa, False, debug, float(case["tanfovx"]),
                     float(case["tanfovy"]), float(case["scale_modifier"]))
    bufs = L.Buffers(geom=geom.data_ptr() if P else None, geom_bytes=geom.numel(),
                     binning=binning.data_ptr() if n else None, binning_bytes=binning.numel(),
                     image=img.data_ptr(), image_bytes=img.numel())
    sv = L.StateView()
    L.check(lib.gs2d_state_view_get(C.byref(s), C.byref(bufs), n, C.byref(sv)), "state_view")
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    i32, f32 = torch.int32, torch.float32
    counts = (C.c_uint64 * 3)()
    L.check(lib.gs2d_footprint_audit(C.byref(s), C.byref(bufs), n, counts, None), "footprint_audit")
    out["footprint_audit"] = np.array(list(counts), dtype=np.int64)   # [exact hits, mask bits, missed hits]
    if P:
        splat = _raw(sv.splat, (P, 20), f32, dev).cpu().numpy()
        out.update(
            tiles_touched=_raw(sv.tiles_touched, (P,), i32, dev).cpu().numpy().view(np.uint32),
            point_offsets=_raw(sv.point_offsets, (P,), i32, dev).cpu().numpy().view(np.uint32),
            depths=_raw(sv.depths, (P,), f32, dev).cpu().numpy(),
            transMat=np.concatenate([splat[:, 0:3], splat[:, 4:7], splat[:, 8:11]], axis=1),
            means2D=np.stack([splat[:, 3], splat[:, 7]], axis=1),
            normal_opacity=np.concatenate([splat[:, 12:15], splat[:, 11:12]], axis=1),
            rgb=splat[:, 16:19].copy())
    if n:
        out.update(
            keys_sorted=_raw(sv.keys_sorted, (n,), torch.int64, dev).cpu().numpy().view(np.uint64),
            point_list=_raw(sv.point_list, (n,), i32, dev).cpu().numpy().view(np.uint32))
        if debug:
            out.update(
                keys_unsorted=_raw(sv.keys_unsorted, (n,), torch.int64, dev).cpu().numpy().view(np.uint64),
                values_unsorted=_raw(sv.values_unsorted, (n,), i32, dev).cpu().numpy().view(np.uint32))
    out.update(
        ranges=_raw(sv.ranges, (tiles, 2), i32, dev).cpu().numpy().view(np.uint32),
        n_contrib=_raw(sv.n_contrib, (2, H, W), i32, dev).cpu().numpy().view(np.uint32),
        final_T=_raw(sv.final_T, (3, H, W), f32, dev).cpu().numpy(),
        median_depth=_raw(sv.median_depth, (H, W), f32, dev).cpu().numpy(),
        depth_std=_raw(sv.depth_std, (H, W), f32, dev).cpu().numpy())
    if backward:
        g = GC.rasterize_gaussians_backward(
            d["bg"], d["means3D"], radii, d["colors"], d["scales"], d["rotations"], float(case["scale_modifier"]),
            d["transMat"], d["viewmatrix"], d["projmatrix"], float(case["tanfovx"]), float(case["tanfovy"]),
            d["dL_dcolor"], d["dL_dothers"], d["shs"], int(case["sh_degree"]), d["campos"], geom, n, binning, img,
            use_sa, debug)
        torch.cuda.synchronize()
        for nm, t in zip(GRAD_NAMES, g):
            out["grad_" + nm] = t.cpu().numpy()
    return out


def run_ref(case, dev="cuda:0"):
    """Same dict from the UNMODIFIED reference CUDA rasterizer (oracle/_ref/libgs2d_ref.so)."""
    import refgpu
    dev = torch.device(dev)
    d = to_device(case, dev)
    W, H = case["W"], case["H"]
    nz = lambda t: t if t.numel() else None
    kw = dict(bg=d["bg"], means3D=d["means3D"], viewmatrix=d["viewmatrix"], projmatrix=d["projmatrix"],
              campos=d["campos"], W=W, H=H, tanfovx=float(case["tanfovx"]), tanfovy=float(case["tanfovy"]),
              colors=nz(d["colors"]), shs=nz(d["shs"]), scales=nz(d["scales"]), rotations=nz(d["rotations"]),
              transMat=nz(d["transMat"]), scale_modifier=float(case["scale_modifier"]),
              sh_degree=int(case["sh_degree"]), use_sa=bool(case["use_sa"]))
    ref = refgpu.RefRasterizer()
    n, color, others, radii = ref.forward(opacities=d["opacities"], **kw)
    torch.cuda.synchronize()
    out = dict(num_rendered=np.int64(n), color=color.cpu().numpy(), others=others.cpu().numpy(),
               radii=radii.cpu().numpy())
    st = ref.state()
    u32 = {"tiles_touched", "point_offsets", "values_unsorted", "point_list", "ranges", "n_contrib"}
    for k, v in st.items():
        a = v.cpu().numpy()
        if k in u32:
            a = a.view(np.uint32)
        if k in ("keys_unsorted", "keys_sorted"):
            a = a.view(np.uint64)
        out[k] = a
    g = ref.backward(radii=radii, dL_dcolor=d["dL_dcolor"], dL_dothers=d["dL_dothers"], **kw)
    torch.cuda.synchronize()
    for nm in GRAD_NAMES:
        out["grad_" + nm] = g[nm].cpu().numpy()
    return out
