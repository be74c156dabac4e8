"""CPU: the SOURCE TEXT of the forward preprocess stage (csrc/preprocess.cu with the real common.cuh / math.cuh /
exact.cuh: preprocess_forward_kernel, the single-pass look-back scan and its tile-range CTA) executed on the host under
tests/emu/cuda_emu.h and held to the golden vectors of the UNMODIFIED compiled reference (tests/golden/*.npz, produced
on a B200): radii, tiles touched, offsets, instance count and tile ranges exactly; depths, transMat, means2D and normals
of the visible surfels BIT FOR BIT.

That a host can reproduce these bits at all is the point of csrc/exact.cuh: every product that decides a ceil() /
truncation is written with explicit round-to-nearest intrinsics in the operation order of the reference's compiled
kernel, which map one-to-one onto the host's correctly rounded fmaf / mul / add / sqrt / div (built with
-ffp-contract=off).  The one operation of that chain a host cannot reproduce is the device's approximate rsqrt
(MUFU.RSQ, what glm::normalize compiles to) that normalises the quaternion: transMat, means2D and the normal inherit
its last bit, so they are compared bit for bit only when the transform is given precomputed (case transmat_pre) and
to 2e-6 of the largest magnitude otherwise (measured: <= 7.2e-7); the integer outputs agree exactly in every case.
A pre-check for edits of the preprocess kernel where there is no GPU; the GPU tests remain the parity tests proper."""
import ctypes as C
import os
import shutil
import sys

import numpy as np
import pytest

import cases
from gs2d_b200 import _lib as L

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))
import build_emu  # noqa: E402

pytestmark = pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++ (C++20) for the host build of the kernels")


@pytest.fixture(scope="module")
def lib():
    lib = C.CDLL(build_emu.build_preprocess())
    lib.emu_forward_preprocess.restype = C.c_int
    lib.emu_forward_preprocess.argtypes = [C.POINTER(L.Settings), C.POINTER(L.Inputs)] + [C.c_void_p] * 8
    return lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _run(lib, case):
    f32 = lambda a: None if a is None else np.ascontiguousarray(a, np.float32)
    P, W, H = case["means3D"].shape[0], case["W"], case["H"]
    M = 0 if case["shs"] is None else case["shs"].shape[1]
    s = L.Settings(P=P, image_width=W, image_height=H, sh_degree=case["sh_degree"], sh_coeffs=M,
                   use_sa=int(case["use_sa"]), prefiltered=0, debug=0, tanfovx=float(case["tanfovx"]),
                   tanfovy=float(case["tanfovy"]), scale_modifier=float(case["scale_modifier"]), reserved_=0.0)
    keep = {k: f32(case[k]) for k in ("bg", "means3D", "shs", "colors", "opacities", "scales", "rotations", "transMat",
                                      "viewmatrix", "projmatrix", "campos")}
    inp = L.Inputs(background=_ptr(keep["bg"]), means3D=_ptr(keep["means3D"]), shs=_ptr(keep["shs"]),
                   colors_precomp=_ptr(keep["colors"]), opacities=_ptr(keep["opacities"]), scales=_ptr(keep["scales"]),
                   rotations=_ptr(keep["rotations"]), transMat_precomp=_ptr(keep["transMat"]),
                   viewmatrix=_ptr(keep["viewmatrix"]), projmatrix=_ptr(keep["projmatrix"]), campos=_ptr(keep["campos"]))
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    out = dict(radii=np.full(P, -7, np.int32), depths=np.zeros(P, np.float32), splat=np.zeros((P, 20), np.float32),
               tiles_touched=np.zeros(P, np.uint32), point_offsets=np.zeros(P, np.uint32),
               ranges=np.zeros((tiles, 2), np.uint32), tile_count=np.zeros(tiles, np.uint32))
    n = C.c_uint32(0)
    rc = lib.emu_forward_preprocess(C.byref(s), C.byref(inp), *[_ptr(out[k]) for k in
                                    ("radii", "depths", "splat", "tiles_touched", "point_offsets", "ranges", "tile_count")],
                                    C.cast(C.byref(n), C.c_void_p))
    assert rc == L.GS2D_OK
    sp = out["splat"]
    out.update(num_rendered=int(n.value), transMat=np.concatenate([sp[:, 0:3], sp[:, 4:7], sp[:, 8:11]], axis=1),
               means2D=np.stack([sp[:, 3], sp[:, 7]], axis=1), normal=sp[:, 12:15], opacity=sp[:, 11], rgb=sp[:, 16:19])
    return out


def _bits_differ(a, b):
    return int((np.ascontiguousarray(a, np.float32).view(np.uint32) != np.ascontiguousarray(b, np.float32).view(np.uint32)).sum())


@pytest.mark.parametrize("name", cases.GOLDEN_CASES)
def test_preprocess_kernel_text_reproduces_the_compiled_reference(lib, name):
    g = cases.load_golden(name)
    assert g is not None, f"tests/golden/{name}.npz missing"
    case = cases.case_from_golden(g)
    got = _run(lib, case)
    # integer stages: exact
    assert np.array_equal(got["radii"], g["ref_radii"])
    assert np.array_equal(got["tiles_touched"], g["ref_tiles_touched"])
    assert np.array_equal(got["point_offsets"], g["ref_point_offsets"])
    assert got["num_rendered"] == int(g["ref_num_rendered"]) == int(got["tile_count"].sum())
    ranges = g["ref_ranges"].astype(np.int64)
    assert np.array_equal(got["tile_count"].astype(np.int64), ranges[:, 1] - ranges[:, 0])
    touched = got["tile_count"] > 0            # the reference leaves untouched tiles at (0, 0)
    assert np.array_equal(got["ranges"][touched].astype(np.int64), ranges[touched])
    # per-surfel float state of the visible surfels
    vis = g["ref_radii"] > 0
    assert _bits_differ(got["depths"][vis], g["ref_depths"][vis]) == 0
    assert _bits_differ(got["opacity"][vis], g["ref_normal_opacity"][vis, 3]) == 0
    pairs = ((got["transMat"][vis], g["ref_transMat"][vis]), (got["means2D"][vis], g["ref_means2D"][vis]),
             (got["normal"][vis], g["ref_normal_opacity"][vis, :3]))
    for a, b in pairs:
        if case["transMat"] is not None:          # no quaternion, no device rsqrt: every bit
            assert _bits_differ(a, b) == 0
        elif b.size:
            assert float(np.abs(a - b).max()) <= 2e-6 * float(np.abs(b).max())


def test_look_back_scan_over_many_blocks_and_oracle_agreement(lib):
    """20 000 surfels = 10 scan CTAs chained by look-back + the tile-range CTA: offsets are the inclusive scan of the
    tile counts, ranges partition the instances by tile; and the C restatement agrees on the integers up to its
    documented ulp-boundary radii (oracle/gs2d_oracle.c header: <= 0.1 % of the surfels)."""
    import cpu as O
    case = cases.room_sa()
    from gs2d_b200 import scene
    big = scene.room_scene(20_000, cases._intr(case["W"], case["H"]), seed=21)
    case.update({k: big[k] for k in ("means3D", "scales", "rotations", "opacities", "colors")})
    got = _run(lib, case)
    tt = got["tiles_touched"].astype(np.int64)
    assert np.array_equal(np.cumsum(tt), got["point_offsets"].astype(np.int64))
    assert got["num_rendered"] == int(tt.sum()) == int(got["tile_count"].sum())
    cnt = got["tile_count"].astype(np.int64)
    r = got["ranges"].astype(np.int64)
    assert np.array_equal(r[:, 1] - r[:, 0], cnt) and np.array_equal(r[:, 0], np.cumsum(cnt) - cnt)
    assert np.all((got["radii"] > 0) == (tt > 0))
    a = (case["means3D"], case["opacities"], case["viewmatrix"], case["projmatrix"], case["W"], case["H"], case["bg"])
    st = O.forward(*a, scales=case["scales"], rotations=case["rotations"], colors=case["colors"], campos=case["campos"],
                   scale_modifier=float(case["scale_modifier"]), use_sa=case["use_sa"])
    nd = int((np.asarray(st["radii"]) != got["radii"]).sum())
    assert nd <= 20, nd


# ---- backward preprocess (csrc/preprocess_bwd.cu) ----------------------------------------------------------------------
def _backward(lib, case, fwd, rows):
    f32 = lambda a: None if a is None else np.ascontiguousarray(a, np.float32)
    P, W, H = case["means3D"].shape[0], case["W"], case["H"]
    s = L.Settings(P=P, image_width=W, image_height=H, sh_degree=0, sh_coeffs=0, use_sa=int(case["use_sa"]), prefiltered=0,
                   debug=0, tanfovx=float(case["tanfovx"]), tanfovy=float(case["tanfovy"]),
                   scale_modifier=float(case["scale_modifier"]), reserved_=0.0)
    keep = {k: f32(case[k]) for k in ("bg", "means3D", "colors", "opacities", "scales", "rotations", "viewmatrix",
                                      "projmatrix", "campos")}
    inp = L.Inputs(background=_ptr(keep["bg"]), means3D=_ptr(keep["means3D"]), shs=None, colors_precomp=_ptr(keep["colors"]),
                   opacities=_ptr(keep["opacities"]), scales=_ptr(keep["scales"]), rotations=_ptr(keep["rotations"]),
                   transMat_precomp=None, viewmatrix=_ptr(keep["viewmatrix"]), projmatrix=_ptr(keep["projmatrix"]),
                   campos=_ptr(keep["campos"]))
    out = dict(means2D=np.zeros((P, 3), np.float32), colors=np.zeros((P, 3), np.float32), opacity=np.zeros((P, 1), np.float32),
               means3D=np.zeros((P, 3), np.float32), transMat=np.zeros((P, 9), np.float32), scales=np.zeros((P, 2), np.float32),
               rotations=np.zeros((P, 4), np.float32))
    gr = L.Grads(dL_dmeans2D=_ptr(out["means2D"]), dL_dcolors=_ptr(out["colors"]), dL_dopacity=_ptr(out["opacity"]),
                 dL_dmeans3D=_ptr(out["means3D"]), dL_dtransMat=_ptr(out["transMat"]), dL_dsh=None,
                 dL_dscales=_ptr(out["scales"]), dL_drotations=_ptr(out["rotations"]))
    lib.emu_backward_preprocess.restype = C.c_int
    lib.emu_backward_preprocess.argtypes = [C.POINTER(L.Settings), C.POINTER(L.Inputs)] + [C.c_void_p] * 5 + [C.POINTER(L.Grads)]
    rows = np.ascontiguousarray(rows, np.float32)
    rc = lib.emu_backward_preprocess(C.byref(s), C.byref(inp), _ptr(fwd["radii"]), _ptr(fwd["splat"]), _ptr(fwd["tiles_touched"]),
                                     _ptr(fwd["point_offsets"]), _ptr(rows), C.byref(gr))
    assert rc == L.GS2D_OK
    return out


@pytest.mark.parametrize("name", ["room_sa", "room_dist", "scale_mod"])
def test_backward_preprocess_kernel_text_equals_the_transliteration(lib, name):
    """The per-surfel sums of the compositing backward pass (taken from the pure-PyTorch transliteration, which is
    pinned by the golden gradients of the compiled reference: tests/test_torch_cpu.py) are spread over each surfel's
    gradient rows; the kernel text must reduce them and finish the chain rule (CR/backward.cu:466-664) like the
    transliteration of the same lines does: 2e-5 of the largest magnitude per tensor, fp32 on both sides."""
    import torch
    import torch_cpu as T
    g = cases.load_golden(name)
    case = cases.case_from_golden(g)
    fwd = _run(lib, case)
    W, H = case["W"], case["H"]
    a = (case["means3D"], case["opacities"], case["viewmatrix"], case["projmatrix"], W, H, case["bg"])
    st = T.forward(*a, scales=case["scales"], rotations=case["rotations"], colors=case["colors"], campos=case["campos"],
                   scale_modifier=float(case["scale_modifier"]), use_sa=case["use_sa"])
    assert np.array_equal(st["radii"].numpy(), fwd["radii"])                   # same surfels, same rows
    with torch.no_grad():
        acc = T.render_backward(W, H, st, case["bg"], case["dL_dcolor"], case["dL_dothers"], case["use_sa"])
        fx = np.float32(W) / (np.float32(2.0) * np.float32(case["tanfovx"]))
        fy = np.float32(H) / (np.float32(2.0) * np.float32(case["tanfovy"]))
        Wb, Hb = int(np.float32(fx) * np.float32(case["tanfovx"]) * np.float32(2)), int(np.float32(fy) * np.float32(case["tanfovy"]) * np.float32(2))
        want = T.preprocess_backward(st, acc, case["means3D"], case["scales"], case["rotations"], case["viewmatrix"],
                                     case["projmatrix"], Wb, Hb)
    # rows: [0:9] dL/dT, [9:11] dL/dmean2D, [11] dL/dopacity, [12:15] dL/dnormal, [15:18] dL/dcolor (csrc/common.cuh)
    P = case["means3D"].shape[0]
    per = np.zeros((P, 20), np.float32)
    per[:, 0:9] = acc["transMat"].numpy(); per[:, 9:11] = acc["means2D"].numpy()[:, :2]; per[:, 11] = acc["opacity"].numpy()[:, 0]
    per[:, 12:15] = acc["normal3D"].numpy(); per[:, 15:18] = acc["colors"].numpy()
    rows = np.zeros((fwd["num_rendered"], 20), np.float32)
    tt = fwd["tiles_touched"].astype(np.int64)
    first = fwd["point_offsets"].astype(np.int64) - tt
    for i in np.nonzero(tt)[0]:
        if tt[i] == 1:
            rows[first[i]] = per[i]
        else:                                   # two exact halves, first and last row of the surfel
            rows[first[i]] = 0.5 * per[i]
            rows[first[i] + tt[i] - 1] = 0.5 * per[i]
    got = _backward(lib, case, fwd, rows)
    for k in ("means3D", "scales", "rotations", "transMat", "colors", "opacity", "means2D"):
        w = want[k].numpy().reshape(got[k].shape)
        scale = float(np.abs(w).max())
        assert scale > 0 and float(np.abs(got[k] - w).max()) <= 2e-5 * scale, k
