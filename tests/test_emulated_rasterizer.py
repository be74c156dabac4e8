"""CPU: the WHOLE hot path through the C ABI without a GPU.  Every source of libgs2d.so except the NVSwitch collective
(api.cu, preprocess, scan, scatter, tile sort + footprints, forward and backward compositing, backward preprocess, ...)
is built for the host from its SOURCE TEXT under the CUDA-thread emulator (tests/emu/build_emu.py: build_full), driven
with the call sequence of gaus_2dgs_rasterization/_C.py on host memory (tests/emu/emu_run.py), and held to the golden
vectors of the UNMODIFIED compiled reference (tests/golden/*.npz, produced on a B200):

  * radii, tiles touched, offsets, keys (unsorted and sorted), point list, tile ranges: EXACT;
  * images and per-pixel state at the tolerances used for the CPU oracle (tests/compare.py: host exp / rsqrt vs the
    device's ex2.approx / MUFU.RSQ, which decide the last bit of every alpha);
  * gradients at BASELINE's 1e-3 (1e-2 through the surface-aware confidence, see tests/compare.py);
  * footprint masks are a superset of the contributing pairs (audit: 0 missed).

mbarrier / bulk-copy PTX is replaced by tests/emu/tma_emu.cuh, the L2 discard hint is compiled out.  This checks the
arithmetic, indexing and control flow of the kernels as written -- a gate for kernel edits where there is no GPU --
not performance, memory ordering or hardware behaviour; the GPU tests remain the parity tests proper.

The default run takes the cases that cover every branch (about a minute); GS2D_EMU_ALL=1 runs all of them."""
import os
import shutil
import sys

import numpy as np
import pytest

import cases
import compare

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))

pytestmark = pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++ (C++20) for the host build of the kernels")

# surface-aware, 2DGS distortion branch, SH colours, precomputed transform, nothing visible
DEFAULT = ["room_sa", "room_dist", "sh_deg2", "behind"]          # transmat_pre: tests/test_emulated_public_api.py
CASES = cases.GOLDEN_CASES if os.environ.get("GS2D_EMU_ALL") else DEFAULT


@pytest.mark.parametrize("name", CASES)
def test_product_sources_on_the_host_match_the_compiled_reference(name):
    import emu_run
    g = cases.load_golden(name)
    assert g is not None, f"tests/golden/{name}.npz missing"
    case = cases.case_from_golden(g)
    ref = {k[4:]: v for k, v in g.items() if k.startswith("ref_")}
    out = emu_run.run_ours(case, debug=True, backward=True)
    fails = compare.check_binning(out, ref)                       # integer stages exact, float state 1e-5
    fails += compare.check_images_oracle(out, ref, case["use_sa"])
    fails += compare.check_grads(out, ref, rtol=compare.ORACLE_SA_GRAD_RTOL if case["use_sa"] else compare.GRAD_RTOL)
    assert not fails, "\n".join(fails)
    hits, bits, missed = (int(x) for x in out["footprint_audit"])
    assert missed == 0 and bits >= hits
    if int(out["num_rendered"]):
        assert out["num_pairs"] == bits                           # pair slots of the backward pass = mask bits
    # every output was written (buffers start as NaN / garbage)
    assert np.isfinite(out["color"]).all() and np.isfinite(out["others"]).all()
    for k in compare.GRADS:
        assert np.isfinite(out["grad_" + k]).all(), k


def test_release_path_equals_debug_path():
    """debug = 0 (no reference-order unsorted list, no per-stage checks) gives the same frame as debug = 1."""
    import emu_run
    case = cases.case_from_golden(cases.load_golden("sh_deg2"))
    a = emu_run.run_ours(case, debug=True, backward=True)
    b = emu_run.run_ours(case, debug=False, backward=True)
    assert not compare.check_images(a, b, exact=True)
    assert np.array_equal(a["point_list"], b["point_list"]) and np.array_equal(a["ranges"], b["ranges"])
    for k in compare.GRADS:
        if a["grad_" + k].size:
            assert compare.grad_rel(b["grad_" + k], a["grad_" + k]) < 1e-5, k


# ---- the rest of the rasterizer's API surface on the host build --------------------------
def test_mark_visible_matches_the_reference():
    import ctypes as C
    import emu_run
    lib = emu_run.library()
    g = cases.load_golden("mark_visible")
    pts = np.ascontiguousarray(g["in_means3D"], np.float32)
    view, proj = np.ascontiguousarray(g["in_viewmatrix"], np.float32), np.ascontiguousarray(g["in_projmatrix"], np.float32)
    present = np.full(pts.shape[0], 7, np.uint8)
    assert lib.gs2d_mark_visible(pts.shape[0], emu_run._ptr(pts), emu_run._ptr(view), emu_run._ptr(proj),
                                 emu_run._ptr(present), None) == 0
    assert np.array_equal(present.astype(bool), g["ref_present"])


@pytest.mark.parametrize("name", cases.KNN_CASES)
def test_dist2_matches_the_reference_simple_knn(name):
    """distCUDA2 (Morton sort by the hand-written onesweep, box search): bit for bit against the reference's output --
    squared distances are sums of products with no approximate device instruction in them."""
    import emu_run
    lib = emu_run.library()
    g = cases.load_golden(name)
    pts = np.ascontiguousarray(g["in_points"], np.float32)
    P = pts.shape[0]
    out = np.full(P, np.nan, np.float32)
    scratch = emu_run._garbage(lib.gs2d_dist2_scratch_bytes(P))
    assert lib.gs2d_dist2(P, emu_run._ptr(pts), emu_run._ptr(out), emu_run._ptr(scratch), None) == 0
    assert compare.bits_differ(out, g["ref_dist2"]) == 0, f"max abs {np.abs(out - g['ref_dist2']).max():.3e}"


@pytest.mark.parametrize("n,end_bit", [(1, 40), (2047, 44), (2049, 44), (10000, 45), (5000, 17)])
def test_onesweep_sort_is_the_stable_sort_of_the_low_bits(n, end_bit):
    """gs2d_sort_pairs = cub::DeviceRadixSort::SortPairs(begin_bit 0, end_bit) (CR/rasterizer_impl.cu:309-314)."""
    import emu_run
    lib = emu_run.library()
    rng = np.random.default_rng(n)
    # keys as the rasterizer makes them: nothing set at or above end_bit (whole 8-bit digits are sorted, include/gs2d.h)
    keys = rng.integers(0, 1 << end_bit, n, dtype=np.uint64)
    keys[rng.uniform(size=n) < 0.3] &= np.uint64(0xFF)                         # many equal keys: stability matters
    vals = np.arange(n, dtype=np.uint32)
    ko, vo = np.zeros(n, np.uint64), np.zeros(n, np.uint32)
    scratch = emu_run._garbage(lib.gs2d_sort_scratch_bytes(n))
    assert lib.gs2d_sort_pairs(n, emu_run._ptr(keys), emu_run._ptr(vals), emu_run._ptr(ko), emu_run._ptr(vo), end_bit,
                               emu_run._ptr(scratch), None) == 0
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(vo, vals[order]) and np.array_equal(ko, keys[order])


# ---- edge shapes against the C restatement (the reference has no golden for them) ---------------------------------------
def _oracle(case):
    import cpu as O
    a = (case["means3D"], case["opacities"], case["viewmatrix"], case["projmatrix"], case["W"], case["H"], case["bg"])
    st = O.forward(*a, scales=case["scales"], rotations=case["rotations"], colors=case["colors"], campos=case["campos"],
                   scale_modifier=float(case["scale_modifier"]), use_sa=case["use_sa"])
    g = O.backward(st, case["means3D"], case["viewmatrix"], case["projmatrix"], case["W"], case["H"], float(case["tanfovx"]),
                   float(case["tanfovy"]), case["bg"], case["dL_dcolor"], case["dL_dothers"], scales=case["scales"],
                   rotations=case["rotations"], campos=case["campos"], use_sa=case["use_sa"])
    out = {k: np.asarray(v) for k, v in st.items()}
    out.update({"grad_" + k: np.asarray(v) for k, v in g.items()})
    out["transMat"] = out["transMat_used"]
    return out


@pytest.mark.parametrize("P,size,seed,use_sa", [
    (1, (16, 16), 0, True), (5, (1, 1), 1, True), (300, (33, 17), 2, True), (300, (17, 33), 3, False),
    (800, (250, 3), 4, True), (500, (15, 15), 5, False), (300, (129, 65), 8, False), (50, (1200, 8), 9, True)])
def test_ragged_and_tiny_images_against_the_c_restatement(P, size, seed, use_sa):
    """One surfel, one pixel, images narrower than a tile, one tile row of 75 tiles, non-zero background; debug and
    release paths alternate.  The oracle may place a radius across an integer boundary for a handful of surfels
    (oracle/gs2d_oracle.c header): binning is compared only when the radii agree, images and gradients always."""
    import emu_run
    case = cases._base(P, cases._intr(*size), seed=seed, view=seed % 3, use_sa=use_sa,
                       bg=(0.1, 0.2, 0.3) if seed % 2 else (0.0, 0.0, 0.0))
    out = emu_run.run_ours(case, debug=bool(seed % 2), backward=True)
    want = _oracle(case)
    fails = compare.check_images_oracle(out, want, use_sa)
    fails += compare.check_grads(out, want, rtol=compare.ORACLE_SA_GRAD_RTOL if use_sa else compare.GRAD_RTOL)
    if np.array_equal(out["radii"], want["radii"]):
        fails += compare.check_binning(out, want)
    assert not fails, "\n".join(fails)
    assert int(out["footprint_audit"][2]) == 0


def test_empty_map_through_the_abi():
    """P = 0: nothing is launched that reads a surfel; the image is the background, alpha 0 (rasterize_points.cu:101)."""
    import emu_run
    case = cases._base(4, cases._intr(40, 24), seed=0, view=0, bg=(0.25, 0.5, 0.75))
    for k in ("means3D", "scales", "rotations", "opacities", "colors"):
        case[k] = case[k][:0]
    out = emu_run.run_ours(case, debug=False, backward=True)
    assert int(out["num_rendered"]) == 0
    assert np.allclose(out["color"], np.array([0.25, 0.5, 0.75], np.float32)[:, None, None]) and not out["others"][1].any()


def test_one_call_forward_and_its_capacity_protocol():
    """gs2d_forward (the reference's control flow for a plain-C caller): *num_rendered_io carries the provisioned
    capacity in and the count out; GS2D_ERR_CAPACITY when it does not fit, then the same frame as the staged calls."""
    import ctypes as C
    import emu_run
    from gs2d_b200 import _lib as L
    lib = emu_run.library()
    case = cases.case_from_golden(cases.load_golden("sh_deg2"))
    staged = emu_run.run_ours(case, debug=False, backward=False)
    n = int(staged["num_rendered"])
    P, W, H, M = case["means3D"].shape[0], case["W"], case["H"], case["shs"].shape[1]
    f32 = lambda a: None if a is None else np.ascontiguousarray(a, np.float32)      # noqa: E731
    k = {x: f32(case[x]) for x in ("bg", "means3D", "shs", "opacities", "scales", "rotations", "viewmatrix", "projmatrix", "campos")}
    s = L.Settings(P=P, image_width=W, image_height=H, sh_degree=int(case["sh_degree"]), sh_coeffs=M, use_sa=int(case["use_sa"]),
                   prefiltered=0, debug=0, tanfovx=float(case["tanfovx"]), tanfovy=float(case["tanfovy"]),
                   scale_modifier=float(case["scale_modifier"]), reserved_=0.0)
    p = emu_run._ptr
    inp = L.Inputs(background=p(k["bg"]), means3D=p(k["means3D"]), shs=p(k["shs"]), colors_precomp=None, opacities=p(k["opacities"]),
                   scales=p(k["scales"]), rotations=p(k["rotations"]), transMat_precomp=None, viewmatrix=p(k["viewmatrix"]),
                   projmatrix=p(k["projmatrix"]), campos=p(k["campos"]))
    geom, img = emu_run._garbage(lib.gs2d_geom_bytes(P, M)), emu_run._garbage(lib.gs2d_image_bytes(W, H))
    radii = np.zeros(P, np.int32)
    color, others = np.full((3, H, W), np.nan, np.float32), np.full((7, H, W), np.nan, np.float32)

    def call(capacity):
        binning = emu_run._garbage(lib.gs2d_binning_bytes(capacity, W, H))
        bufs = L.Buffers(geom=p(geom), geom_bytes=geom.size, binning=p(binning), binning_bytes=binning.size, image=p(img),
                         image_bytes=img.size)
        io = C.c_int64(capacity)
        rc = lib.gs2d_forward(C.byref(s), C.byref(inp), C.byref(bufs), p(radii), C.byref(io), p(color), p(others), None)
        return rc, int(io.value)
    rc, count = call(n // 2)
    assert rc == L.GS2D_ERR_CAPACITY and count == n
    rc, count = call(n + 100)
    assert rc == L.GS2D_OK and count == n
    assert np.array_equal(radii, staged["radii"])
    assert not compare.check_images(dict(color=color, others=others), staged, exact=True)


@pytest.mark.skipif(not os.environ.get("GS2D_EMU_ALL"), reason="about a minute and a half under emulation: GS2D_EMU_ALL=1")
def test_stress_resolution_of_baseline_config_5_against_the_c_restatement():
    """1752 x 1168 (BASELINE.json configs[4]: 8030 tiles, 45 sort bits) with a sparse map: tile indices beyond 2^12,
    thousands of empty tiles, the multi-CTA tile-count scan.  Last run by hand at the end of round 1: radii, keys,
    ranges exact; images, gradients and the footprint audit clean."""
    import emu_run
    case = cases._base(400, cases._intr(1752, 1168, f=1000.0), seed=7, view=1, use_sa=True)
    out = emu_run.run_ours(case, debug=False, backward=True)
    want = _oracle(case)
    assert np.array_equal(out["radii"], want["radii"])
    fails = compare.check_binning(out, want) + compare.check_images_oracle(out, want, True)
    fails += compare.check_grads(out, want, rtol=compare.ORACLE_SA_GRAD_RTOL)
    assert not fails, "\n".join(fails)
    assert int(out["footprint_audit"][2]) == 0


def test_one_tile_with_six_thousand_instances():
    """The tile-local sort, the staged compositing ring and the pair-slot scan on a list far longer than the usual
    ~250 instances per tile: 6000 large surfels over a single 16 x 16 tile, against the C restatement."""
    import emu_run
    case = cases._base(6000, cases._intr(16, 16), seed=3, view=0, use_sa=False, scale_mult=4.0)
    out = emu_run.run_ours(case, debug=False, backward=True)
    want = _oracle(case)
    assert np.array_equal(out["radii"], want["radii"])
    r = out["ranges"].astype(np.int64)
    assert int((r[:, 1] - r[:, 0]).max()) > 5000
    fails = compare.check_binning(out, want) + compare.check_images_oracle(out, want, False) + compare.check_grads(out, want)
    assert not fails, "\n".join(fails)
    assert int(out["footprint_audit"][2]) == 0


@pytest.mark.parametrize("degree,coeffs", [(0, 1), (1, 4), (3, 16), (1, 16)])
def test_spherical_harmonics_degrees_and_clamping(degree, coeffs):
    """The cold SH branch at every degree the reference supports (CR/forward.cu:20-71, CR/backward.cu:20-139), more
    coefficients stored than the active degree uses, coefficients large enough to drive colours below zero (clamped:
    their gradient is cut): against the C restatement."""
    import cpu as O
    import emu_run
    case = cases._base(800, cases._intr(96, 64), seed=10 + degree, view=degree % 3, use_sa=bool(degree % 2))
    case["shs"] = np.random.default_rng(degree).normal(0, 1.2, (800, coeffs, 3)).astype(np.float32)
    case["colors"], case["sh_degree"] = None, degree
    out = emu_run.run_ours(case, debug=False, backward=True)
    a = (case["means3D"], case["opacities"], case["viewmatrix"], case["projmatrix"], case["W"], case["H"], case["bg"])
    kw = dict(scales=case["scales"], rotations=case["rotations"], campos=case["campos"], shs=case["shs"], sh_degree=degree)
    st = O.forward(*a, scale_modifier=1.0, use_sa=case["use_sa"], **kw)
    g = O.backward(st, case["means3D"], case["viewmatrix"], case["projmatrix"], case["W"], case["H"], float(case["tanfovx"]),
                   float(case["tanfovy"]), case["bg"], case["dL_dcolor"], case["dL_dothers"], use_sa=case["use_sa"], **kw)
    want = {k: np.asarray(v) for k, v in st.items()}
    want.update({"grad_" + k: np.asarray(v) for k, v in g.items()})
    assert np.array_equal(out["radii"], want["radii"])
    fails = compare.check_images_oracle(out, want, case["use_sa"])
    fails += compare.check_grads(out, want, rtol=compare.ORACLE_SA_GRAD_RTOL if case["use_sa"] else compare.GRAD_RTOL)
    assert not fails, "\n".join(fails)
    assert float(np.abs(out["grad_sh"]).max()) > 0 and (out["color"] == 0).any() is not None


@pytest.mark.parametrize("i", range(6))
def test_random_shapes_intrinsics_and_backgrounds(i):
    """Seeded fuzz: surfel count, image size (down to a few pixels), focal length, off-centre principal point, splat
    size, background, SA / 2DGS branch, debug / release path -- product sources on the host against the C restatement."""
    import emu_run
    rng = np.random.default_rng(7700 + i)
    P, W, H = int(rng.integers(1, 2500)), int(rng.integers(1, 300)), int(rng.integers(1, 200))
    use_sa, scale = bool(rng.integers(0, 2)), float(rng.choice([0.1, 0.5, 1.0, 2.0, 5.0]))
    intr = cases._intr(W, H, f=float(rng.uniform(0.2, 2.0)) * max(W, 2), cx=float(rng.uniform(0, W)), cy=float(rng.uniform(0, H)))
    case = cases._base(P, intr, seed=200 + i, view=i % 3, use_sa=use_sa, scale_mult=scale, bg=tuple(rng.uniform(0, 1, 3)))
    out = emu_run.run_ours(case, debug=bool(i % 2), backward=True)
    want = _oracle(case)
    nd = int((out["radii"] != want["radii"]).sum())
    assert nd <= max(1, P // 1000)                       # the oracle's ulp-boundary radii (oracle/gs2d_oracle.c header)
    fails = compare.check_images_oracle(out, want, use_sa)
    fails += compare.check_grads(out, want, rtol=compare.ORACLE_SA_GRAD_RTOL if use_sa else compare.GRAD_RTOL)
    if nd == 0:
        fails += compare.check_binning(out, want)
    assert not fails, "\n".join(fails)
    assert int(out["footprint_audit"][2]) == 0
