"""CPU tests of the drop-in boundary: libgs2d.so loads, exports every symbol include/gs2d.h declares, its size
queries and argument validation behave; the python packages mirror the reference's names.  No GPU compute."""
import ctypes as C
import os
import re

import pytest

from gs2d_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_functions():
    src = open(os.path.join(ROOT, "include", "gs2d.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gs2d_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(L.LIB_PATH)
    names = _header_functions()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/gs2d.h but not exported"
    # and the ctypes mirror covers the header exactly
    assert sorted(L.PROTOTYPES) == names


def test_abi_version_and_strerror():
    lib = L.load()
    assert lib.gs2d_abi_version() == 4
    assert lib.gs2d_strerror(0).decode().lower().startswith("ok") or lib.gs2d_strerror(0)
    for code in (-1, -2, -3, -4):
        assert lib.gs2d_strerror(code).decode()


def test_size_queries_monotone_and_aligned():
    lib = L.load()
    assert lib.gs2d_geom_bytes(0, 0) >= 0
    g1, g2 = lib.gs2d_geom_bytes(1000, 0), lib.gs2d_geom_bytes(500_000, 0)
    assert 0 < g1 < g2
    assert g2 >= 500_000 * (4 + 80 + 8 + 16 + 4 + 4 + 3)
    i1 = lib.gs2d_image_bytes(1200, 680)
    assert i1 >= 1200 * 680 * (12 + 8 + 4 + 4) + 75 * 43 * 8
    b1, b2 = lib.gs2d_binning_bytes(1000, 1200, 680), lib.gs2d_binning_bytes(1_000_000, 1200, 680)
    assert 0 < b1 < b2
    assert lib.gs2d_backward_scratch_bytes(1000, 5000, 90000, 1200, 680) >= 5000 * 80 + 90000 * 16
    assert lib.gs2d_dist2_scratch_bytes(1000) > 1000 * 16
    assert lib.gs2d_sort_scratch_bytes(1 << 20) > 0


def test_argument_validation_without_gpu():
    """Invalid calls are rejected before anything is launched (so this runs on a CPU-only box)."""
    lib = L.load()
    s = L.Settings(P=10, image_width=64, image_height=64)
    assert lib.gs2d_forward_preprocess(None, None, None, None, None, None) == L.GS2D_ERR_INVALID
    i = L.Inputs()   # all NULL
    b = L.Buffers()
    assert lib.gs2d_forward_preprocess(C.byref(s), C.byref(i), C.byref(b), None, None, None) == L.GS2D_ERR_INVALID
    assert lib.gs2d_forward_render(C.byref(s), C.byref(i), C.byref(b), 0, None, None, None, None) == L.GS2D_ERR_INVALID
    assert lib.gs2d_mark_visible(-1, None, None, None, None, None) == L.GS2D_ERR_INVALID
    assert lib.gs2d_mark_visible(0, None, None, None, None, None) == L.GS2D_OK
    assert lib.gs2d_dist2(-1, None, None, None, None) == L.GS2D_ERR_INVALID
    assert lib.gs2d_dist2(0, None, None, None, None) == L.GS2D_OK
    assert lib.gs2d_dist2(5, None, None, None, None) == L.GS2D_ERR_INVALID
    assert lib.gs2d_sort_pairs(-1, None, None, None, None, 32, None, None) == L.GS2D_ERR_INVALID
    assert lib.gs2d_sort_pairs(10, None, None, None, None, 32, None, None) == L.GS2D_ERR_INVALID
    assert lib.gs2d_sort_pairs(10, None, None, None, None, 0, None, None) == L.GS2D_ERR_INVALID
    with pytest.raises(L.Gs2dError):
        L.check(L.GS2D_ERR_CAPACITY, "x")


def test_python_surface_matches_reference_names():
    import gaus_2dgs_rasterization as G
    fields = ("image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
              "sh_degree", "campos", "use_sa", "prefiltered", "debug")
    assert G.GaussianRasterizationSettings._fields == fields      # RAST/gaus_2dgs_rasterization/__init__.py:163-176
    assert hasattr(G.GaussianRasterizer, "markVisible") and hasattr(G.GaussianRasterizer, "forward")
    for n in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):   # RAST/ext.cpp:15-19
        assert callable(getattr(G._C, n))
    from simple_knn._C import distCUDA2                                                   # KNN/ext.cpp:15-17
    assert callable(distCUDA2)


def test_rasterizer_argument_errors():
    """Same exceptions as the reference for inconsistent arguments (__init__.py:198-202)."""
    import torch
    import gaus_2dgs_rasterization as G
    z = torch.zeros(3)
    st = G.GaussianRasterizationSettings(8, 8, 1.0, 1.0, z, 1.0, torch.eye(4), torch.eye(4), 0, z, True, False, False)
    r = G.GaussianRasterizer(st)
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(m, m, torch.zeros(4, 1), scales=torch.zeros(4, 2), rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="scale/rotation pair or precomputed"):
        r(m, m, torch.zeros(4, 1), colors_precomp=m)
    with pytest.raises(RuntimeError):          # CPU tensors are rejected, there is no CPU path
        r(m, m, torch.zeros(4, 1), colors_precomp=m, scales=torch.zeros(4, 2), rotations=torch.zeros(4, 4))


def test_product_never_touches_the_oracle():
    """The shipped package must not import, link or execute anything under oracle/ -- nor the host emulation of its
    own kernels under tests/emu (a CPU pre-check for the tests, never a fallback)."""
    pkg = os.path.join(ROOT, "gaus-slam_b200")
    bad = re.compile(r"libgs2d_oracle|libgs2d_ref|gs2d_oracle\.c|import\s+refgpu|import\s+cpu\b|from\s+oracle|oracle[/\\]|"
                     r"[\"']oracle[\"']|libgs2d_emu|cuda_emu|build_emu|host_only|tests[/\\]emu")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(d, f), errors="ignore").read()
                m = bad.search(txt)
                assert m is None, f"{os.path.join(d, f)} references the oracle: {m.group(0)!r}"


def test_next_row_entry_points_validate_their_arguments():
    """gs2d_map_* / gs2d_pose_transform_* / loss: invalid calls are rejected before anything is launched."""
    lib = L.load()
    ms, ap = L.MapState(), L.AdamParams()
    assert lib.gs2d_map_activate(4, 3, C.byref(ms), None, None, None, None) == L.GS2D_ERR_INVALID      # scale_cols
    assert lib.gs2d_map_activate(0, 2, C.byref(ms), None, None, None, None) == L.GS2D_OK               # nothing to do
    assert lib.gs2d_map_activate(4, 2, C.byref(ms), None, None, None, None) == L.GS2D_ERR_INVALID      # NULL tensors
    assert lib.gs2d_map_adam_step(4, 2, C.byref(ms), None, C.byref(ap), None, None, None, None) == L.GS2D_ERR_INVALID
    assert lib.gs2d_map_adam_step(-1, 2, None, None, None, None, None, None, None) == L.GS2D_ERR_INVALID
    assert lib.gs2d_pose_transform_scratch_bytes() >= 12 * 4 * 148
    # exactly one pose parametrisation: none given / both given
    one = C.c_void_p(16)          # any non-NULL value: validation happens before a launch
    assert lib.gs2d_pose_transform_forward(4, one, None, None, None, None, one, None, None) == L.GS2D_ERR_INVALID
    assert lib.gs2d_pose_transform_forward(4, one, None, one, one, one, one, None, None) == L.GS2D_ERR_INVALID
    assert lib.gs2d_pose_transform_forward(4, one, one, one, one, None, one, None, None) == L.GS2D_ERR_INVALID   # rotations w/o output
    assert lib.gs2d_pose_transform_forward(0, None, None, None, None, None, None, None, None) == L.GS2D_OK
    assert lib.gs2d_pose_transform_backward(4, one, one, one, None, one, None, one, one, None, None, None) == L.GS2D_ERR_INVALID  # no scratch
    assert lib.gs2d_pose_transform_backward(4, one, one, one, None, one, one, None, one, None, None, None) == L.GS2D_ERR_INVALID  # no dL_dcam_rot
    assert lib.gs2d_mapping_loss_forward(0, 8, None, None, None, None, 0, None, None, None, None) == L.GS2D_ERR_INVALID
    # map compaction: scratch = one total + one offset per 256-surfel block
    assert lib.gs2d_map_compact_scratch_bytes(0) == 4 and lib.gs2d_map_compact_scratch_bytes(257) == 4 * 3
    assert lib.gs2d_map_compact_scratch_bytes(-5) == 0
    assert lib.gs2d_map_compact_plan(-1, one, one, None) == L.GS2D_ERR_INVALID
    assert lib.gs2d_map_compact_plan(4, one, None, None) == L.GS2D_ERR_INVALID                             # no scratch
    assert lib.gs2d_map_compact_plan(4, None, one, None) == L.GS2D_ERR_INVALID                             # no mask
    dst = L.MapState()
    assert lib.gs2d_map_compact_move(0, 2, None, C.byref(ms), C.byref(dst), one, None) == L.GS2D_OK        # nothing to do
    assert lib.gs2d_map_compact_move(4, 3, one, C.byref(ms), C.byref(dst), one, None) == L.GS2D_ERR_INVALID
    assert lib.gs2d_map_compact_move(4, 2, one, C.byref(ms), C.byref(dst), one, None) == L.GS2D_ERR_INVALID  # NULL parameters
    for st in (ms, dst):
        for k in ("xyz", "opacity", "scaling", "rotation", "rgb"):
            setattr(st, k, 16 if st is ms else 32)
    dst.exp_avg[2] = 32                                                                                    # moment on one side only
    assert lib.gs2d_map_compact_move(4, 2, one, C.byref(ms), C.byref(dst), one, None) == L.GS2D_ERR_INVALID
    dst.exp_avg[2] = 0
    dst.rgb = ms.rgb                                                                                       # in place
    assert lib.gs2d_map_compact_move(4, 2, one, C.byref(ms), C.byref(dst), one, None) == L.GS2D_ERR_INVALID


def test_next_row_wrappers_refuse_cpu_tensors_and_unbuilt_branches():
    """No CPU path anywhere in the product: the python wrappers raise before touching CUDA."""
    import torch
    from gs2d_b200 import loss, optim, tracking
    P = 8
    raw = [torch.zeros((P, 3)), torch.zeros((P, 1)), torch.zeros((P, 2)), torch.ones((P, 4)), torch.zeros((P, 3))]
    args = dict(xyz_lr=1e-4, opacity_lr=5e-2, scaling_lr=1e-3, rotation_lr=1e-3, rgb_lr=2.5e-3)
    with pytest.raises(RuntimeError, match="CUDA"):
        optim.MapOptimizer(*raw, args)
    with pytest.raises(RuntimeError, match="CUDA"):
        tracking.pose_transform(raw[0], raw[3], cam_rot=torch.tensor([1.0, 0, 0, 0]), cam_trans=torch.zeros(3))
    with pytest.raises(Exception, match="exactly one"):
        tracking.pose_transform(raw[0], raw[3])
    img, am = torch.zeros((3, 8, 8)), torch.zeros((7, 8, 8))
    with pytest.raises(NotImplementedError):
        loss.tracking_loss(img, am, img, torch.zeros((8, 8)), use_normal_loss=True)
    with pytest.raises(RuntimeError, match="CUDA"):      # the outlier mask is built since round 2; still no CPU path
        loss.tracking_loss(img, am, img, torch.zeros((8, 8)), ignore_outliners=True)
    with pytest.raises(RuntimeError, match="CUDA"):      # edge growth is built since round 2; still no CPU path
        loss.mapping_loss(img, am, img, torch.zeros((8, 8)), use_edge_growth=True)
    with pytest.raises(NotImplementedError):
        loss.mapping_loss(img, am, img, torch.zeros((8, 8)), use_normal_loss=True)
    with pytest.raises(RuntimeError):
        loss.mapping_loss(img, am, img, torch.zeros((8, 8)))


def test_header_is_plain_c_and_a_c_client_links(tmp_path):
    """The boundary is a C ABI: include/gs2d.h compiles as strict C99 (no C++ / torch / CUDA types in a signature),
    a C program that references EVERY declared entry point links against libgs2d.so, and the calls that need no
    GPU answer (version, sizes, argument validation)."""
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("needs gcc")
    names = _header_functions()
    src = os.path.join(tmp_path, "client.c")
    with open(src, "w") as f:
        f.write('#include <stdio.h>\n#include "gs2d.h"\n')
        f.write("typedef void (*fn)(void);\nstatic const fn table[] = {\n")
        f.write("".join(f"\t(fn){n},\n" for n in names))
        f.write("};\nint main(void) {\n"
                "\tif (sizeof(table) / sizeof(table[0]) != %d) return 2;\n"
                "\tif (gs2d_abi_version() != GS2D_ABI_VERSION) return 3;\n"
                "\tif (gs2d_map_compact_scratch_bytes(257) != 12) return 4;\n"
                "\tif (gs2d_map_activate(4, 3, 0, 0, 0, 0, 0) == GS2D_OK) return 5;\n"
                '\tprintf("%%s\\n", gs2d_strerror(GS2D_ERR_INVALID));\n\treturn 0;\n}\n' % len(names))
    exe = os.path.join(tmp_path, "client")
    libdir = os.path.dirname(L.LIB_PATH)
    r = subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror",
                        "-I", os.path.join(ROOT, "include"), src, "-o", exe, "-L", libdir, "-lgs2d",
                        "-Wl,-rpath," + libdir], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip(), (r.returncode, r.stdout, r.stderr)


def test_every_kernel_id_has_a_name():
    """bench.py indexes its per-kernel table by these names: the name table must cover the whole id range."""
    lib = L.load()
    n = lib.gs2d_num_kernels()
    names = [lib.gs2d_kernel_name(i) for i in range(n)]
    assert n >= 20 and all(names) and len(set(names)) == n
    assert lib.gs2d_kernel_name(n) == b"?" and lib.gs2d_kernel_name(-1) == b"?"
