"""CPU: the drop-in package itself -- `GaussianRasterizationSettings`, `GaussianRasterizer`, the autograd bridge and the
ctypes glue of gaus_2dgs_rasterization, loaded host-only (tests/emu/host_only.py: load_rasterizer) over the host build
of the whole library -- called the way render/render_2dgs.py calls the reference, and held to the golden vectors of
the unmodified compiled reference: images through `forward`, every gradient through `torch.autograd.backward`."""
import os
import shutil
import sys

import numpy as np
import pytest
import torch

import cases
import compare

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))

pytestmark = pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++ (C++20) for the host build of the kernels")


@pytest.fixture(scope="module")
def G():
    import host_only
    return host_only.load_rasterizer()


def _t(a, rg=False):
    return None if a is None else torch.tensor(np.ascontiguousarray(a, np.float32), requires_grad=rg)


def _settings(G, case, debug=False):
    return G.GaussianRasterizationSettings(
        image_height=case["H"], image_width=case["W"], tanfovx=float(case["tanfovx"]), tanfovy=float(case["tanfovy"]),
        bg=_t(case["bg"]), scale_modifier=float(case["scale_modifier"]), viewmatrix=_t(case["viewmatrix"]).view(4, 4),
        projmatrix=_t(case["projmatrix"]).view(4, 4), sh_degree=case["sh_degree"], campos=_t(case["campos"]),
        use_sa=case["use_sa"], prefiltered=False, debug=debug)


@pytest.mark.parametrize("name", cases.GOLDEN_CASES if os.environ.get("GS2D_EMU_ALL") else ["sh_deg2", "transmat_pre"])
def test_public_api_forward_and_autograd_match_the_compiled_reference(G, name):
    g = cases.load_golden(name)
    case = cases.case_from_golden(g)
    P = case["means3D"].shape[0]
    leaves = dict(means3D=_t(case["means3D"], True), means2D=torch.zeros((P, 3), requires_grad=True),
                  opacity=_t(case["opacities"], True))
    kw = {}
    if case["shs"] is not None:
        leaves["sh"] = _t(case["shs"], True); kw["shs"] = leaves["sh"]
    else:
        leaves["colors"] = _t(case["colors"], True); kw["colors_precomp"] = leaves["colors"]
    if case["transMat"] is not None:
        leaves["transMat"] = _t(case["transMat"], True); kw["cov3D_precomp"] = leaves["transMat"]
    else:
        leaves["scales"] = _t(case["scales"], True); leaves["rotations"] = _t(case["rotations"], True)
        kw.update(scales=leaves["scales"], rotations=leaves["rotations"])
    color, radii, allmap = G.GaussianRasterizer(_settings(G, case))(means3D=leaves["means3D"], means2D=leaves["means2D"],
                                                                    opacities=leaves["opacity"], **kw)
    ref = {k[4:]: v for k, v in g.items() if k.startswith("ref_")}
    assert np.array_equal(radii.numpy(), ref["radii"])
    got = dict(color=color.detach().numpy(), others=allmap.detach().numpy())
    assert not compare.check_images_oracle(got, ref, case["use_sa"])
    torch.autograd.backward([color, allmap], [_t(case["dL_dcolor"]), _t(case["dL_dothers"])])
    rtol = compare.ORACLE_SA_GRAD_RTOL if case["use_sa"] else compare.GRAD_RTOL
    for k, leaf in leaves.items():
        want = ref["grad_" + k]
        if want.size:
            assert compare.grad_rel(leaf.grad.numpy().reshape(want.shape), want) <= rtol, k


def test_empty_map_through_the_public_api(G):
    case = cases._base(4, cases._intr(40, 24), seed=0, view=0, bg=(0.25, 0.5, 0.75))
    z = lambda *s: torch.zeros(s, requires_grad=True)      # noqa: E731
    m3, m2, op, col, sc, rot = z(0, 3), z(0, 3), z(0, 1), z(0, 3), z(0, 2), z(0, 4)
    color, radii, allmap = G.GaussianRasterizer(_settings(G, case))(means3D=m3, means2D=m2, opacities=op, colors_precomp=col,
                                                                    scales=sc, rotations=rot)
    assert color.shape == (3, 24, 40) and allmap.shape == (7, 24, 40) and radii.shape == (0,)
    assert torch.allclose(color, torch.tensor([0.25, 0.5, 0.75])[:, None, None].expand(3, 24, 40))
    assert float(allmap[1].detach().abs().max()) == 0.0
    (color.sum() + allmap.sum()).backward()
    assert m3.grad is None or m3.grad.shape == (0, 3)


def test_argument_errors_are_the_reference_ones(G):
    """render/render_2dgs.py relies on these (RAST/gaus_2dgs_rasterization/__init__.py:199-206, rasterize_points.cu:62)."""
    case = cases.room_sa()
    rast = G.GaussianRasterizer(_settings(G, case))
    P = case["means3D"].shape[0]
    m3, m2, op = _t(case["means3D"]), torch.zeros((P, 3)), _t(case["opacities"])
    col, sc, rot = _t(case["colors"]), _t(case["scales"]), _t(case["rotations"])
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        rast(means3D=m3, means2D=m2, opacities=op, scales=sc, rotations=rot)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed"):
        rast(means3D=m3, means2D=m2, opacities=op, colors_precomp=col)
    with pytest.raises(RuntimeError, match="num_points, 3"):
        rast(means3D=m3[:, :2], means2D=m2, opacities=op, colors_precomp=col, scales=sc, rotations=rot)


def test_early_scatter_arena_paths_give_the_same_frame(G):
    """Release mode overlaps the host's wait for num_rendered with the scatter into an arena sized from earlier frames
    (gs2d_forward_scatter_early / _render_scattered).  Three paths -- no arena yet, arena too small (a bigger scene
    arrives), arena large enough -- must produce the same frame bit for bit."""
    def frame(case):
        P = case["means3D"].shape[0]
        color, radii, allmap = G.GaussianRasterizer(_settings(G, case))(
            means3D=_t(case["means3D"]), means2D=torch.zeros((P, 3)), opacities=_t(case["opacities"]),
            colors_precomp=_t(case["colors"]), scales=_t(case["scales"]), rotations=_t(case["rotations"]))
        return color.numpy().copy(), allmap.numpy().copy(), radii.numpy().copy()
    G._C._arenas.clear()
    small = cases._base(300, cases._intr(64, 48), seed=3, view=0)
    big = cases._base(1500, cases._intr(64, 48), seed=4, view=1, scale_mult=2.0)
    frame(small)                                   # no arena yet -> one is sized from this frame
    assert len(G._C._arenas) == 1
    key = next(iter(G._C._arenas))
    G._C._arenas[key] = (G._C._arenas[key][0][:64], 64)     # the smallest real arena holds 2^18 words: shrink it by hand
    a = frame(big)                                 # arena too small: scatter after the wait, arena grows
    assert G._C._arenas[key][1] > 64
    b = frame(big)                                 # early scatter into the arena
    for x, y in zip(a, b):
        assert np.array_equal(x.view(np.uint32) if x.dtype == np.float32 else x, y.view(np.uint32) if y.dtype == np.float32 else y)


def test_debug_mode_dumps_the_arguments_on_failure_like_the_reference(G, tmp_path, monkeypatch):
    """debug=True: a failing forward leaves `snapshot_fw.dump` (the CPU copy of its arguments) in the working
    directory and re-raises (RAST/gaus_2dgs_rasterization/__init__.py:83-93)."""
    monkeypatch.chdir(tmp_path)
    case = cases.room_sa()
    P = case["means3D"].shape[0]
    rast = G.GaussianRasterizer(_settings(G, case, debug=True))
    with pytest.raises(RuntimeError, match="num_points, 3"):
        rast(means3D=_t(case["means3D"])[:, :2], means2D=torch.zeros((P, 3)), opacities=_t(case["opacities"]),
             colors_precomp=_t(case["colors"]), scales=_t(case["scales"]), rotations=_t(case["rotations"]))
    dump = torch.load(os.path.join(tmp_path, "snapshot_fw.dump"), weights_only=False)
    assert isinstance(dump, tuple) and len(dump) == 20 and dump[1].shape == (P, 2) and dump[13] == case["W"]
    # and a healthy debug frame (per-stage checks on) equals the release frame
    a = rast(means3D=_t(case["means3D"]), means2D=torch.zeros((P, 3)), opacities=_t(case["opacities"]),
             colors_precomp=_t(case["colors"]), scales=_t(case["scales"]), rotations=_t(case["rotations"]))
    b = G.GaussianRasterizer(_settings(G, case))(means3D=_t(case["means3D"]), means2D=torch.zeros((P, 3)),
                                                 opacities=_t(case["opacities"]), colors_precomp=_t(case["colors"]),
                                                 scales=_t(case["scales"]), rotations=_t(case["rotations"]))
    assert torch.equal(a[0], b[0]) and torch.equal(a[2], b[2]) and torch.equal(a[1], b[1])


@pytest.mark.parametrize("name", cases.KNN_CASES)
def test_dist_cuda2_through_the_drop_in_module(name):
    """`from simple_knn._C import distCUDA2` as scene/Gaussians.py:8 uses it: bit for bit the reference's output."""
    import host_only
    knn = host_only.load_simple_knn()
    g = cases.load_golden(name)
    got = knn.distCUDA2(torch.tensor(g["in_points"])).numpy()
    assert got.dtype == np.float32 and compare.bits_differ(got, g["ref_dist2"]) == 0
    assert knn.distCUDA2(torch.zeros((0, 3))).shape == (0,)
    with pytest.raises(RuntimeError, match="num_points, 3"):
        knn.distCUDA2(torch.zeros((5, 2)))
