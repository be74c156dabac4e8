"""CPU: the oracle (oracle/gs2d_oracle.c, the C restatement) against the golden fixtures that the UNMODIFIED
reference CUDA rasterizer wrote on a B200 (tests/golden/*.npz, made by tests/golden/make_golden.py).
This is what pins the oracle (the reference itself has no tests or golden vectors)."""
import numpy as np
import pytest

import cases
import compare
import cpu as O


def _ref(g):
    return {k[4:]: v for k, v in g.items() if k.startswith("ref_")}


def _run_oracle(case):
    st = O.forward(case["means3D"], case["opacities"], case["viewmatrix"], case["projmatrix"], case["W"], case["H"],
                   case["bg"], scales=case["scales"], rotations=case["rotations"], transMat=case["transMat"],
                   colors=case["colors"], shs=case["shs"], sh_degree=case["sh_degree"], campos=case["campos"],
                   scale_modifier=float(case["scale_modifier"]), use_sa=case["use_sa"])
    g = O.backward(st, case["means3D"], case["viewmatrix"], case["projmatrix"], case["W"], case["H"],
                   float(case["tanfovx"]), float(case["tanfovy"]), case["bg"], case["dL_dcolor"], case["dL_dothers"],
                   scales=case["scales"], rotations=case["rotations"], transMat=case["transMat"], shs=case["shs"],
                   sh_degree=case["sh_degree"], campos=case["campos"], use_sa=case["use_sa"])
    out = dict(st)
    out["transMat"] = st["transMat_used"]
    for k, v in g.items():
        out["grad_" + k] = v
    return out


@pytest.mark.parametrize("name", cases.GOLDEN_CASES)
def test_oracle_matches_reference_outputs(name):
    g = cases.load_golden(name)
    assert g is not None, f"tests/golden/{name}.npz missing"
    case = cases.case_from_golden(g)
    ref = _ref(g)
    got = _run_oracle(case)
    # integer stages: exact (the restatement spells out the reference's fused multiply-adds)
    fails = compare.check_binning(got, ref)
    fails += compare.check_images_oracle(got, ref, case["use_sa"])
    fails += compare.check_grads(got, ref, rtol=compare.ORACLE_SA_GRAD_RTOL if case["use_sa"] else compare.GRAD_RTOL)
    assert not fails, "\n".join(fails)


@pytest.mark.parametrize("name", cases.GOLDEN_CASES)
def test_oracle_stages_from_reference_state(name):
    """Each oracle stage fed with the REFERENCE's previous-stage output, so a stage cannot hide behind another."""
    g = cases.load_golden(name)
    case = cases.case_from_golden(g)
    ref = _ref(g)
    W, H = case["W"], case["H"]
    st = dict(radii=ref["radii"], means2D=ref["means2D"], depths=ref["depths"], tiles_touched=ref["tiles_touched"])
    b = O.binning(st, W, H)
    assert b["num_rendered"] == int(ref["num_rendered"])
    if b["num_rendered"]:
        for k in ("point_offsets", "keys_unsorted", "values_unsorted", "keys_sorted", "point_list", "ranges"):
            assert np.array_equal(b[k], ref[k]), k
    feats = case["colors"] if case["colors"] is not None else None
    if feats is None:
        return
    r = O.render_forward(W, H, ref["ranges"], ref["point_list"] if b["num_rendered"] else np.zeros(1, np.uint32),
                         ref["means2D"], feats, ref["transMat"], ref["normal_opacity"], case["bg"], case["use_sa"])
    # same per-pixel operation order as the reference kernel: images agree to the last few ulps
    assert not compare.check_images_oracle(r, ref, case["use_sa"])
    for k in ("color", "final_T"):       # well-conditioned channels: a few ulps
        assert np.abs(r[k] - ref[k]).max() <= 1e-5, k
    assert np.array_equal(r["n_contrib"][0], ref["n_contrib"][0])


def test_sort_pairs_is_stable_lsd():
    rng = np.random.default_rng(0)
    keys = (rng.integers(0, 1 << 12, size=20000).astype(np.uint64) << np.uint64(32)) | \
        rng.integers(0, 1 << 32, size=20000).astype(np.uint64)
    keys[::7] = keys[0]
    vals = np.arange(20000, dtype=np.uint32)
    ks, vs = O.sort_pairs(keys, vals, 44)
    order = np.argsort(keys & np.uint64((1 << 44) - 1), kind="stable")
    assert np.array_equal(ks, keys[order]) and np.array_equal(vs, vals[order])


def test_higher_msb():
    # CR/rasterizer_impl.cu:35-50 for the tile counts of the BASELINE configs
    assert O.higher_msb(75 * 43) == 12 and O.higher_msb(110 * 73) == 13
    assert O.higher_msb(1) == 1 and O.higher_msb(4096) == 13


@pytest.mark.parametrize("name", cases.KNN_CASES)
def test_oracle_dist2(name):
    g = cases.load_golden(name)
    assert g is not None
    got = O.dist2(g["in_points"])
    assert compare.bits_differ(got, g["ref_dist2"]) == 0


def test_oracle_mark_visible():
    g = cases.load_golden("mark_visible")
    got = O.mark_visible(g["in_means3D"], g["in_viewmatrix"])
    assert np.array_equal(got, g["ref_present"])
