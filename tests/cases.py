"""Small deterministic parity cases shared by tests/golden/make_golden.py and the tests.

Every case is a dict of numpy float32 inputs plus scalar settings in the vocabulary of the reference's
`_C.rasterize_gaussians` call (RAST/rasterize_points.cu:39-61).  The golden fixtures under tests/golden/
store these inputs AND the outputs the unmodified reference CUDA rasterizer produced for them on a B200,
so the tests never depend on regenerating the inputs bit-for-bit.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "gaus-slam_b200"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

from gs2d_b200 import scene  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def _intr(W, H, f=None, cx=None, cy=None):
    f = f if f is not None else 0.5 * W
    return dict(W=W, H=H, fx=float(f), fy=float(f), cx=(W - 1) / 2 if cx is None else cx,
                cy=(H - 1) / 2 if cy is None else cy)


def _base(P, intr, seed, view, n_views=3, scale_mult=1.0, use_sa=True, bg=(0.0, 0.0, 0.0)):
    sc = scene.room_scene(P, intr, seed=seed, scale_mult=scale_mult)
    cam = scene.view_sweep(n_views, intr, seed=seed)[view]
    gc, go = scene.image_grads(intr["W"], intr["H"], seed=seed)
    return dict(
        W=intr["W"], H=intr["H"], tanfovx=np.float32(cam.tanfovx), tanfovy=np.float32(cam.tanfovy),
        viewmatrix=cam.viewmatrix.reshape(16).copy(), projmatrix=cam.projmatrix.reshape(16).copy(),
        campos=cam.campos.copy(), bg=np.asarray(bg, np.float32), use_sa=bool(use_sa), scale_modifier=np.float32(1.0),
        sh_degree=0, means3D=sc["means3D"], scales=sc["scales"], rotations=sc["rotations"],
        opacities=sc["opacities"], colors=sc["colors"], shs=None, transMat=None,
        dL_dcolor=gc, dL_dothers=go)


def room_sa():
    """Yawed view of the room, surface-aware depth (the configuration gaus-slam always uses)."""
    return _base(3000, _intr(160, 96), seed=1, view=1)


def room_dist():
    """Same scene, 2DGS distortion branch (use_sa=False), non-zero background."""
    return _base(3000, _intr(160, 96), seed=1, view=2, use_sa=False, bg=(0.2, 0.5, 0.7))


def ragged_sa():
    """Image size not a multiple of 16, principal point off-centre, focal length for which the backward's
    float re-derivation of W,H (CR/backward.cu:641-642) is exercised."""
    c = _base(2500, _intr(171, 93, f=90.3, cx=80.2, cy=47.9), seed=2, view=1, bg=(0.1, 0.0, 0.3))
    return c


def dense_sa():
    """Large splats: several hundred instances per tile, so tiles need more than one staging batch and
    most pixels terminate early (T < 1e-4)."""
    c = _base(4000, _intr(96, 64), seed=3, view=0, scale_mult=5.0)
    c["opacities"] = np.clip(c["opacities"] * 1.5, 0.0, 1.0).astype(np.float32)
    return c


def sh_deg2():
    """Colours from spherical harmonics (degree 2, 9 coefficients): the cold SH branch
    (CR/forward.cu:20-71, CR/backward.cu:20-139)."""
    c = _base(1500, _intr(128, 80), seed=4, view=1)
    rng = np.random.default_rng(44)
    c["shs"] = (rng.normal(0.0, 0.4, size=(1500, 9, 3))).astype(np.float32)
    c["colors"] = None
    c["sh_degree"] = 2
    return c


def scale_mod():
    """scale_modifier != 1 (forward applies it, backward ignores it: CR/backward.cu:504)."""
    c = _base(2000, _intr(128, 80), seed=5, view=0)
    c["scale_modifier"] = np.float32(1.7)
    return c


def behind():
    """Every surfel behind the near plane: num_rendered == 0, images are background only."""
    c = _base(500, _intr(64, 48), seed=6, view=0, bg=(0.3, 0.6, 0.9))
    c["means3D"] = c["means3D"].copy()
    c["means3D"][:, 2] = -np.abs(c["means3D"][:, 2]) - 0.5
    return c


CASES = dict(room_sa=room_sa, room_dist=room_dist, ragged_sa=ragged_sa, dense_sa=dense_sa, sh_deg2=sh_deg2,
             scale_mod=scale_mod, behind=behind)

# precomputed-transMat case is derived from room_sa's golden transMat inside make_golden.py ("transmat_pre")
GOLDEN_CASES = list(CASES) + ["transmat_pre"]


def knn_points(name):
    rng = np.random.default_rng(90)
    if name == "knn_room":
        return scene.room_scene(5000, _intr(160, 96), seed=9)["means3D"]
    if name == "knn_small":
        return rng.normal(size=(37, 3)).astype(np.float32)
    if name == "knn_dups":
        p = rng.uniform(-2, 2, size=(3000, 3)).astype(np.float32)
        p[100:200] = p[0:100]            # exact duplicates -> zero distances
        p[1500:] += np.float32(5.0)      # second cluster, positive octant
        return p
    if name == "knn_positive":
        return rng.uniform(1.0, 3.0, size=(2048, 3)).astype(np.float32)  # min>0: the init={0,0,0} quirk
    raise KeyError(name)


KNN_CASES = ["knn_room", "knn_small", "knn_dups", "knn_positive"]


def golden_path(name):
    return os.path.join(GOLDEN_DIR, name + ".npz")


def load_golden(name):
    """-> dict of numpy arrays (inputs under 'in_*', reference outputs under 'ref_*'), or None if absent."""
    p = golden_path(name)
    if not os.path.exists(p):
        return None
    with np.load(p, allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


def case_from_golden(g):
    """Rebuild the input dict from a golden file."""
    c = {}
    for k, v in g.items():
        if k.startswith("in_"):
            c[k[3:]] = v
    for k in ("shs", "colors", "transMat", "scales", "rotations"):
        if k not in c or c[k].size == 0:
            c[k] = None
    for k in ("W", "H", "sh_degree"):
        c[k] = int(c[k])
    c["use_sa"] = bool(c["use_sa"])
    return c
