"""Generate the golden fixtures under tests/golden/ from the UNMODIFIED reference CUDA code.

Run on a GPU box (the reference has no CPU path):
    python tests/golden/make_golden.py [--out gpurun_out/golden]
It executes oracle/_ref/libgs2d_ref.so -- the reference's forward.cu / backward.cu / rasterizer_impl.cu /
simple_knn.cu compiled as they are for sm_100a by oracle/build.py -- on the small cases of tests/cases.py and
stores inputs (`in_*`) and reference outputs (`ref_*`) as compressed .npz.  The files committed under
tests/golden/ are exactly what this script wrote on a B200; the CPU oracle (oracle/gs2d_oracle.c) and the
product (libgs2d.so) are both tested against them.

Per-surfel state of culled surfels (radii == 0) is uninitialised memory in the reference; it is zeroed
here so the fixtures are deterministic.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

import cases  # noqa: E402
import gpu_run  # noqa: E402
import refgpu  # noqa: E402

INPUT_KEYS = ["W", "H", "tanfovx", "tanfovy", "viewmatrix", "projmatrix", "campos", "bg", "use_sa", "scale_modifier",
              "sh_degree", "means3D", "scales", "rotations", "opacities", "colors", "shs", "transMat", "dL_dcolor",
              "dL_dothers"]


def pack(case, ref):
    z = {}
    for k in INPUT_KEYS:
        v = case[k]
        z["in_" + k] = np.zeros(0, np.float32) if v is None else np.asarray(v)
    vis = ref["radii"] > 0
    for k in ("depths", "means2D", "transMat", "normal_opacity"):
        a = ref[k].copy()
        a[~vis] = 0
        ref[k] = a
    for k, v in ref.items():
        z["ref_" + k] = np.asarray(v)
    return z


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=cases.GOLDEN_DIR)
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)
    assert torch.cuda.is_available() and refgpu.available()
    room_ref = None
    for name in cases.GOLDEN_CASES:
        if name == "transmat_pre":
            # precomputed transform matrices: feed room_sa's reference transMat back in (scales/rotations absent)
            case = cases.room_sa()
            tm = room_ref["transMat"].copy()
            tm[room_ref["radii"] <= 0] = 0
            case.update(transMat=tm, scales=None, rotations=None)
        else:
            case = cases.CASES[name]()
        ref = gpu_run.run_ref(case)
        if name == "room_sa":
            room_ref = {k: np.array(v) for k, v in ref.items()}
        if name == "transmat_pre":
            ref["transMat"] = case["transMat"].copy()   # the reference never copies a precomputed transMat into its state
        z = pack(case, ref)
        np.savez_compressed(os.path.join(a.out, name + ".npz"), **z)
        print(f"{name:14s} P={case['means3D'].shape[0]} {case['W']}x{case['H']} num_rendered={int(ref['num_rendered'])} "
              f"visible={int((ref['radii'] > 0).sum())} bytes={os.path.getsize(os.path.join(a.out, name + '.npz'))}")
    for name in cases.KNN_CASES:
        pts = cases.knn_points(name)
        d = refgpu.dist_cuda2(torch.from_numpy(pts).cuda())
        torch.cuda.synchronize()
        np.savez_compressed(os.path.join(a.out, name + ".npz"), in_points=pts, ref_dist2=d.cpu().numpy())
        print(f"{name:14s} P={pts.shape[0]} mean={float(d.mean()):.6g}")
    # markVisible
    case = cases.room_sa()
    rng = np.random.default_rng(17)
    near = rng.uniform(-1.0, 1.0, size=(2000, 3)).astype(np.float32)      # straddles the z = 0.2 plane
    case["means3D"] = np.concatenate([case["means3D"], cases.behind()["means3D"], near], axis=0)
    dev = gpu_run.to_device(case)
    vis = refgpu.mark_visible(dev["means3D"], dev["viewmatrix"], dev["projmatrix"])
    np.savez_compressed(os.path.join(a.out, "mark_visible.npz"), in_means3D=case["means3D"],
                        in_viewmatrix=case["viewmatrix"], in_projmatrix=case["projmatrix"],
                        ref_present=vis.cpu().numpy())
    print("mark_visible", int(vis.sum()), "of", vis.numel())


if __name__ == "__main__":
    main()
