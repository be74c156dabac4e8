"""Stage-by-stage parity report: libgs2d (product) vs the compiled reference (oracle/_ref) on a GPU.

Run on the GPU box:  python tests/parity_report.py [--P 100000] [--size 1200x680] [--views 2] [--no-sa]
Prints, for every stage, how many elements differ and the worst deviation; writes JSON to
gpurun_out/parity_<tag>.json.  This is a diagnostic script; the pass/fail tests are tests/test_gpu_*.py.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gaus-slam_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

from gs2d_b200 import _lib as L, scene  # noqa: E402
import gaus_2dgs_rasterization._C as GC  # noqa: E402
import refgpu  # noqa: E402


def t(a, dev="cuda"):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def view_state(lib, s, bufs, I, dev):
    sv = L.StateView()
    L.check(lib.gs2d_state_view_get(C.byref(s), C.byref(bufs), I, C.byref(sv)), "state_view")
    return sv


def raw(ptr, shape, dtype, dev):
    return refgpu._view(ptr, shape, dtype, dev)


def cmp_exact(name, a, b, rep):
    a, b = a.cpu(), b.cpu()
    nd = int((a != b).sum())
    rep[name] = dict(n=a.numel(), differ=nd)
    print(f"  {name:28s} n={a.numel():>10d} differ={nd}")
    return nd


def cmp_float(name, a, b, rep, bits=True):
    a, b = a.float().cpu(), b.float().cpu()
    d = (a - b).abs()
    d = torch.nan_to_num(d, nan=float("inf"))
    mx = float(d.max()) if d.numel() else 0.0
    nbit = int((a.view(torch.int32) != b.view(torch.int32)).sum()) if bits else -1
    denom = float(b.abs().max()) if b.numel() else 0.0
    rep[name] = dict(n=a.numel(), max_abs=mx, bit_differ=nbit, ref_absmax=denom)
    print(f"  {name:28s} n={a.numel():>10d} max_abs={mx:.3e} bit_differ={nbit} ref_absmax={denom:.3e}")
    return mx


def rel_err(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30)), float((a - b).abs().max()), float(b.abs().max())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--P", type=int, default=100000)
    ap.add_argument("--size", default="1200x680")
    ap.add_argument("--views", type=int, default=2)
    ap.add_argument("--no-sa", action="store_true")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--scale-mult", type=float, default=1.0)
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    W, H = map(int, a.size.split("x"))
    intr = dict(scene.REPLICA)
    intr.update(W=W, H=H, fx=600.0 * W / 1200, fy=600.0 * W / 1200, cx=(W - 1) / 2, cy=(H - 1) / 2)
    dev = torch.device("cuda:0")
    lib = L.load()
    use_sa = not a.no_sa
    sc = scene.room_scene(a.P, intr, seed=a.seed, scale_mult=a.scale_mult)
    cams = scene.view_sweep(a.views, intr, seed=a.seed)
    gcol, goth = scene.image_grads(W, H, seed=a.seed)
    means, scales, rots, opac, cols = (t(sc[k]) for k in ("means3D", "scales", "rotations", "opacities", "colors"))
    bg = torch.zeros(3, device=dev)
    gcol, goth = t(gcol), t(goth)
    empty = torch.empty(0, device=dev)
    report = dict(P=a.P, W=W, H=H, use_sa=use_sa, views=[])
    ref = refgpu.RefRasterizer()

    for vi, cam in enumerate(cams):
        print(f"== view {vi}  (P={a.P}, {W}x{H}, use_sa={use_sa})")
        rep = {}
        view, proj, campos = t(cam.viewmatrix), t(cam.projmatrix), t(cam.campos)
        # ---- reference
        n_ref, c_ref, o_ref, r_ref = ref.forward(bg=bg, means3D=means, opacities=opac, viewmatrix=view,
                                                 projmatrix=proj, campos=campos, W=W, H=H, tanfovx=cam.tanfovx,
                                                 tanfovy=cam.tanfovy, colors=cols, scales=scales, rotations=rots,
                                                 use_sa=use_sa)
        torch.cuda.synchronize()
        rs = ref.state()
        g_ref = ref.backward(bg=bg, means3D=means, radii=r_ref, viewmatrix=view, projmatrix=proj, campos=campos,
                             W=W, H=H, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, dL_dcolor=gcol, dL_dothers=goth,
                             colors=cols, scales=scales, rotations=rots, use_sa=use_sa)
        torch.cuda.synchronize()
        # ---- ours (debug=True keeps the unsorted keys for inspection)
        n, c, o, r, geom, binning, img = GC.rasterize_gaussians(
            bg, means, cols, opac, scales, rots, 1.0, empty, view, proj, cam.tanfovx, cam.tanfovy, H, W, empty, 0,
            campos, use_sa, False, True)
        torch.cuda.synchronize()
        g = GC.rasterize_gaussians_backward(bg, means, r, cols, scales, rots, 1.0, empty, view, proj, cam.tanfovx,
                                            cam.tanfovy, gcol, goth, empty, 0, campos, geom, n, binning, img,
                                            use_sa, False)
        torch.cuda.synchronize()
        rep["num_rendered"] = dict(ours=n, ref=n_ref)
        print(f"  num_rendered ours={n} ref={n_ref}  visible={int((r_ref > 0).sum())}")

        s = GC._settings(a.P, W, H, 0, 0, use_sa, False, True, cam.tanfovx, cam.tanfovy, 1.0)
        bufs = L.Buffers(geom=geom.data_ptr(), geom_bytes=geom.numel(), binning=binning.data_ptr() if n else None,
                         binning_bytes=binning.numel(), image=img.data_ptr(), image_bytes=img.numel())
        sv = view_state(lib, s, bufs, n, dev)
        P = a.P
        vis = (r_ref > 0)
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        # ---- stage 1
        cmp_exact("radii", r, r_ref, rep)
        cmp_exact("tiles_touched", raw(sv.tiles_touched, (P,), torch.int32, dev), rs["tiles_touched"], rep)
        cmp_exact("point_offsets", raw(sv.point_offsets, (P,), torch.int32, dev), rs["point_offsets"], rep)
        splat = raw(sv.splat, (P, 20), torch.float32, dev)
        mine_T = torch.cat([splat[:, 0:3], splat[:, 4:7], splat[:, 8:11]], dim=1)
        cmp_float("depths(visible)", raw(sv.depths, (P,), torch.float32, dev)[vis], rs["depths"][vis], rep)
        cmp_float("transMat(visible)", mine_T[vis], rs["transMat"][vis], rep)
        cmp_float("means2D(visible)", torch.stack([splat[:, 3], splat[:, 7]], 1)[vis], rs["means2D"][vis], rep)
        cmp_float("normal(visible)", splat[:, 12:15][vis], rs["normal_opacity"][:, :3][vis], rep)
        # ---- stage 2 (binning)
        if n == n_ref and n > 0:
            cmp_exact("keys_unsorted", raw(sv.keys_unsorted, (n,), torch.int64, dev), rs["keys_unsorted"], rep)
            cmp_exact("values_unsorted", raw(sv.values_unsorted, (n,), torch.int32, dev), rs["values_unsorted"], rep)
            cmp_exact("keys_sorted", raw(sv.keys_sorted, (n,), torch.int64, dev), rs["keys_sorted"], rep)
            cmp_exact("point_list", raw(sv.point_list, (n,), torch.int32, dev), rs["point_list"], rep)
        cmp_exact("ranges", raw(sv.ranges, (tiles, 2), torch.int32, dev), rs["ranges"], rep)
        # ---- images
        cmp_float("out_color", c, c_ref, rep)
        names = ["depth", "alpha", "normal_x", "normal_y", "normal_z", "median_depth", "distortion"]
        for k, nm in enumerate(names):
            cmp_float("allmap." + nm, o[k], o_ref[k], rep)
        nc_m, nc_r = raw(sv.n_contrib, (2, H, W), torch.int32, dev), rs["n_contrib"]
        if cmp_exact("n_contrib", nc_m, nc_r, rep):
            for ch in range(2):
                bad = (nc_m[ch] != nc_r[ch]).nonzero()
                print(f"    channel {ch}: {bad.shape[0]} differ; samples (y,x,ours,ref):",
                      [(int(y), int(x), int(nc_m[ch, y, x]), int(nc_r[ch, y, x])) for y, x in bad[:6].tolist()])
        cmp_float("final_T", raw(sv.final_T, (3, H, W), torch.float32, dev), rs["final_T"], rep)
        cmp_float("depth_std", raw(sv.depth_std, (H, W), torch.float32, dev), rs["depth_std"], rep)
        # ---- gradients
        names = ["means2D", "colors", "opacity", "means3D", "transMat", "sh", "scales", "rotations"]
        for nm, mine in zip(names, g):
            if mine.numel() == 0:
                continue
            rel, mx, refmx = rel_err(mine, g_ref[nm])
            finite = bool(torch.isfinite(mine).all())
            rep["grad." + nm] = dict(rel_l2=rel, max_abs=mx, ref_absmax=refmx, finite=finite)
            print(f"  grad.{nm:23s} rel_l2={rel:.3e} max_abs={mx:.3e} ref_absmax={refmx:.3e} finite={finite}")
        # instances per tile
        rng_ = rs["ranges"].cpu().long()
        cnt = (rng_[:, 1] - rng_[:, 0])
        ncon = rs["n_contrib"][0].float()
        rep["stats"] = dict(inst_per_tile_mean=float(cnt.float().mean()), inst_per_tile_max=int(cnt.max()),
                            last_contrib_mean=float(ncon.mean()))
        print(f"  stats: {rep['stats']}")
        report["views"].append(rep)

    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    tag = a.tag or f"P{a.P}_{W}x{H}_{'sa' if use_sa else 'dist'}"
    with open(os.path.join(ROOT, "gpurun_out", f"parity_{tag}.json"), "w") as f:
        json.dump(report, f, indent=1)


if __name__ == "__main__":
    main()
