#!/usr/bin/env python
"""Pair statistics of the bench workload (design input for the compositing kernels; GPU only).

Reads the per-pixel hit words the forward pass leaves for the backward pass (`hit_pix`, csrc/common.cuh) and
prints: footprint pairs, contributing pairs, list lengths per pixel, lane efficiency of the per-pixel walks for
the thread<->pixel layouts in use and for a length-sorted layout, pairs per 128-instance batch, tile sizes.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gaus-slam_b200"))
from gs2d_b200 import scene  # noqa: E402
import gaus_2dgs_rasterization._C as GC  # noqa: E402

ALIGN = 256


def carve(sizes):
    off, out = 0, []
    for nbytes in sizes:
        off = (off + ALIGN - 1) // ALIGN * ALIGN
        out.append(off)
        off += nbytes
    return out


def pct(a):
    return {"mean": float(a.mean()), "p50": float(np.percentile(a, 50)), "p90": float(np.percentile(a, 90)),
            "p99": float(np.percentile(a, 99)), "max": int(a.max())}


def render(P, view, dev):
    intr = scene.REPLICA
    W, H = intr["W"], intr["H"]
    sc = scene.room_scene(P, intr, seed=0)
    cam = scene.view_sweep(8, intr, seed=0)[view]
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)
    e = torch.empty(0, device=dev)
    r = GC.rasterize_gaussians(
        torch.zeros(3, device=dev), t(sc["means3D"]), t(sc["colors"]), t(sc["opacities"]), t(sc["scales"]),
        t(sc["rotations"]), 1.0, e, t(cam.viewmatrix.reshape(16)), t(cam.projmatrix.reshape(16)), float(cam.tanfovx),
        float(cam.tanfovy), H, W, e, 0, t(cam.campos), True, False, False)
    torch.cuda.synchronize()
    return W, H, r


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 500_000
    view = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    dev = torch.device("cuda:0")
    W, H, (n, color, others, radii, geom, binning, img) = render(P, view, dev)
    I = int(n)
    gx, gy = (W + 15) // 16, (H + 15) // 16
    tiles, px = gx * gy, W * H
    # byte offsets of the carved state buffers (csrc/api.cu: ImageState::carve, BinningState::carve)
    io = carve([12 * px, 8 * px, 4 * px, 4 * px, 8 * tiles])
    ranges = img[io[4]:io[4] + 8 * tiles].view(torch.int32).cpu().numpy().reshape(tiles, 2).astype(np.int64)
    hp_words = (I // 32 + 8 * tiles + 8) * 512
    bo = carve([8 * I, 8 * I, 4 * I, 4 * I, 80 * I, 4 * (8 * I + 64), 4 * hp_words])
    hp = binning[bo[6]:bo[6] + 4 * hp_words].view(torch.int32).cpu().numpy().view(np.uint32).reshape(-1, 2, 256)
    nt = ranges[:, 1] - ranges[:, 0]
    out = {"P": P, "view": view, "I": I, "tiles": tiles, "tile_n": pct(nt)}
    for k in (128, 256, 384):
        out["tile_n"][f"frac_le{k}"] = float((nt <= k).mean())
    tot_foot = tot_used = 0
    walk = {"rows": 0, "blk8x4": 0, "sorted": 0, "sorted_tile": 0, "tileperm_batch128": 0, "tileperm_batch256": 0,
            "blk8x4_batch256": 0, "blk8x4_tile": 0, "ideal": 0}
    per_pixel_len, batch_pairs = [], []
    # phase B of the backward pass (a group of lanes per instance, lanes over the instance's pair slots): warp iterations
    phb = {"grp8": 0, "grp8_sorted": 0, "grp4": 0, "grp4_sorted": 0, "grp16_sorted": 0, "adaptive_sorted": 0, "ideal": 0}
    inst_slots = []
    lane = np.arange(256)
    # thread -> row-major pixel id of the 8x4 block layout (pair.cuh thread_lx / thread_ly with BLOCK = 1)
    blk_pix = ((((lane >> 6) << 2) | ((lane >> 3) & 3)) << 4) | ((((lane >> 5) & 1) << 3) | (lane & 7))
    for tix in range(tiles):
        nn = int(nt[tix])
        if nn == 0:
            continue
        nb = (nn + 127) // 128
        base = (int(ranges[tix, 0]) >> 5) + 8 * tix
        w = hp[base:base + 4 * nb]
        tot_used += int(np.bitwise_count(w[:, 0, :]).sum())
        foot = np.bitwise_count(w[:, 1, :]).astype(np.int64)
        tot_foot += int(foot.sum())
        fb = foot.reshape(nb, 4, 256).sum(1)            # [batch, pixel]
        per_pixel_len.append(fb.sum(0))
        batch_pairs.extend(fb.sum(1).tolist())
        walk["rows"] += int(fb.reshape(nb, 8, 32).max(2).sum()) * 32
        walk["blk8x4"] += int(fb[:, blk_pix].reshape(nb, 8, 32).max(2).sum()) * 32
        walk["sorted"] += int(np.sort(fb, axis=1).reshape(nb, 8, 32).max(2).sum()) * 32
        walk["sorted_tile"] += int(np.sort(fb.sum(0)).reshape(8, 32).max(1).sum()) * 32
        # one pixel -> thread permutation per tile (by whole-tile list length), walk still in lockstep per batch
        order = np.argsort(fb.sum(0), kind="stable")
        walk["tileperm_batch128"] += int(fb[:, order].reshape(nb, 8, 32).max(2).sum()) * 32
        fb2 = fb if nb % 2 == 0 else np.concatenate([fb, np.zeros((1, 256), fb.dtype)])
        fb2 = fb2.reshape(-1, 2, 256).sum(1)
        walk["tileperm_batch256"] += int(fb2[:, order].reshape(-1, 8, 32).max(2).sum()) * 32
        walk["blk8x4_batch256"] += int(fb2[:, blk_pix].reshape(-1, 8, 32).max(2).sum()) * 32
        walk["blk8x4_tile"] += int(fb.sum(0)[blk_pix].reshape(8, 32).max(1).sum()) * 32
        walk["ideal"] += int(fb.sum())
        # slots per instance: pixels whose footprint word has the instance's bit set
        bits = (w[:, 1, :, None] >> np.arange(32, dtype=np.uint32)[None, None, :]) & 1     # [word, pixel, bit]
        sl = bits.sum(1).reshape(-1)[:nn].astype(np.int64)
        inst_slots.append(sl)
        phb["ideal"] += int(sl.sum())

        def warp_iters(s_, lanes):
            per_warp = 32 // lanes
            it = -(-s_ // lanes)
            pad = (-len(it)) % per_warp
            it = np.concatenate([it, np.zeros(pad, it.dtype)]).reshape(-1, per_warp)
            return int(it.max(1).sum()) * 32
        phb["grp8"] += warp_iters(sl, 8)
        phb["grp4"] += warp_iters(sl, 4)
        ss = np.sort(sl)[::-1]
        phb["grp8_sorted"] += warp_iters(ss, 8)
        phb["grp4_sorted"] += warp_iters(ss, 4)
        phb["grp16_sorted"] += warp_iters(ss, 16)
        # lanes per instance chosen by size class (<= 4 slots: 4 lanes ... > 32: 32 lanes), classes processed separately
        tot = 0
        for lanes, lo, hi in ((4, 0, 8), (8, 8, 24), (16, 24, 64), (32, 64, 1 << 30)):
            c = ss[(ss > lo) & (ss <= hi)]
            if len(c):
                tot += warp_iters(c, lanes)
        phb["adaptive_sorted"] += tot
    bp = np.array(batch_pairs)
    out.update(footprint_pairs=tot_foot, used_pairs=tot_used, used_over_footprint=tot_used / max(tot_foot, 1),
               pairs_per_instance=tot_foot / max(I, 1), pixel_list_len=pct(np.concatenate(per_pixel_len)),
               batch_pairs=pct(bp),
               lane_efficiency={k: walk["ideal"] / v for k, v in walk.items() if k != "ideal" and v},
               phase_b_lane_efficiency={k: phb["ideal"] / v for k, v in phb.items() if k != "ideal" and v},
               slots_per_instance=pct(np.concatenate(inst_slots)))
    for k in (2048, 3072, 4096):
        out["batch_pairs"][f"frac_gt{k}"] = float((bp > k).mean())
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
