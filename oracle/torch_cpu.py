"""TEST / BASELINE INFRASTRUCTURE ONLY (never imported by the product): a pure-PyTorch CPU transliteration of the
reference's 2DGS surfel rasterizer, forward and hand-written backward (BASELINE.md baseline B1, SURVEY.md 8c form 2).

Why hand-written backward: the reference's backward is NOT the analytic gradient of its forward (surface-aware
confidence from the final per-pixel statistics, CR/backward.cu:349; dL_dmean2D overwritten by the densification
proxy, :660-663), so torch autograd over the forward would not reproduce it.

Vectorisation: per-surfel stages over [P]; compositing over [tiles, 256] pixel lanes, one python iteration per
"k-th instance of every tile" (tiles are ordered by list length so the active tiles are always a prefix and every
iteration works on slices).  The python loop is as long as the longest tile list, not the number of instances.
Per-instance gradients are summed over the 256 pixels of the tile and scattered with index_add_ (deterministic).

Covers what gaus-slam calls (colors_precomp, scales + rotations, sh_degree 0: render/render_2dgs.py:16-30) with
use_sa on or off; the SH and precomputed-transMat branches raise NotImplementedError (the C restatement has them).
Checked against the C restatement (oracle/gs2d_oracle.c, itself pinned by the golden vectors of the compiled
reference) in tests/test_torch_cpu.py: same instance lists up to fp32 rounding of the bounding boxes, images 1e-4,
gradients 1e-3.  Plain fp32 torch ops: no attempt at the compiled kernel's FMA placement (that is the C oracle's job).
"""
from __future__ import annotations

import numpy as np
import torch

TILE = 16
NEAR_N, FAR_N, FILTER_INV_SQUARE = 0.2, 100.0, 100.0


def _t(a, dtype=torch.float32):
    return torch.as_tensor(np.ascontiguousarray(a) if isinstance(a, np.ndarray) else a, dtype=dtype)


def _xform_point(p, m):     # transformPoint4x3 (CR/auxiliary.h:78-86), m[4c + r]
    return torch.stack([m[12 + r] + p[:, 0] * m[r] + p[:, 1] * m[4 + r] + p[:, 2] * m[8 + r] for r in range(3)], 1)


def _xform_vec(v, m):       # transformVec4x3 (:100-108)
    return torch.stack([v[:, 0] * m[r] + v[:, 1] * m[4 + r] + v[:, 2] * m[8 + r] for r in range(3)], 1)


def _xform_vec_T(v, m):     # transformVec4x3Transpose (:110-118)
    return torch.stack([m[4 * r] * v[:, 0] + m[4 * r + 1] * v[:, 1] + m[4 * r + 2] * v[:, 2] for r in range(3)], 1)


def _quat_to_rot(q):
    """quat_to_rotmat (CR/auxiliary.h:212-234): columns R0, R1, R2 of the rotation, q = (w, x, y, z) normalised here."""
    s = torch.rsqrt((q * q).sum(1, keepdim=True))
    w, x, y, z = (q * s).unbind(1)
    R0 = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y + w * z), 2 * (x * z - w * y)], 1)
    R1 = torch.stack([2 * (x * y - w * z), 1 - 2 * (x * x + z * z), 2 * (y * z + w * x)], 1)
    R2 = torch.stack([2 * (x * z + w * y), 2 * (y * z - w * x), 1 - 2 * (x * x + y * y)], 1)
    return R0, R1, R2


def _trunc_clamp(v, hi):
    """min(hi, max(0, (int)v)) with C truncation (getRect, CR/auxiliary.h:66-76)."""
    return torch.clamp(torch.trunc(torch.nan_to_num(v, nan=0.0, posinf=2e9, neginf=-2e9)).to(torch.int64), 0, hi)


def preprocess(means3D, opacities, viewmatrix, projmatrix, W, H, scales, rotations, scale_modifier=1.0):
    """preprocessCUDA (CR/forward.cu:150-253) over all surfels at once."""
    p, op = _t(means3D), _t(opacities).reshape(-1)
    vm, pm = _t(viewmatrix).reshape(16), _t(projmatrix).reshape(16)
    sc, q = _t(scales), _t(rotations)
    P = p.shape[0]
    gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    pv = _xform_point(p, vm)
    ok = pv[:, 2] > 0.2
    R0, R1, R2 = _quat_to_rot(q)
    L0, L1 = R0 * (scale_modifier * sc[:, 0:1]), R1 * (scale_modifier * sc[:, 1:2])
    # compute_transmat (:75-115): rows of T = (splat2world)^T * world2ndc * ndc2pix
    X = []
    for c in range(4):
        col = torch.stack([pm[c], pm[4 + c], pm[8 + c]])
        X.append(torch.stack([L0 @ col, L1 @ col, pm[12 + c] + p @ col], 1))
    Tu = (W - 1) * 0.5 * X[3] + W * 0.5 * X[0]
    Tv = (H - 1) * 0.5 * X[3] + H * 0.5 * X[1]
    Tw = X[3]
    normal = _xform_vec(R2, vm)
    pdotn = (pv * normal).sum(1)
    ok &= pdotn != 0
    normal = normal * torch.where(pdotn < 0, 1.0, -1.0)[:, None]
    # compute_aabb at cutoff 3 (:119-147)
    tvec = torch.tensor([9.0, 9.0, -1.0])
    d = (Tw * Tw * tvec).sum(1)
    ok &= d != 0
    f = tvec / torch.where(d == 0, torch.ones_like(d), d)[:, None]
    cx, cy = (f * Tu * Tw).sum(1), (f * Tv * Tw).sum(1)
    ex = torch.sqrt(torch.clamp(cx * cx - (f * Tu * Tu).sum(1), min=1e-4))
    ey = torch.sqrt(torch.clamp(cy * cy - (f * Tv * Tv).sum(1), min=1e-4))
    radius = torch.ceil(torch.maximum(ex, ey))
    x0, y0 = _trunc_clamp((cx - radius) / TILE, gx), _trunc_clamp((cy - radius) / TILE, gy)
    x1, y1 = _trunc_clamp((cx + radius + TILE - 1) / TILE, gx), _trunc_clamp((cy + radius + TILE - 1) / TILE, gy)
    touched = (x1 - x0) * (y1 - y0)
    ok &= touched > 0
    z = torch.zeros(())
    radii = torch.where(ok, torch.nan_to_num(radius, nan=0.0, posinf=2e9).to(torch.int64), 0).to(torch.int32)
    return dict(radii=radii, means2D=torch.where(ok[:, None], torch.stack([cx, cy], 1), z),
                depths=torch.where(ok, pv[:, 2], z),
                transMat=torch.where(ok[:, None], torch.cat([Tu, Tv, Tw], 1), z),
                normal_opacity=torch.where(ok[:, None], torch.cat([normal, op[:, None]], 1), z),
                tiles_touched=torch.where(ok, touched, 0), rect=torch.stack([x0, y0, x1, y1], 1), visible=ok)


def binning(st, W, H):
    """InclusiveSum + duplicateWithKeys + stable SortPairs + identifyTileRanges (CR/rasterizer_impl.cu:70-138, :283-323)."""
    gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    tt = st["tiles_touched"]
    offs = torch.cumsum(tt, 0)
    n = int(offs[-1]) if tt.numel() else 0
    idx = torch.repeat_interleave(torch.arange(tt.numel()), tt)              # surfel of every instance, emission order
    first = (offs - tt)[idx]
    e = torch.arange(n) - first                                               # position inside the surfel's rect
    r = st["rect"][idx]
    w = r[:, 2] - r[:, 0]
    yy = torch.div(e, torch.clamp(w, min=1), rounding_mode="floor")
    tile = (r[:, 1] + yy) * gx + r[:, 0] + (e - yy * w)                       # y outer, x inner
    dbits = st["depths"].contiguous().view(torch.int32).to(torch.int64)[idx] & 0xFFFFFFFF
    keys = (tile << 32) | dbits
    ks, order = torch.sort(keys, stable=True)
    pl = idx[order]
    counts = torch.bincount(ks >> 32, minlength=gx * gy) if n else torch.zeros(gx * gy, dtype=torch.int64)
    end = torch.cumsum(counts, 0)
    ranges = torch.stack([end - counts, end], 1)
    ranges[counts == 0] = 0                                                   # untouched tiles stay (0, 0): the memset
    return dict(num_rendered=n, point_offsets=offs, keys_unsorted=keys, values_unsorted=idx, keys_sorted=ks,
                point_list=pl, ranges=ranges)


class _Lanes:
    """[tiles, 256] pixel lanes with the tiles ordered by list length (longest first): at step k the tiles that
    still have a k-th instance are the prefix [:active(k)]."""

    def __init__(self, ranges, W, H):
        gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
        n = (ranges[:, 1] - ranges[:, 0])
        self.order = torch.argsort(n, descending=True, stable=True)
        self.start = ranges[self.order, 0]
        self.n = n[self.order]
        self.K = int(self.n[0]) if self.n.numel() else 0
        # active(k) = number of tiles with n > k
        self.active = (self.n[None, :] > torch.arange(max(self.K, 1))[:, None]).sum(1)
        tx, ty = self.order % gx, self.order // gx
        l = torch.arange(TILE * TILE)
        self.px = (tx[:, None] * TILE + (l % TILE)[None, :])
        self.py = (ty[:, None] * TILE + (l // TILE)[None, :])
        self.inside = (self.px < W) & (self.py < H)
        self.pid = torch.where(self.inside, self.py * W + self.px, 0)
        self.pxf, self.pyf = self.px.to(torch.float32), self.py.to(torch.float32)
        self.W, self.H = W, H

    def scatter(self, lanes, fill=0.0):
        """[tiles, 256] lane values -> [H, W] image."""
        out = torch.full((self.H * self.W,), fill, dtype=lanes.dtype)
        out[self.pid[self.inside]] = lanes[self.inside]
        return out.view(self.H, self.W)

    def gather(self, img):
        """[H, W] image -> [tiles, 256] lanes (0 outside the image)."""
        return torch.where(self.inside, img.reshape(-1)[self.pid], torch.zeros((), dtype=img.dtype))


def _pair(L, a, ids, tm, m2d, no):
    """The per-(pixel, instance) front half of the inner loop (CR/forward.cu:355-386, CR/backward.cu:283-316) for the
    instances `ids` [a] of the first `a` tiles: everything both passes need."""
    t = tm[ids]
    Tu, Tv, Tw = t[:, None, 0:3], t[:, None, 3:6], t[:, None, 6:9]
    px, py = L.pxf[:a, :, None], L.pyf[:a, :, None]
    k = px * Tw - Tu
    l = py * Tw - Tv
    p = torch.cross(k, l, dim=2)
    pz = p[..., 2]
    safe = torch.where(pz == 0, torch.ones_like(pz), pz)
    sx, sy = p[..., 0] / safe, p[..., 1] / safe
    rho3d = sx * sx + sy * sy
    dx, dy = m2d[ids, 0, None] - L.pxf[:a], m2d[ids, 1, None] - L.pyf[:a]
    rho2d = FILTER_INV_SQUARE * (dx * dx + dy * dy)
    rho = torch.minimum(rho3d, rho2d)
    in3d = rho3d <= rho2d
    depth = torch.where(in3d, sx * Tw[..., 0] + sy * Tw[..., 1] + Tw[..., 2], Tw[..., 2].expand_as(sx))
    power = -0.5 * rho
    G = torch.exp(power)
    opa = no[ids, 3, None]
    alpha = torch.clamp(opa * G, max=0.99)
    ok = (pz != 0) & (depth >= NEAR_N) & (power <= 0) & (alpha >= 1.0 / 255.0)
    return dict(k=k, l=l, pz=safe, sx=sx, sy=sy, in3d=in3d, dx=dx, dy=dy, depth=depth, G=G, alpha=alpha, ok=ok,
                Tw=Tw, opa=opa)


def render_forward(W, H, ranges, point_list, means2D, features, transMat, normal_opacity, bg, use_sa=True):
    """renderCUDA forward (CR/forward.cu:258-467) for all tiles at once, one step per list position."""
    L = _Lanes(ranges, W, H)
    nt = ranges.shape[0]
    bg = _t(bg)
    zf = lambda: torch.zeros((nt, TILE * TILE))      # noqa: E731
    T = torch.ones((nt, TILE * TILE))
    C, N = [zf() for _ in range(3)], [zf() for _ in range(3)]
    D, D2, M1, M2, dist, med = zf(), zf(), zf(), zf(), zf(), zf()
    medc = torch.zeros((nt, TILE * TILE), dtype=torch.int64)
    last = torch.zeros((nt, TILE * TILE), dtype=torch.int64)
    done = ~L.inside
    for kk in range(L.K):
        a = int(L.active[kk])
        if bool(done[:a].all()):
            continue
        ids = point_list[L.start[:a] + kk]
        f = _pair(L, a, ids, transMat, means2D, normal_opacity)
        Ta, alpha, depth = T[:a], f["alpha"], f["depth"]
        cand = f["ok"] & ~done[:a]
        test_T = Ta * (1 - alpha)
        stop = cand & (test_T < 0.0001)
        done[:a] |= stop
        c = cand & ~stop
        w = alpha * Ta
        is_med = c & (Ta > 0.5)
        med[:a] = torch.where(is_med, depth, med[:a])
        medc[:a] = torch.where(is_med, kk + 1, medc[:a])
        cf = c.to(torch.float32)
        if use_sa:
            md = med[:a]
            exp_std = torch.clamp((D2[:a] - 2 * D[:a] * md) / (1 - Ta) + md * md, min=1e-7)
            conf = torch.exp(-((md - depth) ** 2) / (4 * exp_std))
            depth = torch.where(D[:a] > 0, conf * depth + (1 - conf) * md, depth)
            D[:a] += torch.where(c, depth * w, 0.0)
            D2[:a] += torch.where(c, depth * depth * w, 0.0)
        else:
            A = 1 - Ta
            m = FAR_N / (FAR_N - NEAR_N) * (1 - NEAR_N / depth)
            dist[:a] += torch.where(c, (m * m * A + M2[:a] - 2 * m * M1[:a]) * w, 0.0)
            D[:a] += torch.where(c, depth * w, 0.0)
            M1[:a] += torch.where(c, m * w, 0.0)
            M2[:a] += torch.where(c, m * m * w, 0.0)
        wc = w * cf
        for ch in range(3):
            N[ch][:a] += normal_opacity[ids, ch, None] * wc
            C[ch][:a] += features[ids, ch, None] * wc
        T[:a] = torch.where(c, test_T, Ta)
        last[:a] = torch.where(c, kk + 1, last[:a])
    var = D2 - 2 * med * D + med * med * (1 - T)
    out = dict(
        final_T=torch.stack([L.scatter(T), L.scatter(M1), L.scatter(M2)]),
        n_contrib=torch.stack([L.scatter(last), L.scatter(medc)]),
        color=torch.stack([L.scatter(C[ch] + T * bg[ch]) for ch in range(3)]),
        others=torch.stack([L.scatter(D), L.scatter(1 - T)] + [L.scatter(N[ch]) for ch in range(3)] +
                           [L.scatter(med), L.scatter(var if use_sa else dist)]),
        median_depth=L.scatter(med), depth_std=L.scatter(var))
    return out


def forward(means3D, opacities, viewmatrix, projmatrix, W, H, bg, scales=None, rotations=None, transMat=None,
            colors=None, shs=None, sh_degree=0, campos=None, scale_modifier=1.0, use_sa=True):
    """Whole forward pass (same dictionary keys as oracle/cpu.py: forward)."""
    if transMat is not None or shs is not None or colors is None or scales is None or rotations is None:
        raise NotImplementedError("torch transliteration: colors_precomp + scales/rotations only (what gaus-slam calls)")
    with torch.no_grad():
        st = preprocess(means3D, opacities, viewmatrix, projmatrix, W, H, scales, rotations, scale_modifier)
        st.update(binning(st, W, H))
        feats = _t(colors)
        st.update(render_forward(W, H, st["ranges"], st["point_list"], st["means2D"], feats, st["transMat"],
                                 st["normal_opacity"], bg, use_sa))
        st["features"] = feats
    return st


def render_backward(W, H, st, bg, dL_dcolor, dL_dothers, use_sa=True):
    """renderCUDA backward (CR/backward.cu:143-463): back to front over the list positions; per-instance gradients are
    the sums over the 256 pixels of the tile, scattered with index_add_ instead of atomics."""
    ranges, pl = st["ranges"], st["point_list"]
    tm, m2d, no, feats = st["transMat"], st["means2D"], st["normal_opacity"], st["features"]
    L = _Lanes(ranges, W, H)
    P = tm.shape[0]
    bg = _t(bg)
    gpix = [L.gather(_t(dL_dcolor)[ch]) for ch in range(3)]
    go = [L.gather(_t(dL_dothers)[ch]) for ch in range(7)]
    dL_ddepth, dL_daccum, dL_dn, dL_dmed, dL_dreg = go[0], go[1], go[2:5], go[5], go[6]
    T_final = L.gather(st["final_T"][0])
    final_D, final_D2 = L.gather(st["final_T"][1]), L.gather(st["final_T"][2])
    final_A = 1 - T_final
    last_c, med_c = L.gather(st["n_contrib"][0]), L.gather(st["n_contrib"][1])
    mm, mstd = L.gather(st["median_depth"]), L.gather(st["depth_std"])
    bg_dot = bg[0] * gpix[0] + bg[1] * gpix[1] + bg[2] * gpix[2]
    inv4std = 1.0 / (4 * torch.clamp(mstd / (1 - T_final), min=1e-7))
    nt = ranges.shape[0]
    zf = lambda: torch.zeros((nt, TILE * TILE))      # noqa: E731
    T = T_final.clone()
    acc_c, last_col = [zf() for _ in range(3)], [zf() for _ in range(3)]
    acc_n, last_n = [zf() for _ in range(3)], [zf() for _ in range(3)]
    acc_d, last_d, acc_a, last_alpha, last_dT = zf(), zf(), zf(), zf(), zf()
    g = dict(transMat=torch.zeros((P, 9)), means2D=torch.zeros((P, 3)), normal3D=torch.zeros((P, 3)),
             opacity=torch.zeros((P, 1)), colors=torch.zeros((P, 3)))
    for kk in range(L.K - 1, -1, -1):
        a = int(L.active[kk])
        ids = pl[L.start[:a] + kk]
        f = _pair(L, a, ids, tm, m2d, no)
        c = f["ok"] & (kk < last_c[:a]) & L.inside[:a]        # contributor (0-based kk) < last_contributor
        if not bool(c.any()):
            continue
        cf = c.to(torch.float32)
        alpha, G, c_d = f["alpha"], f["G"], f["depth"]
        Ta = torch.where(c, T[:a] / (1 - alpha), T[:a])
        T[:a] = Ta
        w = alpha * Ta
        la = last_alpha[:a]
        dL_dalpha = torch.zeros_like(w)
        gsum = {}
        for ch in range(3):
            col = feats[ids, ch, None]
            acc_c[ch][:a] = torch.where(c, la * last_col[ch][:a] + (1 - la) * acc_c[ch][:a], acc_c[ch][:a])
            last_col[ch][:a] = torch.where(c, col.expand_as(w), last_col[ch][:a])
            dL_dalpha = dL_dalpha + (col - acc_c[ch][:a]) * gpix[ch][:a]
            gsum[("colors", ch)] = (w * gpix[ch][:a] * cf).sum(1)
        conf = torch.ones_like(w)
        if use_sa:
            conf = torch.where(Ta < 0.5, torch.exp(-(c_d - mm[:a]) ** 2 * inv4std[:a]), conf)
            c_d = c_d * conf + mm[:a] * (1 - conf)
        dL_dz = torch.where((kk + 1) == med_c[:a], dL_dmed[:a], torch.zeros_like(w))
        if use_sa:
            dL_dweight = (c_d - mm[:a]) ** 2 * dL_dreg[:a]
            dL_dz = dL_dz + conf * 2.0 * w * (c_d - mm[:a]) * dL_dreg[:a]
        else:
            m_d = FAR_N / (FAR_N - NEAR_N) * (1 - NEAR_N / c_d)
            dmd_dd = (FAR_N * NEAR_N) / ((FAR_N - NEAR_N) * c_d * c_d)
            dL_dweight = (final_D2[:a] + m_d * m_d * final_A[:a] - 2 * m_d * final_D[:a]) * dL_dreg[:a]
            dL_dz = dL_dz + 2.0 * (Ta * alpha) * (m_d * final_A[:a] - final_D[:a]) * dL_dreg[:a] * dmd_dd
        dL_dalpha = dL_dalpha + dL_dweight - last_dT[:a]
        last_dT[:a] = torch.where(c, dL_dweight * alpha + (1 - alpha) * last_dT[:a], last_dT[:a])
        acc_d[:a] = torch.where(c, la * last_d[:a] + (1 - la) * acc_d[:a], acc_d[:a])
        last_d[:a] = torch.where(c, c_d, last_d[:a])
        dL_dalpha = dL_dalpha + (c_d - acc_d[:a]) * dL_ddepth[:a]
        acc_a[:a] = torch.where(c, la + (1 - la) * acc_a[:a], acc_a[:a])
        dL_dalpha = dL_dalpha + (1 - acc_a[:a]) * dL_daccum[:a]
        for ch in range(3):
            nrm = no[ids, ch, None]
            acc_n[ch][:a] = torch.where(c, la * last_n[ch][:a] + (1 - la) * acc_n[ch][:a], acc_n[ch][:a])
            last_n[ch][:a] = torch.where(c, nrm.expand_as(w), last_n[ch][:a])
            dL_dalpha = dL_dalpha + (nrm - acc_n[ch][:a]) * dL_dn[ch][:a]
            gsum[("normal3D", ch)] = (w * dL_dn[ch][:a] * cf).sum(1)
        dL_dalpha = dL_dalpha * Ta
        last_alpha[:a] = torch.where(c, alpha, la)
        dL_dalpha = dL_dalpha + (-T_final[:a] / (1 - alpha)) * bg_dot[:a]
        dL_dG = f["opa"] * dL_dalpha
        dL_dz = dL_dz + conf * alpha * Ta * dL_ddepth[:a]
        # geometric chain (CR/backward.cu:412-460)
        sx, sy, Tw, k, l = f["sx"], f["sy"], f["Tw"], f["k"], f["l"]
        m3 = (c & f["in3d"]).to(torch.float32)
        m2 = (c & ~f["in3d"]).to(torch.float32)
        dsx = (dL_dG * -G * sx + dL_dz * Tw[..., 0]) / f["pz"]
        dsy = (dL_dG * -G * sy + dL_dz * Tw[..., 1]) / f["pz"]
        dp = torch.stack([dsx, dsy, -(dsx * sx + dsy * sy)], 2)
        dk = torch.cross(l, dp, dim=2)
        dl = torch.cross(dp, k, dim=2)
        px, py = L.pxf[:a], L.pyf[:a]
        dT = torch.zeros((a, 9))
        for r in range(3):
            dT[:, r] = (-dk[..., r] * m3).sum(1)
            dT[:, 3 + r] = (-dl[..., r] * m3).sum(1)
        dT[:, 6] = ((px * dk[..., 0] + py * dl[..., 0] + dL_dz * sx) * m3).sum(1)
        dT[:, 7] = ((px * dk[..., 1] + py * dl[..., 1] + dL_dz * sy) * m3).sum(1)
        dT[:, 8] = ((px * dk[..., 2] + py * dl[..., 2] + dL_dz) * m3).sum(1) + (dL_dz * m2).sum(1)
        g["transMat"].index_add_(0, ids, dT)
        dm = torch.stack([(dL_dG * (-G * FILTER_INV_SQUARE * f["dx"]) * m2).sum(1),
                          (dL_dG * (-G * FILTER_INV_SQUARE * f["dy"]) * m2).sum(1), torch.zeros(a)], 1)
        g["means2D"].index_add_(0, ids, dm)
        g["opacity"].index_add_(0, ids, (G * dL_dalpha * cf).sum(1, keepdim=True))
        g["colors"].index_add_(0, ids, torch.stack([gsum[("colors", ch)] for ch in range(3)], 1))
        g["normal3D"].index_add_(0, ids, torch.stack([gsum[("normal3D", ch)] for ch in range(3)], 1))
    return g


def preprocess_backward(st, g, means3D, scales, rotations, viewmatrix, projmatrix, W, H):
    """preprocessCUDA backward + compute_transmat_aabb (CR/backward.cu:466-664) over all visible surfels at once.
    W, H here are the backward's own float re-derivation (int(focal * tan_fov * 2), :641-642), passed in."""
    p, sc, q = _t(means3D), _t(scales), _t(rotations)
    vm, pm = _t(viewmatrix).reshape(16), _t(projmatrix).reshape(16)
    P = p.shape[0]
    vis = st["radii"] > 0
    R0, R1, R2 = _quat_to_rot(q)
    L0, L1 = R0 * sc[:, 0:1], R1 * sc[:, 1:2]                 # scale_modifier ignored, like the reference (:504)
    one, zero = torch.ones((P, 1)), torch.zeros((P, 1))
    Mm = torch.stack([torch.cat([L0, zero], 1), torch.cat([L1, zero], 1), torch.cat([p, one], 1)], 1)     # [P,3,4]
    n2p = torch.tensor([[W / 2.0, 0.0, 0.0, (W - 1) / 2.0], [0.0, H / 2.0, 0.0, (H - 1) / 2.0], [0.0, 0.0, 0.0, 1.0]])
    A = pm.view(4, 4).t().contiguous()                        # A[k][r] = projmatrix[k + 4r]
    Pm = n2p @ A                                              # Pm[c][r] = sum_k A[k][r] * n2p[c][k]
    T = torch.einsum("prk,ck->pcr", Mm, Pm)                   # T[c] = (Tu | Tv | Tw)
    dT_raw = g["transMat"].view(P, 3, 3)
    dT = dT_raw.clone()
    gmx, gmy = g["means2D"][:, 0], g["means2D"][:, 1]
    Tu, Tv, Tw = T[:, 0], T[:, 1], T[:, 2]
    dist = Tw[:, 0] ** 2 + Tw[:, 1] ** 2 - Tw[:, 2] ** 2
    has = ((gmx != 0) | (gmy != 0)) & vis
    f = 1 / torch.where(has, dist, torch.ones_like(dist))
    sgn = torch.tensor([1.0, 1.0, -1.0])
    add = torch.zeros_like(dT)
    add[:, 0] = gmx[:, None] * (f[:, None] * Tw * sgn)
    add[:, 1] = gmy[:, None] * (f[:, None] * Tw * sgn)
    fac = f[:, None] - sgn * 2 * (f * f)[:, None] * Tw * Tw   # (f - 2 f^2 Tw^2), and (f + 2 f^2 Tw^2) for z
    add[:, 2] = (gmx[:, None] * Tu + gmy[:, None] * Tv) * fac * sgn
    dT = dT + add * has[:, None, None].to(torch.float32)
    dM = torch.einsum("kr,pkc->pcr", Pm, dT)                  # dM[c][r] = sum_k Pm[k][r] * dT[k][c]
    normal = _xform_vec(R2, vm)
    pv = _xform_point(p, vm)
    mult = torch.where(-(pv * normal).sum(1) > 0, 1.0, -1.0)
    dtn = _xform_vec_T(g["normal3D"], vm) * mult[:, None]
    dRS = [dM[:, 0, :3], dM[:, 1, :3], dtn]
    vR = [dRS[0] * sc[:, 0:1], dRS[1] * sc[:, 1:2], dRS[2]]
    s = torch.rsqrt((q * q).sum(1))
    w, x, y, z = (q * s[:, None]).unbind(1)
    v = lambda i, j: vR[i][:, j]      # noqa: E731
    dq = torch.stack([
        2 * (x * (v(1, 2) - v(2, 1)) + y * (v(2, 0) - v(0, 2)) + z * (v(0, 1) - v(1, 0))),
        2 * (-2 * x * (v(1, 1) + v(2, 2)) + y * (v(0, 1) + v(1, 0)) + z * (v(0, 2) + v(2, 0)) + w * (v(1, 2) - v(2, 1))),
        2 * (x * (v(0, 1) + v(1, 0)) - 2 * y * (v(0, 0) + v(2, 2)) + z * (v(1, 2) + v(2, 1)) + w * (v(2, 0) - v(0, 2))),
        2 * (x * (v(0, 2) + v(2, 0)) + y * (v(1, 2) + v(2, 1)) - 2 * z * (v(0, 0) + v(1, 1)) + w * (v(0, 1) - v(1, 0)))], 1)
    vf = vis[:, None].to(torch.float32)
    depth = st["transMat"][:, 8]
    out = dict(g)
    out["rotations"] = dq * vf
    out["scales"] = torch.stack([(dRS[0] * R0).sum(1), (dRS[1] * R1).sum(1)], 1) * vf
    out["means3D"] = dM[:, 2, :3] * vf
    # densification proxy (:660-663) from the RAW accumulated dL/dT
    out["means2D"] = torch.stack([dT_raw[:, 0, 2] * depth * 0.5 * W, dT_raw[:, 1, 2] * depth * 0.5 * H,
                                  torch.zeros(P)], 1) * vf
    out["sh"] = torch.zeros((P, 0, 3))
    return out


def backward(st, means3D, viewmatrix, projmatrix, W, H, tanfovx, tanfovy, bg, dL_dcolor, dL_dothers, scales=None,
             rotations=None, transMat=None, shs=None, sh_degree=0, campos=None, use_sa=True):
    """Whole backward pass from a forward() state (same keys as oracle/cpu.py: backward)."""
    if transMat is not None or shs is not None or scales is None or rotations is None:
        raise NotImplementedError("torch transliteration: colors_precomp + scales/rotations only (what gaus-slam calls)")
    with torch.no_grad():
        g = render_backward(W, H, st, bg, dL_dcolor, dL_dothers, use_sa)
        focal_x = np.float32(W) / (np.float32(2.0) * np.float32(tanfovx))
        focal_y = np.float32(H) / (np.float32(2.0) * np.float32(tanfovy))
        Wb = int(np.float32(focal_x) * np.float32(tanfovx) * np.float32(2))
        Hb = int(np.float32(focal_y) * np.float32(tanfovy) * np.float32(2))
        return preprocess_backward(st, g, means3D, scales, rotations, viewmatrix, projmatrix, Wb, Hb)
