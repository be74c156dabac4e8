"""Device timing of the "next"-row kernels (f1 pose transform, f3 map optimiser step) against the HBM
roofline, with the torch code they replace timed beside them.  CUDA events, warm-up, inputs larger than L2 (or
rotated over several sets).  Prints one JSON line; run it under ncu (--set full -k regex:map_step|pose_transform) for
the DRAM counters.  usage: python scripts/next_rows_bench.py [--P 500000] [--iters 50]"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gaus-slam_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, ROOT)


def timed(fn, iters, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3       # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--P", type=int, default=500_000)
    ap.add_argument("--iters", type=int, default=50)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    from gs2d_b200 import optim, tracking, scene
    import bench_loops
    import pose_ref
    peak = 6550.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    P = a.P
    sc = scene.room_scene(P, scene.REPLICA, seed=0)
    raw = bench_loops.raw_from_scene(sc, torch, dev)
    out = {"P": P, "hbm_peak_gbs": peak}

    # ---- f3: one optimiser step (392 B per surfel)
    opt = optim.MapOptimizer(*[raw[k] for k in bench_loops.GROUPS], bench_loops.TRAINING_ARGS)
    opt.flat_grad.normal_(0, 1e-4)
    us = timed(lambda: opt.step(), a.iters)
    out["map_adam_step"] = {"us": us, "alg_bytes": 392 * P, "gbs": 392 * P / us / 1e3, "frac_hbm": 392 * P / us / 1e3 / peak}
    us = timed(lambda: opt.activate(), a.iters)
    out["map_activate"] = {"us": us, "alg_bytes": 56 * P, "gbs": 56 * P / us / 1e3, "frac_hbm": 56 * P / us / 1e3 / peak}
    # what it replaces: torch activations + autograd of the activations + torch.optim.Adam (foreach)
    p = {k: torch.nn.Parameter(v.clone()) for k, v in raw.items()}
    lrk = dict(xyz="xyz_lr", opacity="opacity_lr", scaling="scaling_lr", rotation="rotation_lr", rgb="rgb_lr")
    adam = torch.optim.Adam([{"params": [p[k]], "lr": bench_loops.TRAINING_ARGS[lrk[k]]} for k in bench_loops.GROUPS], lr=0.0, eps=1e-15)
    g = {k: torch.randn_like(v) * 1e-4 for k, v in raw.items()}

    def torch_step():
        adam.zero_grad(set_to_none=True)
        rp = [p["xyz"], torch.sigmoid(p["opacity"]), torch.exp(p["scaling"]), torch.nn.functional.normalize(p["rotation"]), p["rgb"]]
        torch.autograd.backward(rp, [g["xyz"], g["opacity"], g["scaling"], g["rotation"], g["rgb"]])
        adam.step()
    out["torch_activations_plus_adam"] = {"us": timed(torch_step, a.iters)}

    # ---- f1: pose transform, 56 B per surfel forward, 24 B backward (+ 12 B with dL/dxyz); 8 sets > L2 at P = 500k
    sets = []
    for v in range(8):
        x = raw["xyz"].clone() + 0.001 * v
        r = torch.tensor(sc["rotations"], device=dev)
        sets.append((x, r, torch.randn_like(x) * 1e-4))
    q = torch.nn.Parameter(torch.tensor([0.9995, 0.01, -0.02, 0.015], device=dev))
    t = torch.nn.Parameter(torch.tensor([0.03, -0.02, 0.05], device=dev))
    it = {"i": 0}

    def ours(backward):
        def f():
            x, r, gg = sets[it["i"] % 8]; it["i"] += 1
            m, rr = tracking.pose_transform(x, r, cam_rot=q, cam_trans=t)
            if backward:
                q.grad = t.grad = None
                m.backward(gg)
        return f

    def ref(backward):
        def f():
            x, r, gg = sets[it["i"] % 8]; it["i"] += 1
            w2c = pose_ref.get_transform_matrix(q, t)
            m, rr = pose_ref.pose_transform_torch(x, r, w2c)
            if backward:
                q.grad = t.grad = None
                m.backward(gg)
        return f
    uf = timed(ours(False), a.iters)
    ufb = timed(ours(True), a.iters)
    out["pose_transform_forward"] = {"us": uf, "alg_bytes": 56 * P, "gbs": 56 * P / uf / 1e3, "frac_hbm": 56 * P / uf / 1e3 / peak,
                                     "note": "python + launch overhead included (one launch per call)"}
    out["pose_transform_fwd_bwd"] = {"us": ufb, "alg_bytes": 80 * P}
    out["torch_pose_forward"] = {"us": timed(ref(False), a.iters)}
    out["torch_pose_fwd_bwd"] = {"us": timed(ref(True), a.iters)}
    # per-kernel device time from the library's own event hooks
    from gs2d_b200 import _lib
    import ctypes as C
    lib = _lib.load()
    lib.gs2d_timing_enable(1)
    for _ in range(a.iters):
        ours(True)()
        opt.step()
    torch.cuda.synchronize()
    lib.gs2d_timing_enable(0)
    nk = lib.gs2d_num_kernels()
    tot = (C.c_double * nk)(); cnt = (C.c_int64 * nk)()
    lib.gs2d_timing_collect(tot, cnt)
    out["kernel_us"] = {lib.gs2d_kernel_name(i).decode(): {"avg_us": 1e3 * tot[i] / cnt[i], "launches": cnt[i]} for i in range(nk) if cnt[i]}
    ks = out["kernel_us"].get("map_adam_step")
    if ks:
        ks["gbs"] = 392 * P / ks["avg_us"] / 1e3
        ks["frac_hbm"] = ks["gbs"] / peak
    kp = out["kernel_us"].get("pose_transform")
    if kp:      # forward (56 B) and backward (24 B) launches averaged together
        kp["gbs"] = 40 * P / kp["avg_us"] / 1e3
        kp["frac_hbm"] = kp["gbs"] / peak
    # ---- f3, second half: prune 30 % of the map (plan + move of 15 arrays) vs 15 boolean-index launches in torch
    try:
        keep = torch.rand(P, device=dev) < 0.7
        kept = int(keep.sum())
        state = {k: (opt.raw[k].clone(), opt.exp_avg[k].clone(), opt.exp_avg_sq[k].clone()) for k in bench_loops.GROUPS}

        def ours_prune():
            for k in bench_loops.GROUPS:
                opt.raw[k], opt.exp_avg[k], opt.exp_avg_sq[k] = (t.clone() for t in state[k])
            opt._resize()
            opt.prune(keep)

        def torch_prune():
            for k in bench_loops.GROUPS:
                _ = [t[keep] for t in state[k]]

        def restore_only():
            for k in bench_loops.GROUPS:
                opt.raw[k], opt.exp_avg[k], opt.exp_avg_sq[k] = (t.clone() for t in state[k])
            opt._resize()
        n = max(3, a.iters // 4)
        t_ours, t_restore, t_torch = timed(ours_prune, n), timed(restore_only, n), timed(torch_prune, n)
        alg = 2 * P + 312 * kept
        out["map_compact_prune"] = {"us": t_ours - t_restore, "kept": kept, "alg_bytes": alg,
                                    "gbs": alg / max(t_ours - t_restore, 1e-3) / 1e3,
                                    "note": "MapOptimizer.prune incl. allocation of the new tensors, the 4-byte "
                                            "read-back and re-activation; restore cost subtracted"}
        out["torch_boolean_index_x15"] = {"us": t_torch}
    except Exception as e:      # first GPU run of the compaction kernels: report, do not lose the other numbers
        out["map_compact_prune"] = {"error": f"{type(e).__name__}: {e}"}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
