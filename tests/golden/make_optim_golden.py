"""Golden vectors for the fused map optimiser step from the reference's OWN `Gaussians` class.

Run in the authoring container (needs /root/reference; CPU only):  python tests/golden/make_optim_golden.py

`scene/Gaussians.py` is loaded UNMODIFIED from the reference checkout as a single module (not through `scene/__init__`,
which pulls in the frame / renderer code and with it the CUDA rasterizer).  Its imports of packages that are absent here
and that the optimiser path never touches (pytorch3d, plyfile, simple_knn) are replaced by empty stub modules.  A
`Gaussians(config)` is constructed by the reference's constructor, given raw parameters as nn.Parameters (what
`create_from_pcd` ends with, scene/Gaussians.py:107-118), and then the reference's code runs:

    training_setup()                        scene/Gaussians.py:121-137   param groups + torch.optim.Adam(l, lr=0.0, eps=1e-15)
    get_render_params                       scene/Gaussians.py:327-347   sigmoid / exp (tiled when isotropic) / F.normalize
    loss.backward(); optimizer.step(); optimizer.zero_grad(set_to_none=True)     slam/Frontend.py:116-128

with `loss` replaced by a given gradient of the render parameters (torch.autograd.backward with grad_tensors), which is
exactly what the rasterizer's backward hands to autograd.

Output: tests/golden/optim_<case>.npz with in_<group> (raw parameters), in_grad (steps x [13 P], flat layout of
gs2d_b200/mapping.py), in_lr (5 doubles) and, after every step s = 1..S: ref_<group>_s, ref_m_<group>_s, ref_v_<group>_s,
ref_opacities_s, ref_scales_s, ref_rotations_s.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE = "/root/reference"
GROUPS = ("xyz", "opacity", "scaling", "rotation", "rgb")
LR_KEYS = ("xyz_lr", "opacity_lr", "scaling_lr", "rotation_lr", "rgb_lr")
FLAT = (("means3D", 3), ("opacities", 1), ("scales", 2), ("rotations", 4), ("colors_precomp", 3))
TRAINING_ARGS = dict(xyz_lr=0.0001, feature_lr=0.0025, rgb_lr=0.0025, rotation_lr=0.001, opacity_lr=0.05,
                     scaling_lr=0.001)                                           # configs/replica/config.py:125-132
STEPS = 4


def import_reference_gaussians():
    def stub(name, **attrs):
        m = types.ModuleType(name)
        sys.modules[name] = m
        for k, v in attrs.items():
            setattr(m, k, v)
        return m
    stub("pytorch3d")
    sys.modules["pytorch3d"].transforms = stub("pytorch3d.transforms")
    stub("plyfile", PlyData=object, PlyElement=object)
    stub("simple_knn")
    stub("simple_knn._C", distCUDA2=None)
    sys.path.insert(0, REFERENCE)
    spec = importlib.util.spec_from_file_location("reference_gaussians", os.path.join(REFERENCE, "scene", "Gaussians.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)

    class TorchOnCpu:
        """The module's `torch`, with the `device="cuda"` of zeros / tensor redirected to the CPU (the per-surfel
        statistics re-created after pruning / densification, scene/Gaussians.py:295-297, :527-529)."""
        def __getattr__(self, name):
            return getattr(torch, name)

        @staticmethod
        def zeros(*a, **kw):
            kw.pop("device", None)
            return torch.zeros(*a, **kw)

        @staticmethod
        def tensor(*a, **kw):
            kw.pop("device", None)
            return torch.tensor(*a, **kw)
    mod.torch = TorchOnCpu()
    return mod.Gaussians


def raw_parameters(P, scale_cols, seed):
    rng = np.random.default_rng(seed)
    return dict(xyz=rng.normal(0, 2, (P, 3)).astype(np.float32), opacity=rng.normal(0, 1.5, (P, 1)).astype(np.float32),
                scaling=rng.normal(-4, 0.5, (P, scale_cols)).astype(np.float32),
                rotation=(rng.normal(0, 1, (P, 4)) * rng.uniform(0.5, 2.0, (P, 1))).astype(np.float32),
                rgb=rng.uniform(0, 1, (P, 3)).astype(np.float32))


def gradients(P, seed):
    rng = np.random.default_rng(7919 * seed + 1)
    g = (rng.normal(0, 1, (STEPS, 13 * P)) * 10.0 ** rng.uniform(-6, -2, (STEPS, 13 * P))).astype(np.float32)
    g[rng.uniform(size=g.shape) < 0.3] = 0.0           # surfels a view does not see get exactly zero gradient
    return g


def run_reference(Gaussians, raw, grads, distribution):
    config = dict(render=dict(method="2dgs"), gaussians=dict(gaussian_distribution=distribution, training_args=TRAINING_ARGS))
    gs = Gaussians(config)
    for k in GROUPS:
        setattr(gs, "_" + k, torch.nn.Parameter(torch.tensor(raw[k]).requires_grad_(True)))
    gs.training_setup()
    P = raw["xyz"].shape[0]
    out = {}
    for s in range(1, STEPS + 1):
        rp = gs.get_render_params
        outs, gts, off = [], [], 0
        for name, w in FLAT:
            outs.append(rp[name])
            gts.append(torch.tensor(grads[s - 1, off:off + w * P].reshape(P, w)))
            off += w * P
        torch.autograd.backward(outs, gts)
        gs.optimizer.step()
        gs.optimizer.zero_grad(set_to_none=True)
        for k in GROUPS:
            p = getattr(gs, "_" + k)
            st = gs.optimizer.state[p]
            out[f"ref_{k}_{s}"] = p.detach().numpy().copy()
            out[f"ref_m_{k}_{s}"] = st["exp_avg"].numpy().copy()
            out[f"ref_v_{k}_{s}"] = st["exp_avg_sq"].numpy().copy()
        with torch.no_grad():
            rp = gs.get_render_params
            for k in ("opacities", "scales", "rotations"):
                out[f"ref_{k}_{s}"] = rp[k].numpy().copy()
    return out


def snapshot(gs, out, tag):
    for k in GROUPS:
        p = getattr(gs, "_" + k)
        st = gs.optimizer.state[p]
        out[f"ref_{k}_{tag}"] = p.detach().numpy().copy()
        out[f"ref_m_{k}_{tag}"] = st["exp_avg"].numpy().copy()
        out[f"ref_v_{k}_{tag}"] = st["exp_avg_sq"].numpy().copy()
    with torch.no_grad():
        rp = gs.get_render_params
        for k in ("opacities", "scales", "rotations"):
            out[f"ref_{k}_{tag}"] = rp[k].numpy().copy()


def one_step(gs, flat):
    P = gs._xyz.shape[0]
    rp = gs.get_render_params
    outs, gts, off = [], [], 0
    for name, w in FLAT:
        outs.append(rp[name])
        gts.append(torch.tensor(flat[off:off + w * P].reshape(P, w)))
        off += w * P
    torch.autograd.backward(outs, gts)
    gs.optimizer.step()
    gs.optimizer.zero_grad(set_to_none=True)


def run_surgery(Gaussians, P=300, n_new=77, seed=5):
    """step, step, remove_gaussians_from_mask, step, densification_postfix, step, reset_opacity, step -- the
    optimiser surgery of densification (scene/Gaussians.py:143-184, :280-297, :399-406, :419-433, :513-529), all of it
    the reference's own code."""
    rng = np.random.default_rng(seed)
    raw = raw_parameters(P, 2, seed)
    config = dict(render=dict(method="2dgs"), gaussians=dict(gaussian_distribution="anisotropic", training_args=TRAINING_ARGS))
    gs = Gaussians(config)
    for k in GROUPS:
        setattr(gs, "_" + k, torch.nn.Parameter(torch.tensor(raw[k]).requires_grad_(True)))
    gs.training_setup()
    out = {"in_" + k: raw[k] for k in GROUPS}
    out["in_lr"] = np.array([TRAINING_ARGS[k] for k in LR_KEYS], np.float64)

    def grad(tag):
        n = gs._xyz.shape[0]
        g = (rng.normal(0, 1, 13 * n) * 10.0 ** rng.uniform(-6, -2, 13 * n)).astype(np.float32)
        g[rng.uniform(size=g.shape) < 0.3] = 0.0
        out[f"in_grad_{tag}"] = g
        return g

    for tag in ("1", "2"):
        one_step(gs, grad(tag))
        snapshot(gs, out, tag)
    remove = rng.uniform(size=P) < 0.4
    remove[:3] = True                                   # the first surfels and a whole 256-block boundary region go
    remove[250:262] = True
    out["in_remove_mask"] = remove
    gs.remove_gaussians_from_mask(torch.tensor(remove))
    snapshot(gs, out, "pruned")
    one_step(gs, grad("3"))
    snapshot(gs, out, "3")
    new = raw_parameters(n_new, 2, seed + 1)
    for k in GROUPS:
        out["in_new_" + k] = new[k]
    t = {k: torch.tensor(v) for k, v in new.items()}
    gs.densification_postfix(t["xyz"], t["rgb"], t["opacity"], t["scaling"], t["rotation"])
    snapshot(gs, out, "extended")
    one_step(gs, grad("4"))
    snapshot(gs, out, "4")
    gs.reset_opacity()
    snapshot(gs, out, "reset")
    one_step(gs, grad("5"))
    snapshot(gs, out, "5")
    np.savez_compressed(os.path.join(HERE, "optim_surgery.npz"), **out)
    print("surgery", P, "->", int((~remove).sum()), "->", gs._xyz.shape[0], "surfels")


# name: (P, columns of the raw scaling, gaussian_distribution, seed).  P odd: the sections of the flat gradient are
# not 16-byte aligned (scalar path of the kernel); P % 4 == 0: 128-bit path.
CASES = dict(aniso=(257, 2, "anisotropic", 0), iso=(64, 1, "isotropic", 1), aligned=(512, 2, "anisotropic", 2))


def main():
    Gaussians = import_reference_gaussians()
    for name, (P, cols, dist, seed) in CASES.items():
        raw, grads = raw_parameters(P, cols, seed), gradients(P, seed)
        out = {"in_" + k: raw[k] for k in GROUPS}
        out["in_grad"] = grads
        out["in_lr"] = np.array([TRAINING_ARGS[k] for k in LR_KEYS], np.float64)
        out.update(run_reference(Gaussians, raw, grads, dist))
        np.savez_compressed(os.path.join(HERE, f"optim_{name}.npz"), **out)
        print(name, P, dist, "opacity moved by", float(np.abs(out[f"ref_opacity_{STEPS}"] - raw["opacity"]).max()))
    run_surgery(Gaussians)


if __name__ == "__main__":
    main()
