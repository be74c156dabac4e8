"""TEST INFRASTRUCTURE ONLY (never imported by the product): the reference's map optimiser step, restated.

`MapTorch` is the reference's own machinery run on the CPU: raw parameters as nn.Parameters, the activations of
`Gaussians.get_render_params` (scene/Gaussians.py:17-21, :336-358: sigmoid / exp / F.normalize, exp tiled to two
columns when isotropic) and `torch.optim.Adam(groups, lr=0.0, eps=1e-15)` with the per-group learning rates of
`training_setup` (:124-137).  torch.optim.Adam IS the reference's optimiser, so this pins the parity of the fused
kernel against outputs of the reference's implementation run here.  `map_step_numpy` is a float64 restatement of
the same step with hand-written activation gradients (checks the derivation the CUDA kernel uses).

Pinned: tests/golden/optim_*.npz are the states the reference's own `Gaussians` class (scene/Gaussians.py, imported
unmodified by tests/golden/make_optim_golden.py) reaches over 4 steps; MapTorch reproduces them bit for bit
(tests/test_optim.py).
"""
import numpy as np

GROUPS = ("xyz", "opacity", "scaling", "rotation", "rgb")
LR_KEYS = ("xyz_lr", "opacity_lr", "scaling_lr", "rotation_lr", "rgb_lr")
# flat [13 * P] gradient layout of the render parameters (gs2d_b200/mapping.py)
FLAT = (("means3D", 3), ("opacities", 1), ("scales", 2), ("rotations", 4), ("colors", 3))


def split_flat(flat, P):
    out, off = {}, 0
    for name, w in FLAT:
        out[name] = flat[off:off + w * P].reshape(P, w)
        off += w * P
    return out


class MapTorch:
    """The reference's parameter handling, on whatever device the tensors live on (CPU in the tests here)."""

    def __init__(self, raw, training_args, betas=(0.9, 0.999), eps=1e-15):
        import torch
        self.torch = torch
        self.p = {k: torch.nn.Parameter(torch.as_tensor(raw[k]).clone().requires_grad_(True)) for k in GROUPS}
        groups = [{"params": [self.p[k]], "lr": training_args[lk], "name": k} for k, lk in zip(GROUPS, LR_KEYS)]
        self.optimizer = torch.optim.Adam(groups, lr=0.0, eps=eps, betas=betas)      # scene/Gaussians.py:137

    def render_params(self):
        torch = self.torch
        scales = torch.exp(self.p["scaling"])
        if scales.shape[1] == 1:
            scales = torch.tile(scales, (1, 2))                                      # :340-341
        return dict(means3D=self.p["xyz"], opacities=torch.sigmoid(self.p["opacity"]), scales=scales,
                    rotations=torch.nn.functional.normalize(self.p["rotation"]), colors=self.p["rgb"])

    def step(self, flat_grad, detached=()):
        """Back-propagate a given gradient of the render parameters and step (what loss.backward() +
        optimizer.step() do in Frontend.mapping, slam/Frontend.py:116-128).  `detached`: render parameters that
        are detached before rendering (rotations in Renderer_BA, render/__init__.py:98): their raw parameter gets
        no gradient and torch.optim.Adam skips it."""
        torch = self.torch
        self.optimizer.zero_grad(set_to_none=True)
        rp = self.render_params()
        g = split_flat(torch.as_tensor(flat_grad), self.p["xyz"].shape[0])
        keys = [k for k, _ in FLAT if k not in detached]
        torch.autograd.backward([rp[k] for k in keys], [g[k] for k in keys])
        self.optimizer.step()

    # ---- densification surgery, restated from scene/Gaussians.py (pinned by tests/golden/optim_surgery.npz) ----
    def _swap(self, k, new_param, exp_avg=None, exp_avg_sq=None):
        """Replace group k's Parameter, carrying the optimiser state over (its "step" included)."""
        torch = self.torch
        group = next(g for g in self.optimizer.param_groups if g["name"] == k)
        stored = self.optimizer.state.get(group["params"][0], None)
        if stored is not None:
            stored["exp_avg"], stored["exp_avg_sq"] = exp_avg(stored["exp_avg"]), exp_avg_sq(stored["exp_avg_sq"])
            del self.optimizer.state[group["params"][0]]
        group["params"][0] = torch.nn.Parameter(new_param.requires_grad_(True))
        if stored is not None:
            self.optimizer.state[group["params"][0]] = stored
        self.p[k] = group["params"][0]

    def prune(self, keep):
        """prune_optimizer(mask) (:143-160): parameters and both moments indexed with the same boolean mask."""
        keep = self.torch.as_tensor(keep)
        for k in GROUPS:
            self._swap(k, self.p[k].detach()[keep], lambda m: m[keep], lambda v: v[keep])

    def extend(self, new):
        """cat_tensors_to_optimizer (:162-184): new rows appended, their moments zero."""
        torch = self.torch
        for k in GROUPS:
            ext = torch.as_tensor(new[k])
            self._swap(k, torch.cat((self.p[k].detach(), ext), dim=0),
                       lambda m: torch.cat((m, torch.zeros_like(ext)), dim=0),
                       lambda v: torch.cat((v, torch.zeros_like(ext)), dim=0))

    def replace(self, k, tensor):
        """replace_tensor_to_optimizer (:419-433): new values, moments reset to zero."""
        torch = self.torch
        tensor = torch.as_tensor(tensor).clone()
        self._swap(k, tensor, lambda m: torch.zeros_like(tensor), lambda v: torch.zeros_like(tensor))

    def reset_opacity_values(self):
        """What reset_opacity (:399-406) hands to replace_tensor_to_optimizer: the ACTIVATED opacities, min'ed with
        logit(0.01) -- which is always logit(0.01), since sigmoid > 0 > logit(0.01); kept as the reference has it."""
        torch = self.torch
        new = torch.sigmoid(self.p["opacity"]).detach()
        cap = 0.01 * torch.ones_like(new)
        return torch.min(new, torch.log(cap / (1 - cap)))

    def state(self):
        out = {k: self.p[k].detach().cpu().numpy() for k in GROUPS}
        for k in GROUPS:
            st = self.optimizer.state[self.p[k]]
            if "exp_avg" not in st:          # a parameter that never had a gradient has no state yet
                out["m_" + k] = out["v_" + k] = np.zeros_like(out[k])
                continue
            out["m_" + k] = st["exp_avg"].cpu().numpy()
            out["v_" + k] = st["exp_avg_sq"].cpu().numpy()
        with self.torch.no_grad():
            rp = self.render_params()
        for k in ("opacities", "scales", "rotations"):
            out[k] = rp[k].detach().cpu().numpy()
        return out


def activate_numpy(raw):
    q = raw["rotation"].astype(np.float64)
    n = np.maximum(np.sqrt((q * q).sum(1, keepdims=True)), 1e-12)
    s = np.exp(raw["scaling"].astype(np.float64))
    if s.shape[1] == 1:
        s = np.tile(s, (1, 2))
    return dict(opacities=1.0 / (1.0 + np.exp(-raw["opacity"].astype(np.float64))), scales=s, rotations=q / n)


def map_step_numpy(raw, m, v, flat_grad, lr, step, betas=(0.9, 0.999), eps=1e-15):
    """float64.  raw / m / v: dicts over GROUPS; returns the updated (raw, m, v) and the new activations."""
    P = raw["xyz"].shape[0]
    g = {k: x.astype(np.float64) for k, x in split_flat(np.asarray(flat_grad), P).items()}
    x = {k: raw[k].astype(np.float64) for k in GROUPS}
    act = activate_numpy(raw)
    graw = {"xyz": g["means3D"], "rgb": g["colors"]}
    o = act["opacities"]
    graw["opacity"] = g["opacities"] * (1.0 - o) * o
    gs = g["scales"] if x["scaling"].shape[1] == 2 else g["scales"].sum(1, keepdims=True)
    graw["scaling"] = gs * np.exp(x["scaling"])
    qh = act["rotations"]
    n = np.maximum(np.sqrt((x["rotation"] ** 2).sum(1, keepdims=True)), 1e-12)
    graw["rotation"] = (g["rotations"] - qh * (qh * g["rotations"]).sum(1, keepdims=True)) / n
    b1, b2 = betas
    bc1, bc2 = 1.0 - b1 ** step, 1.0 - b2 ** step
    nr, nm, nv = {}, {}, {}
    for k, l in zip(GROUPS, lr):
        nm[k] = m[k].astype(np.float64) + (graw[k] - m[k]) * (1.0 - b1)
        nv[k] = v[k].astype(np.float64) * b2 + (1.0 - b2) * graw[k] * graw[k]
        nr[k] = x[k] - (l / bc1) * nm[k] / (np.sqrt(nv[k]) / np.sqrt(bc2) + eps)
    return nr, nm, nv, activate_numpy(nr)
