"""Fused map optimiser: the host side of gs2d_map_activate / gs2d_map_adam_step.

Mirrors how the reference drives its map parameters: `Gaussians.training_setup` builds
`torch.optim.Adam(groups, lr=0.0, eps=1e-15)` over the raw tensors `_xyz, _opacity, _scaling, _rotation, _rgb`
with the learning rates of `config['gaussians']['training_args']` (scene/Gaussians.py:124-137), and
`get_render_params` renders `sigmoid(_opacity)`, `exp(_scaling)`, `F.normalize(_rotation)` (:336-358).

    opt = MapOptimizer(xyz, opacity_logit, log_scaling, rotation, rgb, training_args)
    params = opt.render_params()            # means3D / opacities / scales / rotations / colors_precomp (leaves)
    ... forward, loss, backward inside mapping.gradient_arena(opt.flat_grad) (+ all-reduce) ...
    opt.step()                              # chain rule of the activations + Adam + next activations: ONE launch

CUDA only: there is no CPU path (the CPU restatement used by the tests lives outside the package).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib as L
from .mapping import FLAT_WIDTH

GROUPS = ("xyz", "opacity", "scaling", "rotation", "rgb")
LR_KEYS = ("xyz_lr", "opacity_lr", "scaling_lr", "rotation_lr", "rgb_lr")


class MapOptimizer:
    def __init__(self, xyz, opacity, scaling, rotation, rgb, training_args, betas=(0.9, 0.999), eps=1e-15,
                 flat_grad=None):
        raw = dict(xyz=xyz, opacity=opacity, scaling=scaling, rotation=rotation, rgb=rgb)
        P = int(xyz.shape[0])
        want = dict(xyz=[(P, 3)], opacity=[(P, 1)], scaling=[(P, 2), (P, 1)], rotation=[(P, 4)], rgb=[(P, 3)])
        for k, t in raw.items():
            if not t.is_cuda or t.dtype != torch.float32:
                raise RuntimeError(f"{k} must be a float32 CUDA tensor (there is no CPU path)")
            if tuple(t.shape) not in want[k]:
                raise RuntimeError(f"{k} has shape {tuple(t.shape)}, expected one of {want[k]}")
        self.P, self.dev = P, xyz.device
        self.scale_cols = int(scaling.shape[1])
        self.raw = {k: t.detach().contiguous() for k, t in raw.items()}
        self.exp_avg = {k: torch.zeros_like(t) for k, t in self.raw.items()}
        self.exp_avg_sq = {k: torch.zeros_like(t) for k, t in self.raw.items()}
        self.lr = [float(training_args[k]) for k in LR_KEYS]
        self.betas, self.eps, self.step_count = (float(betas[0]), float(betas[1])), float(eps), 0
        self.group_steps = [0] * len(GROUPS)
        if flat_grad is not None and (flat_grad.numel() != FLAT_WIDTH * P or not flat_grad.is_cuda):
            raise RuntimeError("flat_grad must be a CUDA tensor of 13 * P floats")
        self._state = L.MapState()
        self._resize(flat_grad)

    def _resize(self, flat_grad=None):
        """(Re)create what depends on P after the raw tensors changed: gradient buffer, activations, bindings."""
        P = self.P = int(self.raw["xyz"].shape[0])
        self.flat_grad = flat_grad if flat_grad is not None else torch.zeros(FLAT_WIDTH * P, device=self.dev)
        self.opacities = torch.empty((P, 1), device=self.dev)
        self.scales = torch.empty((P, 2), device=self.dev)
        self.rotations = torch.empty((P, 4), device=self.dev)
        self._bind(self._state, self.raw, self.exp_avg, self.exp_avg_sq)
        self.activate()

    @staticmethod
    def _bind(s, raw, exp_avg, exp_avg_sq):
        for k in GROUPS:
            setattr(s, k, raw[k].data_ptr())
        for i, k in enumerate(GROUPS):
            s.exp_avg[i] = exp_avg[k].data_ptr()
            s.exp_avg_sq[i] = exp_avg_sq[k].data_ptr()

    def _stream(self):
        return C.c_void_p(L.raw_stream(torch, self.dev))

    def activate(self):
        """Render parameters of the current raw parameters (get_render_params)."""
        with torch.cuda.device(self.dev):
            L.check(L.load().gs2d_map_activate(self.P, self.scale_cols, C.byref(self._state), self.opacities.data_ptr(),
                                               self.scales.data_ptr(), self.rotations.data_ptr(), self._stream()),
                    "gs2d_map_activate")

    def render_params(self, requires_grad=True):
        """The keyword arguments `render()` of render/render_2dgs.py passes on: fresh leaves over the optimiser's
        own buffers (no copies)."""
        def leaf(t):
            t = t.detach()
            t.requires_grad_(requires_grad)
            return t
        return dict(means3D=leaf(self.raw["xyz"]), opacities=leaf(self.opacities), scales=leaf(self.scales),
                    rotations=leaf(self.rotations), colors_precomp=leaf(self.raw["rgb"]))

    def step(self, flat_grad=None, skip=()):
        """One Adam step from the gradient w.r.t. the render parameters in the flat [13 * P] layout.
        `skip`: groups that have NO gradient in this step -- left untouched, moments included, as torch.optim.Adam
        leaves a parameter whose .grad is None (in Renderer_BA / Renderer_tracking the rotations are detached,
        render/__init__.py:36, :98: pass skip=("rotation",) there)."""
        for k in skip:
            if k not in GROUPS:
                raise KeyError(k)
        g = self.flat_grad if flat_grad is None else flat_grad
        if g.numel() != FLAT_WIDTH * self.P or not g.is_cuda or g.dtype != torch.float32:
            raise RuntimeError("flat_grad must be a float32 CUDA tensor of 13 * P elements")
        self.step_count += 1
        ap = L.AdamParams()
        for i, v in enumerate(self.lr):
            ap.lr[i] = -1.0 if GROUPS[i] in skip else v
            if GROUPS[i] not in skip:         # torch keeps one step counter per parameter
                self.group_steps[i] += 1
            ap.group_step[i] = self.group_steps[i]
        ap.beta1, ap.beta2, ap.eps, ap.step = self.betas[0], self.betas[1], self.eps, self.step_count
        with torch.cuda.device(self.dev):
            L.check(L.load().gs2d_map_adam_step(self.P, self.scale_cols, C.byref(self._state), g.data_ptr(), C.byref(ap),
                                                self.opacities.data_ptr(), self.scales.data_ptr(),
                                                self.rotations.data_ptr(), self._stream()), "gs2d_map_adam_step")

    # ---- densification: the optimiser surgery of scene/Gaussians.py:143-184, :419-433 ------------------------
    def prune(self, keep):
        """`prune_optimizer(mask)` (scene/Gaussians.py:143-160): keep the surfels where `keep` [P] (bool) is set, in
        order, with their Adam moments; the step counter stays.  One plan (count + scan) and ONE move kernel for
        all 15 arrays instead of 15 boolean-index launches.  Returns the number of surfels kept."""
        if keep.dtype != torch.bool or not keep.is_cuda or tuple(keep.shape) != (self.P,):
            raise RuntimeError("keep must be a bool CUDA tensor of P elements")
        keep = keep.contiguous()
        lib = L.load()
        with torch.cuda.device(self.dev):
            scratch = torch.empty(lib.gs2d_map_compact_scratch_bytes(self.P) // 4, dtype=torch.int32, device=self.dev)
            L.check(lib.gs2d_map_compact_plan(self.P, keep.data_ptr(), scratch.data_ptr(), self._stream()),
                    "gs2d_map_compact_plan")
            n = int(scratch[0].item())          # the host has to size the new tensors: one 4-byte read-back
            new = [{k: torch.empty((n, t.shape[1]), device=self.dev) for k, t in d.items()}
                   for d in (self.raw, self.exp_avg, self.exp_avg_sq)]
            if n:
                dst = L.MapState()
                self._bind(dst, *new)
                L.check(lib.gs2d_map_compact_move(self.P, self.scale_cols, keep.data_ptr(), C.byref(self._state),
                                                  C.byref(dst), scratch.data_ptr(), self._stream()),
                        "gs2d_map_compact_move")
        self.raw, self.exp_avg, self.exp_avg_sq = new
        self._resize()
        return n

    def extend(self, xyz, opacity, scaling, rotation, rgb):
        """`cat_tensors_to_optimizer` (scene/Gaussians.py:162-184): append new surfels (raw parameters) with zero
        moments; the step counter (hence the bias correction) stays that of the existing surfels."""
        ext = dict(xyz=xyz, opacity=opacity, scaling=scaling, rotation=rotation, rgb=rgb)
        n = int(xyz.shape[0])
        for k, t in ext.items():
            if not t.is_cuda or t.dtype != torch.float32 or tuple(t.shape) != (n, self.raw[k].shape[1]):
                raise RuntimeError(f"{k}: expected a float32 CUDA tensor of shape {(n, self.raw[k].shape[1])}")
        for k, t in ext.items():
            self.raw[k] = torch.cat((self.raw[k], t.detach()), dim=0)
            self.exp_avg[k] = torch.cat((self.exp_avg[k], torch.zeros_like(t)), dim=0)
            self.exp_avg_sq[k] = torch.cat((self.exp_avg_sq[k], torch.zeros_like(t)), dim=0)
        self._resize()
        return self.P

    def replace(self, name, tensor):
        """`replace_tensor_to_optimizer(tensor, name)` (scene/Gaussians.py:419-433; reset_opacity :399-406,
        transform_gaussians :408-417): new raw values for one group, its moments reset to zero."""
        if name not in GROUPS:
            raise KeyError(name)
        if not tensor.is_cuda or tensor.dtype != torch.float32 or tuple(tensor.shape) != tuple(self.raw[name].shape):
            raise RuntimeError(f"{name}: expected a float32 CUDA tensor of shape {tuple(self.raw[name].shape)}")
        self.raw[name].copy_(tensor.detach())
        self.exp_avg[name].zero_()
        self.exp_avg_sq[name].zero_()
        self.activate()
