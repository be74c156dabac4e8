"""View-sharded mapping / bundle-adjustment plumbing (new: the reference is single-GPU).

One process per GPU, surfel parameters replicated, the keyframe window split across ranks by view; every
rank runs forward + backward for its own views and the map gradient -- one flat fp32 buffer of 13 values per
surfel (xyz 3, opacity 1, scale 2, rotation 4, rgb 3: the five parameter groups of scene/Gaussians.py:124-137)
-- goes through ONE all-reduce per optimiser step.  Binning is per view, so per-view results are independent
of the number of ranks.  Works with any torch.distributed backend (NCCL on the GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

import contextlib

import torch
import torch.distributed as dist

PARAM_ORDER = ("means3D", "opacities", "scales", "rotations", "colors")
PARAM_WIDTH = {"means3D": 3, "opacities": 1, "scales": 2, "rotations": 4, "colors": 3}
FLAT_WIDTH = sum(PARAM_WIDTH.values())   # 13


def view_of(step: int, rank: int, world: int, n_views: int) -> int:
    """Keyframe rendered by `rank` at optimiser step `step`: ranks take consecutive views of the window."""
    return (step * world + rank) % n_views


def views_of_step(step: int, world: int, n_views: int):
    return [view_of(step, r, world, n_views) for r in range(world)]


def flat_sections(P: int):
    """(name, start, stop) of every parameter group inside the flat [13 * P] gradient buffer (structure of arrays:
    each group is contiguous, so kernels can write gradients straight into it and the all-reduce is one call)."""
    out, off = [], 0
    for name in PARAM_ORDER:
        n = PARAM_WIDTH[name] * P
        out.append((name, off, off + n))
        off += n
    return out


def grads_alias_flat(grads: dict, flat: torch.Tensor) -> bool:
    """True when every group gradient already lives in its section of `flat` (gradient_arena)."""
    P = flat.numel() // FLAT_WIDTH
    return all(grads[name] is not None and grads[name].data_ptr() == flat.data_ptr() + 4 * a
               for name, a, _ in flat_sections(P))


def pack_grads(grads: dict, flat: torch.Tensor) -> torch.Tensor:
    """The per-group gradients as one flat buffer: nothing to do when the backward pass wrote them there
    (gradient_arena), else one concatenating copy."""
    if not grads_alias_flat(grads, flat):
        torch.cat([grads[k].reshape(-1) for k in PARAM_ORDER], out=flat)
    return flat


@contextlib.contextmanager
def gradient_arena(flat: torch.Tensor):
    """While active, the rasterizer's backward pass writes dL/d(means3D, opacities, scales, rotations, colors)
    directly into the sections of `flat` ([13 * P] fp32 on the surfels' device): `param.grad` then aliases the
    buffer the all-reduce runs on, with no packing copy in between.  One backward pass per all-reduce (a second
    one would overwrite the first: accumulate over views by all-reducing per view or summing into another
    buffer)."""
    from gaus_2dgs_rasterization import _C
    _C.set_gradient_arena(flat)
    try:
        yield flat
    finally:
        _C.set_gradient_arena(None)


def unpack_grads(flat: torch.Tensor, P: int) -> dict:
    return {name: flat[a:b].view(P, PARAM_WIDTH[name]) for name, a, b in flat_sections(P)}


def allreduce_map_gradient(flat: torch.Tensor, group=None, average: bool = False) -> torch.Tensor:
    """Sum (or mean) of the map gradient over the ranks; a no-op without an initialised process group."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        if average:
            flat /= dist.get_world_size(group)
    return flat
