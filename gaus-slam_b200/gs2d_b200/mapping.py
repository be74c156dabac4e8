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
    """Keyframe rendered by `rank` at optimiser step `step`: ranks take consecutive views of the window (one pass
    over the window partitions it), and every completed pass shifts the assignment by one view -- so that a rank
    walks through ALL views even when the world size divides the window (8 ranks, 8 keyframes: without the shift
    every rank would render the same view for ever, its inputs would stay in L2 and the step time would be pinned to
    the heaviest view)."""
    return (step * world + rank + (step * world) // n_views) % n_views


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
    buffer).  With a buffer from symmetric_gradient_buffer and fused=True the gradients are ADDED into every rank's
    copy inside the NVSwitch by the backward pass itself (see FusedGradient)."""
    from gaus_2dgs_rasterization import _C
    _C.set_gradient_arena(flat)
    try:
        yield flat
    finally:
        _C.set_gradient_arena(None)


class FusedGradient:
    """The all-reduce of the map gradient fused into the backward pass ("fused with its collective").

        fg = FusedGradient(P, device)            # None-like (fg.ok False) without NVSwitch multicast
        every optimiser step:
            fg.begin()                           # zero the local copy, cross-GPU barrier: nobody adds into a dirty buffer
            ... forward + backward of this rank's view(s): the backward-preprocess epilogue multimem.red's the five
                map-parameter gradients into EVERY rank's copy (gs2d_backward_reduce)
            flat = fg.finish()                   # cross-GPU barrier: every rank's copy now holds the sum over ranks

    Several views per rank and step simply add up in the buffer (a bundle-adjustment window): one begin / finish pair
    per optimiser step.  Gradient tensors returned by the rasterizer alias the buffer and are valid after finish()."""

    def __init__(self, P: int, device, group=None):
        self.flat = symmetric_gradient_buffer(P, device, group)
        self.ok = self.flat is not None
        self.P = P
        if self.ok:
            self.hdl = _symmetric[self.flat.data_ptr()][0]
            self.side = torch.cuda.Stream(device=device)
            self.ready = torch.cuda.Event()
        self._active = False

    def begin(self):
        """Zero the local copy and run the first cross-GPU barrier on a SIDE stream (they hide under the forward pass);
        the backward pass waits for them through an event.  Call when everything that reads the previous gradient has
        been enqueued on the current stream."""
        from gaus_2dgs_rasterization import _C
        self.side.wait_stream(torch.cuda.current_stream(self.flat.device))
        with torch.cuda.stream(self.side):
            self.flat.zero_()
            self.hdl.barrier(channel=0)
            self.ready.record(self.side)
        _C.set_gradient_arena(self.flat, multicast_ptr=self.hdl.multicast_ptr, ready_event=self.ready)
        self._active = True

    def finish(self) -> torch.Tensor:
        from gaus_2dgs_rasterization import _C
        self.hdl.barrier(channel=1)
        _C.set_gradient_arena(None)
        self._active = False
        return self.flat


def unpack_grads(flat: torch.Tensor, P: int) -> dict:
    return {name: flat[a:b].view(P, PARAM_WIDTH[name]) for name, a, b in flat_sections(P)}


_symmetric = {}   # data_ptr of a flat buffer from symmetric_gradient_buffer -> (handle, numel, rank, world)


def symmetric_gradient_buffer(P: int, device, group=None):
    """Flat [13 * P] fp32 gradient buffer in NVLink symmetric memory bound to an NVSwitch multicast object
    (torch.distributed._symmetric_memory does the allocation and the binding).  allreduce_map_gradient then
    reduces it IN the switch with libgs2d's own kernel (csrc/collective.cu: multimem.ld_reduce / multimem.st)
    instead of calling NCCL.  Returns None when symmetric memory or multicast is unavailable (e.g. gloo, one
    rank, no NVSwitch): the caller falls back to an ordinary tensor and NCCL."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) < 2:
        return None
    if dist.get_backend(group) != "nccl" or (13 * P) % 4:
        return None
    try:
        import torch.distributed._symmetric_memory as symm_mem
        g = group if group is not None else dist.group.WORLD
        flat = symm_mem.empty(FLAT_WIDTH * P, dtype=torch.float32, device=device)
        hdl = symm_mem.rendezvous(flat, g.group_name)
        if not getattr(hdl, "multicast_ptr", 0):
            return None
    except Exception:      # no symmetric memory on this build / box
        return None
    flat.zero_()
    _symmetric[flat.data_ptr()] = (hdl, flat.numel(), hdl.rank, hdl.world_size)
    return flat


def allreduce_map_gradient(flat: torch.Tensor, group=None, average: bool = False) -> torch.Tensor:
    """Sum (or mean) of the map gradient over the ranks; a no-op without an initialised process group.
    Buffers from symmetric_gradient_buffer go through the in-switch kernel of libgs2d.so, bracketed by two
    cross-GPU barriers (all gradients written / all slices stored); anything else through the backend's
    all_reduce (NCCL on GPUs, gloo in the CPU tests)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        sym = _symmetric.get(flat.data_ptr()) if flat.is_cuda else None
        if sym is not None and sym[1] == flat.numel():
            import ctypes as C
            from . import _lib as L
            hdl, n, rank, world = sym
            hdl.barrier(channel=0)
            st = C.c_void_p(torch.cuda.current_stream(flat.device).cuda_stream)
            L.check(L.load().gs2d_allreduce_multimem(C.c_void_p(hdl.multicast_ptr), n, rank, world, 0, st),
                    "gs2d_allreduce_multimem")
            hdl.barrier(channel=1)
        else:
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        if average:
            flat /= dist.get_world_size(group)
    return flat
