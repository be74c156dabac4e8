"""Drop-in for the reference's pybind module `gaus_2dgs_rasterization._C` (RAST/ext.cpp:15-19).

Same three entry points, same argument order, same return tuples as
RAST/rasterize_points.cu:39-61/:137, :140-165/:238 and :241-260 -- implemented over the C ABI of
libgs2d.so (include/gs2d.h) with raw device pointers taken from the torch tensors.  torch is only
the allocator / stream provider here; all compute is in the hand-written CUDA library.
"""
from __future__ import annotations

import ctypes as C

import torch

from gs2d_b200 import _lib as L

_pinned = {}


class _PairCount:
    """The pair-slot count of ONE forward pass (a by-product like num_rendered, needed to size the backward scratch):
    its own pinned int64 cell, written by a device-to-host copy enqueued right after the footprint scan, and an event
    recorded after the forward's last launch.  The cell goes back to a free list when the holder dies."""
    __slots__ = ("cell", "event")
    _free = []

    def __init__(self):
        self.cell, self.event = _PairCount._free.pop() if _PairCount._free else (torch.zeros(1, dtype=torch.int64).pin_memory(), torch.cuda.Event())
        C.c_int64.from_address(self.cell.data_ptr()).value = -1      # "not there yet"

    def __del__(self):
        try:
            _PairCount._free.append((self.cell, self.event))
        except Exception:       # interpreter shutdown
            pass

    def ptr(self):
        return self.cell.data_ptr()

    def wait(self):
        """The count, as soon as its copy has landed (long before the forward pass ends).  Bounded: once the forward's
        event has completed the copy must be there; a CUDA error of the forward surfaces through the event query."""
        cell = C.c_int64.from_address(self.cell.data_ptr())
        spins = 0
        while cell.value < 0:
            spins += 1
            if spins % 256 == 0 and self.event.query():
                if cell.value < 0:
                    raise RuntimeError("rasterize_gaussians_backward: the forward pass finished without delivering its "
                                       "pair count")
        return int(cell.value)


class _NumRendered(int):
    """num_rendered as the reference returns it (an int), carrying the forward's pair count for the backward pass:
    the autograd bridge keeps this object in ctx and hands it back as `R`."""
    pairs = None


# Optional destination of the five map-parameter gradients (gs2d_b200.mapping.gradient_arena): a flat fp32
# buffer [13 * P] laid out group by group (xyz | opacity | scale | rotation | rgb).  While set, the backward
# pass writes those gradients straight into it, so a view-sharded trainer all-reduces the buffer without a
# packing copy (the "write grads straight into the collective's buffer" item of SURVEY.md 8e).
_grad_arena = None
_grad_arena_multicast = 0     # NVLS multicast address of the arena (0 = an ordinary buffer: gradients are stored)


_grad_arena_ready = None      # event the backward pass waits for before it adds into a multicast arena


def set_gradient_arena(flat, multicast_ptr: int = 0, ready_event=None):
    """flat: [13 * P] fp32 buffer that receives the five map-parameter gradients (None switches the arena off).
    multicast_ptr: when the buffer lives in symmetric memory bound to an NVSwitch multicast object, its multicast
    address -- the backward pass then ADDS its gradients into every rank's copy inside the switch
    (gs2d_backward_reduce: the all-reduce is fused into the backward-preprocess epilogue) instead of storing them;
    the caller zeroes the buffer and runs the cross-GPU barriers (gs2d_b200.mapping.FusedGradient)."""
    global _grad_arena, _grad_arena_multicast, _grad_arena_ready
    _grad_arena = flat
    _grad_arena_multicast = int(multicast_ptr) if flat is not None else 0
    _grad_arena_ready = ready_event if flat is not None else None


def _arena_views(P, dev):
    a = _grad_arena
    if a is None or a.device != dev or a.dtype != torch.float32 or a.numel() != 13 * P or not a.is_contiguous():
        return None
    o = [0, 3 * P, 4 * P, 6 * P, 10 * P, 13 * P]
    return (a[o[0]:o[1]].view(P, 3), a[o[1]:o[2]].view(P, 1), a[o[2]:o[3]].view(P, 2), a[o[3]:o[4]].view(P, 4),
            a[o[4]:o[5]].view(P, 3))


def _stream_ptr(device) -> C.c_void_p:
    return C.c_void_p(L.raw_stream(torch, device))


def _f32(t: torch.Tensor, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32, got {t.dtype}")
    return t.contiguous()


def _ptr(t):
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _pinned_i64(device, stream_id) -> torch.Tensor:
    """One pinned num_rendered cell per device AND stream: two streams of one process (frontend + backend) do not
    share it."""
    key = (str(device), stream_id)
    if key not in _pinned:
        _pinned[key] = torch.zeros(1, dtype=torch.int64).pin_memory()
    return _pinned[key]


def _bucket(nbytes: int) -> int:
    """Round a big scratch size up to a 32 MiB bucket: consecutive views differ a little in num_rendered, and
    identical sizes let torch's caching allocator reuse the block instead of calling cudaMalloc."""
    g = 1 << 25
    return nbytes if nbytes < g else (nbytes + g - 1) // g * g


# Scatter arenas (one per device and stream): capacity follows the largest num_rendered seen, so that the scatter of
# the next forward pass can be launched BEFORE the host has read num_rendered back (the device is busy with it
# while the host wakes up from its wait, allocates the binning buffer and launches the rest).
_arenas = {}
_events = {}


def _arena(dev, stream_id):
    return _arenas.get((str(dev), stream_id))


def _grow_arena(dev, stream_id, n):
    cap = max(1 << 16, int(n * 1.25))
    cap = (cap + (1 << 18) - 1) >> 18 << 18
    t = torch.empty((cap,), dtype=torch.int64, device=dev)
    _arenas[(str(dev), stream_id)] = (t, cap)


def _settings(P, W, H, degree, M, use_sa, prefiltered, debug, tan_fovx, tan_fovy, scale_modifier):
    return L.Settings(P=P, image_width=W, image_height=H, sh_degree=int(degree), sh_coeffs=int(M),
                      use_sa=int(bool(use_sa)), prefiltered=int(bool(prefiltered)), debug=int(bool(debug)),
                      tanfovx=float(tan_fovx), tanfovy=float(tan_fovy), scale_modifier=float(scale_modifier),
                      reserved_=0.0)


def _inputs(background, means3D, sh, colors, opacity, scales, rotations, transMat, view, proj, campos):
    return L.Inputs(background=_ptr(background), means3D=_ptr(means3D), shs=_ptr(sh),
                    colors_precomp=_ptr(colors), opacities=_ptr(opacity), scales=_ptr(scales),
                    rotations=_ptr(rotations), transMat_precomp=_ptr(transMat), viewmatrix=_ptr(view),
                    projmatrix=_ptr(proj), campos=_ptr(campos))


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                        transMat_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height,
                        image_width, sh, degree, campos, use_sa, prefiltered, debug):
    """-> (num_rendered, out_color[3,H,W], out_others[7,H,W], radii[P] i32, geomBuffer, binningBuffer, imgBuffer)"""
    lib = L.load()
    if means3D.ndim != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    dev = means3D.device
    P, H, W = int(means3D.shape[0]), int(image_height), int(image_width)
    background = _f32(background, "background"); means3D = _f32(means3D, "means3D")
    colors = _f32(colors, "colors"); opacity = _f32(opacity, "opacity")
    scales = _f32(scales, "scales"); rotations = _f32(rotations, "rotations")
    transMat_precomp = _f32(transMat_precomp, "transMat_precomp")
    viewmatrix = _f32(viewmatrix, "viewmatrix"); projmatrix = _f32(projmatrix, "projmatrix")
    sh = _f32(sh, "sh"); campos = _f32(campos, "campos")
    M = int(sh.shape[1]) if sh.numel() != 0 and sh.ndim >= 2 else 0

    with torch.cuda.device(dev):
        st = _stream_ptr(dev)
        u8 = dict(dtype=torch.uint8, device=dev)
        out_color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
        out_others = torch.empty((7, H, W), dtype=torch.float32, device=dev)
        radii = torch.empty((P,), dtype=torch.int32, device=dev)
        geom = torch.empty((lib.gs2d_geom_bytes(P, M) if P else 0,), **u8)
        img = torch.empty((lib.gs2d_image_bytes(W, H),), **u8)
        s = _settings(P, W, H, degree, M, use_sa, prefiltered, debug, tan_fovx, tan_fovy, scale_modifier)
        i = _inputs(background, means3D, sh, colors, opacity, scales, rotations, transMat_precomp,
                    viewmatrix, projmatrix, campos)
        bufs = L.Buffers(geom=_ptr(geom), geom_bytes=geom.numel(), binning=None, binning_bytes=0,
                         image=_ptr(img), image_bytes=img.numel())
        sid = L.raw_stream(torch, dev)
        n_host = _pinned_i64(dev, sid)
        L.check(lib.gs2d_forward_preprocess(C.byref(s), C.byref(i), C.byref(bufs), _ptr(radii),
                                            n_host.data_ptr(), st), "gs2d_forward_preprocess")
        ev = _events.get((str(dev), sid))
        if ev is None:
            ev = _events[(str(dev), sid)] = torch.cuda.Event()
        ev.record()           # on the current stream of dev (this block runs under torch.cuda.device(dev))
        arena = _arena(dev, sid) if P and not debug else None
        if arena is not None:
            L.check(lib.gs2d_forward_scatter_early(C.byref(s), C.byref(bufs), arena[0].data_ptr(), arena[1], st),
                    "gs2d_forward_scatter_early")
        # Everything the host can prepare for the second half is prepared BEFORE it waits: the device is idle between the
        # scatter and the tile sort for as long as the host needs from its wake-up to that launch.
        binning = None
        if arena is not None:     # sized for the arena's capacity (>= this frame's count unless the arena is too small)
            binning = torch.empty((_bucket(lib.gs2d_binning_bytes(arena[1], W, H)),), **u8)
        pairs_holder = _PairCount() if P else None
        # the reference's one host wait (num_rendered): the pinned cell is polled (it is negative until the copy has
        # landed) -- a shorter wake-up than an event wait; the event bounds the wait if the forward's stream has faulted
        cell = C.c_int64.from_address(n_host.data_ptr())
        spins = 0
        while P and cell.value < 0:
            spins += 1
            if (spins & 1023) == 0 and ev.query():
                if cell.value < 0:
                    raise RuntimeError("rasterize_gaussians: the preprocess stage finished without delivering num_rendered")
        num_rendered = cell.value if P else 0
        scattered = None
        if arena is not None and num_rendered <= arena[1]:
            scattered = arena[0].data_ptr()
        else:
            if num_rendered and not debug:
                _grow_arena(dev, sid, num_rendered)         # for the next frames
            binning = torch.empty((_bucket(lib.gs2d_binning_bytes(num_rendered, W, H)) if num_rendered else 0,), **u8)
        bufs.binning = _ptr(binning)
        bufs.binning_bytes = binning.numel()
        num_rendered = _NumRendered(num_rendered)
        pairs_ptr = None
        if num_rendered:
            num_rendered.pairs = pairs_holder
            pairs_ptr = num_rendered.pairs.ptr()
        L.check(lib.gs2d_forward_render_scattered(C.byref(s), C.byref(i), C.byref(bufs), int(num_rendered), _ptr(out_color),
                                                  _ptr(out_others), pairs_ptr, scattered, st),
                "gs2d_forward_render")
        if num_rendered:
            num_rendered.pairs.event.record()
    return num_rendered, out_color, out_others, radii, geom, binning, img


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                 transMat_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                                 dL_dout_color, dL_dout_others, sh, degree, campos, geomBuffer, R,
                                 binningBuffer, imageBuffer, use_sa, debug):
    """-> (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dtransMat, dL_dsh, dL_dscales, dL_drotations)"""
    lib = L.load()
    dev = means3D.device
    P = int(means3D.shape[0])
    H, W = int(dL_dout_color.shape[1]), int(dL_dout_color.shape[2])
    background = _f32(background, "background"); means3D = _f32(means3D, "means3D")
    colors = _f32(colors, "colors"); scales = _f32(scales, "scales"); rotations = _f32(rotations, "rotations")
    transMat_precomp = _f32(transMat_precomp, "transMat_precomp")
    viewmatrix = _f32(viewmatrix, "viewmatrix"); projmatrix = _f32(projmatrix, "projmatrix")
    sh = _f32(sh, "sh"); campos = _f32(campos, "campos")
    dL_dout_color = _f32(dL_dout_color, "dL_dout_color"); dL_dout_others = _f32(dL_dout_others, "dL_dout_others")
    M = int(sh.shape[1]) if sh.numel() != 0 and sh.ndim >= 2 else 0

    f32 = dict(dtype=torch.float32, device=dev)
    arena = _arena_views(P, dev) if P else None
    if arena is not None:
        g_means3D, g_opacity, g_scales, g_rot, g_colors = arena
    else:
        g_means3D = torch.empty((P, 3), **f32); g_opacity = torch.empty((P, 1), **f32)
        g_scales = torch.empty((P, 2), **f32); g_rot = torch.empty((P, 4), **f32)
        g_colors = torch.empty((P, 3), **f32)
    g_means2D = torch.empty((P, 3), **f32)
    g_transMat = torch.empty((P, 9), **f32); g_sh = torch.empty((P, M, 3), **f32)
    if P == 0:
        return g_means2D, g_colors, g_opacity, g_means3D, g_transMat, g_sh, g_scales, g_rot

    num_pairs = 0
    if int(R) > 0:
        holder = getattr(R, "pairs", None)
        if holder is not None:
            num_pairs = holder.wait()
        else:
            # R is a plain int (a caller that did not keep the object the forward returned): ask the binning buffer
            with torch.cuda.device(dev):
                cell = torch.zeros(1, dtype=torch.int64).pin_memory()
                s0 = _settings(P, W, H, degree, 0, use_sa, False, debug, tan_fovx, tan_fovy, scale_modifier)
                b0 = L.Buffers(geom=None, geom_bytes=0, binning=_ptr(binningBuffer), binning_bytes=binningBuffer.numel(),
                               image=None, image_bytes=0)
                L.check(lib.gs2d_forward_pair_count(C.byref(s0), C.byref(b0), int(R), cell.data_ptr(), _stream_ptr(dev)),
                        "gs2d_forward_pair_count")
                torch.cuda.current_stream(dev).synchronize()
                num_pairs = int(cell[0])
    with torch.cuda.device(dev):
        st = _stream_ptr(dev)
        scratch = torch.empty((_bucket(lib.gs2d_backward_scratch_bytes(P, int(R), num_pairs, W, H)),),
                              dtype=torch.uint8, device=dev)
        s = _settings(P, W, H, degree, M, use_sa, False, debug, tan_fovx, tan_fovy, scale_modifier)
        # opacities are not an input of backward (the forward stored them in the splat records)
        i = _inputs(background, means3D, sh, colors, means3D, scales, rotations, transMat_precomp,
                    viewmatrix, projmatrix, campos)
        bufs = L.Buffers(geom=_ptr(geomBuffer), geom_bytes=geomBuffer.numel(), binning=_ptr(binningBuffer),
                         binning_bytes=binningBuffer.numel(), image=_ptr(imageBuffer),
                         image_bytes=imageBuffer.numel())
        grads = L.Grads(dL_dmeans2D=_ptr(g_means2D), dL_dcolors=_ptr(g_colors), dL_dopacity=_ptr(g_opacity),
                        dL_dmeans3D=_ptr(g_means3D), dL_dtransMat=_ptr(g_transMat), dL_dsh=_ptr(g_sh),
                        dL_dscales=_ptr(g_scales), dL_drotations=_ptr(g_rot))
        if arena is not None and _grad_arena_multicast:
            if _grad_arena_ready is not None:     # the buffer is zeroed on every rank (side stream + cross-GPU barrier)
                torch.cuda.current_stream(dev).wait_event(_grad_arena_ready)
            L.check(lib.gs2d_backward_reduce(C.byref(s), C.byref(i), C.byref(bufs), _ptr(radii), int(R), num_pairs,
                                             _ptr(dL_dout_color), _ptr(dL_dout_others), C.byref(grads),
                                             C.c_void_p(_grad_arena_multicast), _ptr(scratch), st), "gs2d_backward_reduce")
        else:
            L.check(lib.gs2d_backward(C.byref(s), C.byref(i), C.byref(bufs), _ptr(radii), int(R), num_pairs,
                                      _ptr(dL_dout_color), _ptr(dL_dout_others), C.byref(grads), _ptr(scratch), st),
                    "gs2d_backward")
    return g_means2D, g_colors, g_opacity, g_means3D, g_transMat, g_sh, g_scales, g_rot


def mark_visible(means3D, viewmatrix, projmatrix):
    """-> bool[P]: view-space z > 0.2 (RAST/rasterize_points.cu:241-260)."""
    lib = L.load()
    dev = means3D.device
    P = int(means3D.shape[0])
    means3D = _f32(means3D, "means3D"); viewmatrix = _f32(viewmatrix, "viewmatrix")
    projmatrix = _f32(projmatrix, "projmatrix")
    present = torch.zeros((P,), dtype=torch.bool, device=dev)
    if P:
        with torch.cuda.device(dev):
            L.check(lib.gs2d_mark_visible(P, _ptr(means3D), _ptr(viewmatrix), _ptr(projmatrix),
                                          present.data_ptr(), _stream_ptr(dev)), "gs2d_mark_visible")
    return present
