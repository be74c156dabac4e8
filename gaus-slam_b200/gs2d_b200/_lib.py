"""ctypes view of include/gs2d.h (libgs2d.so).

The library is plain C ABI: this module only mirrors its structs and prototypes.  There is no
CPU fallback anywhere in the product path: if the shared library is missing, or a call returns
a non-zero status, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# GS2D_LIBRARY points at another build of the same ABI (tuning variants); it is still the CUDA library, never a fallback.
LIB_PATH = os.environ.get("GS2D_LIBRARY") or os.path.join(os.path.dirname(_HERE), "lib", "libgs2d.so")

GS2D_OK = 0
GS2D_ERR_INVALID = -1
GS2D_ERR_CUDA = -2
GS2D_ERR_CAPACITY = -3
GS2D_ERR_UNSUPPORTED = -4


class Settings(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("image_width", C.c_int32), ("image_height", C.c_int32),
        ("sh_degree", C.c_int32), ("sh_coeffs", C.c_int32), ("use_sa", C.c_int32),
        ("prefiltered", C.c_int32), ("debug", C.c_int32),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("scale_modifier", C.c_float),
        ("reserved_", C.c_float),
    ]


class Inputs(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "background", "means3D", "shs", "colors_precomp", "opacities", "scales", "rotations",
        "transMat_precomp", "viewmatrix", "projmatrix", "campos")]


class Buffers(C.Structure):
    _fields_ = [("geom", C.c_void_p), ("geom_bytes", C.c_size_t),
                ("binning", C.c_void_p), ("binning_bytes", C.c_size_t),
                ("image", C.c_void_p), ("image_bytes", C.c_size_t)]


class Grads(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dtransMat", "dL_dsh",
        "dL_dscales", "dL_drotations")]


class StateView(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "depths", "splat", "rect", "cull", "tiles_touched", "point_offsets",
        "keys_unsorted", "values_unsorted", "keys_sorted", "point_list", "instances",
        "ranges", "final_T", "n_contrib", "median_depth", "depth_std")]


# name -> (restype, argtypes); also the list of symbols the header declares (checked by tests).
class LossParams(C.Structure):
    _fields_ = [("w_color", C.c_float), ("w_depth", C.c_float), ("w_dist", C.c_float), ("eps", C.c_float),
                ("depth_near", C.c_float), ("depth_far", C.c_float), ("tracking", C.c_int32), ("silmask_th", C.c_float),
                ("edge_thres", C.c_float)]


class MapState(C.Structure):
    _fields_ = [("xyz", C.c_void_p), ("opacity", C.c_void_p), ("scaling", C.c_void_p), ("rotation", C.c_void_p),
                ("rgb", C.c_void_p), ("exp_avg", C.c_void_p * 5), ("exp_avg_sq", C.c_void_p * 5)]


class PoseAdamParams(C.Structure):
    _fields_ = [("lr", C.c_double * 2), ("beta1", C.c_double), ("beta2", C.c_double), ("eps", C.c_double),
                ("step", C.c_int64 * 2)]


class AdamParams(C.Structure):
    _fields_ = [("lr", C.c_double * 5), ("beta1", C.c_double), ("beta2", C.c_double), ("eps", C.c_double),
                ("step", C.c_int64), ("group_step", C.c_int64 * 5)]


ABI_VERSION = 4      # GS2D_ABI_VERSION of the include/gs2d.h these structs mirror (checked when the library is loaded)

PROTOTYPES = {
    "gs2d_geom_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "gs2d_image_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "gs2d_binning_bytes": (C.c_size_t, [C.c_int64, C.c_int32, C.c_int32]),
    "gs2d_forward_preprocess": (C.c_int, [C.POINTER(Settings), C.POINTER(Inputs), C.POINTER(Buffers),
                                          C.c_void_p, C.c_void_p, C.c_void_p]),
    "gs2d_forward_render": (C.c_int, [C.POINTER(Settings), C.POINTER(Inputs), C.POINTER(Buffers),
                                      C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gs2d_forward_scatter_early": (C.c_int, [C.POINTER(Settings), C.POINTER(Buffers), C.c_void_p, C.c_int64, C.c_void_p]),
    "gs2d_forward_render_scattered": (C.c_int, [C.POINTER(Settings), C.POINTER(Inputs), C.POINTER(Buffers),
                                                C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gs2d_forward": (C.c_int, [C.POINTER(Settings), C.POINTER(Inputs), C.POINTER(Buffers), C.c_void_p,
                               C.POINTER(C.c_int64), C.c_void_p, C.c_void_p, C.c_void_p]),
    "gs2d_forward_pair_count": (C.c_int, [C.POINTER(Settings), C.POINTER(Buffers), C.c_int64, C.c_void_p, C.c_void_p]),
    "gs2d_backward_scratch_bytes": (C.c_size_t, [C.c_int32, C.c_int64, C.c_int64, C.c_int32, C.c_int32]),
    "gs2d_backward": (C.c_int, [C.POINTER(Settings), C.POINTER(Inputs), C.POINTER(Buffers), C.c_void_p,
                                C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.POINTER(Grads), C.c_void_p,
                                C.c_void_p]),
    "gs2d_backward_reduce": (C.c_int, [C.POINTER(Settings), C.POINTER(Inputs), C.POINTER(Buffers), C.c_void_p,
                                       C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.POINTER(Grads), C.c_void_p,
                                       C.c_void_p, C.c_void_p]),
    "gs2d_backward_overflowed": (C.c_int, [C.POINTER(Settings), C.POINTER(Buffers), C.c_int64, C.c_void_p, C.c_void_p]),
    "gs2d_mark_visible": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gs2d_dist2_scratch_bytes": (C.c_size_t, [C.c_int32]),
    "gs2d_dist2": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gs2d_state_view_get": (C.c_int, [C.POINTER(Settings), C.POINTER(Buffers), C.c_int64, C.POINTER(StateView)]),
    "gs2d_footprint_audit": (C.c_int, [C.POINTER(Settings), C.POINTER(Buffers), C.c_int64, C.POINTER(C.c_uint64),
                                       C.c_void_p]),
    "gs2d_sort_scratch_bytes": (C.c_size_t, [C.c_int64]),
    "gs2d_sort_pairs": (C.c_int, [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                  C.c_void_p, C.c_void_p]),
    "gs2d_sort_pairs_static": (C.c_int, [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                         C.c_void_p]),
    "gs2d_mapping_loss_scratch_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "gs2d_mapping_loss_forward": (C.c_int, [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_int32, C.POINTER(LossParams), C.c_void_p, C.c_void_p, C.c_void_p]),
    "gs2d_mapping_loss_backward": (C.c_int, [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_int32, C.POINTER(LossParams), C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_void_p]),
    "gs2d_map_activate": (C.c_int, [C.c_int32, C.c_int32, C.POINTER(MapState), C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p]),
    "gs2d_map_adam_step": (C.c_int, [C.c_int32, C.c_int32, C.POINTER(MapState), C.c_void_p, C.POINTER(AdamParams),
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gs2d_map_compact_scratch_bytes": (C.c_size_t, [C.c_int32]),
    "gs2d_map_compact_plan": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "gs2d_map_compact_move": (C.c_int, [C.c_int32, C.c_int32, C.c_void_p, C.POINTER(MapState), C.POINTER(MapState),
                                        C.c_void_p, C.c_void_p]),
    "gs2d_pose_transform_forward": (C.c_int, [C.c_int32] + [C.c_void_p] * 8),
    "gs2d_pose_transform_scratch_bytes": (C.c_size_t, []),
    "gs2d_pose_adam_step": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.POINTER(PoseAdamParams), C.c_void_p]),
    "gs2d_pose_transform_backward": (C.c_int, [C.c_int32] + [C.c_void_p] * 11),
    "gs2d_allreduce_multimem": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "gs2d_timing_enable": (C.c_int, [C.c_int]),
    "gs2d_num_kernels": (C.c_int, []),
    "gs2d_kernel_name": (C.c_char_p, [C.c_int]),
    "gs2d_timing_collect": (C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "gs2d_strerror": (C.c_char_p, [C.c_int]),
    "gs2d_last_cuda_error": (C.c_int, []),
    "gs2d_abi_version": (C.c_int, []),
    "gs2d_launch_count": (C.c_int64, []),
}

_lib = None


def load():
    """Load libgs2d.so (once).  Raises if it has not been built: there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"libgs2d.so not found at {LIB_PATH}: build it with `python gaus-slam_b200/build.py` "
            "(or __graft_entry__.build()); this package has no CPU or PyTorch fallback")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if lib.gs2d_abi_version() != ABI_VERSION:
        raise RuntimeError(f"{LIB_PATH} has ABI version {lib.gs2d_abi_version()}, this package mirrors version "
                           f"{ABI_VERSION} of include/gs2d.h: rebuild it (python gaus-slam_b200/build.py --force)")
    _lib = lib
    return lib


class Gs2dError(RuntimeError):
    pass


def check(status: int, what: str) -> None:
    if status != GS2D_OK:
        lib = load()
        msg = lib.gs2d_strerror(status).decode()
        extra = ""
        if status == GS2D_ERR_CUDA:
            extra = f" (cudaError_t {lib.gs2d_last_cuda_error()})"
        raise Gs2dError(f"{what}: {msg}{extra}")


def raw_stream(torch, device) -> int:
    """The cudaStream_t of torch's current stream on `device` as an integer -- what
    torch.cuda.current_stream(device).cuda_stream returns, without building a Stream object (~3 us a call, five calls
    per mapping iteration on the path the GPU waits for)."""
    idx = device.index
    if idx is None:
        idx = torch.cuda.current_device()
    return int(torch._C._cuda_getCurrentRawStream(idx))
