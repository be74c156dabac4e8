"""Drop-in for `simple_knn._C` (KNN/ext.cpp:15-17, KNN/spatial.cu:15-26): distCUDA2(points[P,3]) -> float[P],
the mean squared distance of every point to its 3 nearest other points.  Implemented by gs2d_dist2 of
libgs2d.so (csrc/knn.cu); torch only provides the memory and the stream.  No CPU path."""
from __future__ import annotations

import ctypes as C
import os
import sys

import torch

_PKG_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _PKG_ROOT not in sys.path:
    sys.path.insert(0, _PKG_ROOT)

from gs2d_b200 import _lib as L  # noqa: E402


def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    lib = L.load()
    if not points.is_cuda:
        raise RuntimeError("points must be a CUDA tensor")
    if points.ndim != 2 or points.shape[1] != 3:
        raise RuntimeError("points must have dimensions (num_points, 3)")
    pts = points.detach().to(torch.float32).contiguous()
    P = int(pts.shape[0])
    # the reference allocates with torch::full({P}, 0.0) (KNN/spatial.cu:20)
    means = torch.zeros((P,), dtype=torch.float32, device=pts.device)
    if P == 0:
        return means
    with torch.cuda.device(pts.device):
        scratch = torch.empty((lib.gs2d_dist2_scratch_bytes(P),), dtype=torch.uint8, device=pts.device)
        st = C.c_void_p(torch.cuda.current_stream(pts.device).cuda_stream)
        L.check(lib.gs2d_dist2(P, pts.data_ptr(), means.data_ptr(), scratch.data_ptr(), st), "gs2d_dist2")
    return means
