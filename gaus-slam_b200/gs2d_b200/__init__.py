"""gs2d_b200 -- host-side support for the B200-native 2D-Gaussian-surfel rasterizer.

  _lib   : ctypes view of include/gs2d.h (libgs2d.so)
  scene  : deterministic synthetic scenes (tests, bench)
  mapping: view-sharded multi-GPU mapping plumbing (flat gradient buffer + all-reduce)

The reference-facing API lives next to this package: `gaus_2dgs_rasterization` and `simple_knn`.
"""
from . import _lib, scene  # noqa: F401
