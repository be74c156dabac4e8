"""Build libgs2d.so (hand-written CUDA for sm_100a + the C ABI of include/gs2d.h) in-tree.

No torch headers, no CUB/thrust, no CUTLASS: plain nvcc on the .cu files under csrc/.
The library lands in gaus-slam_b200/lib/libgs2d.so (git-ignored, shipped to the GPU box by gpurun).
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "lib", "libgs2d.so")
SOURCES = ["api.cu", "preprocess.cu", "binning.cu", "tile_sort.cu", "footprint.cu", "composite_fwd.cu", "composite_bwd.cu",
           "preprocess_bwd.cu", "knn.cu", "loss.cu", "collective.cu", "optim.cu", "pose.cu", "compact.cu", "no_alloc.cu"]
HEADERS = ["common.cuh", "math.cuh", "exact.cuh", "tma.cuh", "pair.cuh", "footprint.cuh", os.path.join("..", "..", "include", "gs2d.h")]
# Kernels whose results only have to meet a tolerance (backward gradients: 1e-3 relative; footprint masks: a
# superset) may use MUFU division / flush-to-zero; everything that decides bit-exact outputs keeps IEEE math.
FAST_MATH = {"composite_bwd.cu": ["-ftz=true", "-prec-div=false", "-prec-sqrt=false"],
             "footprint.cu": ["-ftz=true", "-prec-div=false", "-prec-sqrt=false"],
             "tile_sort.cu": ["-ftz=true", "-prec-div=false", "-prec-sqrt=false"]}
NVCC_FLAGS = ["-std=c++17", "-O3", "-lineinfo", "-Xcompiler", "-fPIC",
              "-gencode", "arch=compute_100a,code=sm_100a"]


def _stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, variant: str = "", defines: tuple = ()) -> str:
    """variant/defines build a tuning variant (lib/libgs2d_<variant>.so with extra -D flags) next to the default."""
    lib = LIB if not variant else os.path.join(HERE, "lib", f"libgs2d_{variant}.so")
    if not variant and not force and not _stale():
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    objdir = os.path.join(HERE, "lib", "obj" + ("_" + variant if variant else ""))
    os.makedirs(objdir, exist_ok=True)

    def one(src: str) -> str:
        obj = os.path.join(objdir, src + ".o")
        cmd = ["nvcc"] + NVCC_FLAGS + ["-D" + d for d in defines] + FAST_MATH.get(src, []) + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
            raise RuntimeError(f"nvcc failed on {src}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=8) as ex:
        objs = list(ex.map(one, SOURCES))
    r = subprocess.run(["nvcc", "-shared", "-o", lib] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"],
                       capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("link failed")
    return lib


if __name__ == "__main__":
    _variant = next((a.split("=", 1)[1] for a in sys.argv if a.startswith("--variant=")), "")
    _defs = tuple(a[2:] for a in sys.argv if a.startswith("-D"))
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, variant=_variant, defines=_defs))
