"""oracle/build.py -- builds the two checkers.  TEST INFRASTRUCTURE, NOT PRODUCT CODE.

1. ``oracle/_build/libgs2d_oracle.so``: the plain-C CPU restatement (oracle/gs2d_oracle.c),
   compiled with gcc (+OpenMP).  Always built.
2. ``oracle/_ref/libgs2d_ref.so``: the UNMODIFIED reference CUDA rasterizer + simple-knn,
   compiled for sm_100a from the sources where they lie under /root/reference, plus the
   C-ABI shim oracle/ref_shim.cu.  Built only where /root/reference exists (the authoring
   container); the GPU box uses the prebuilt file that travels with the snapshot.
   Nothing from /root/reference is copied into the repository; outputs go to oracle/_ref/
   (git-ignored) only.  We do not run the reference's own build system (setup.py / CMake):
   the four .cu files compile directly with nvcc.  Two headers are force-included because
   the reference relies on transitive includes that gcc 13 no longer provides
   (CR/rasterizer_impl.h:24 uses std::uintptr_t; KNN/simple_knn.cu uses FLT_MAX).
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
RAST = os.path.join(REF, "submodules", "gaus_2dgs_rasterization")
CR = os.path.join(RAST, "cuda_rasterizer")
KNN = os.path.join(REF, "submodules", "simple-knn")
GENCODE = ["-gencode", "arch=compute_100a,code=sm_100a"]

REF_SO = os.path.join(HERE, "_ref", "libgs2d_ref.so")
ORACLE_SO = os.path.join(HERE, "_build", "libgs2d_oracle.so")


def _newer(target: str, sources: list[str]) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.exists(s) and os.path.getmtime(s) <= t for s in sources)


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        raise RuntimeError("command failed: " + cmd[0])


def build_oracle(force: bool = False) -> str:
    src = os.path.join(HERE, "gs2d_oracle.c")
    if not force and _newer(ORACLE_SO, [src]):
        return ORACLE_SO
    os.makedirs(os.path.dirname(ORACLE_SO), exist_ok=True)
    # -ffp-contract=off: every fused multiply-add in the oracle is an explicit fmaf().
    _run(["gcc", "-O2", "-std=c11", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off",
          "-fno-fast-math", "-Wall", "-Wno-unused-variable", "-Wno-unused-but-set-variable",
          "-o", ORACLE_SO, src, "-lm"])
    return ORACLE_SO


def reference_available() -> bool:
    return os.path.isdir(CR) and os.path.isdir(KNN)


def build_ref(force: bool = False) -> str | None:
    """Compile the reference CUDA sources + shim into oracle/_ref/libgs2d_ref.so."""
    if not reference_available():
        return REF_SO if os.path.exists(REF_SO) else None
    shim = os.path.join(HERE, "ref_shim.cu")
    srcs = [
        (shim, []),
        (os.path.join(CR, "rasterizer_impl.cu"), ["-include", "cstdint"]),
        (os.path.join(CR, "forward.cu"), []),
        (os.path.join(CR, "backward.cu"), []),
        (os.path.join(KNN, "simple_knn.cu"), ["-include", "cfloat", "-include", "cstdint"]),
    ]
    if not force and _newer(REF_SO, [s for s, _ in srcs]):
        return REF_SO
    outdir = os.path.dirname(REF_SO)
    os.makedirs(outdir, exist_ok=True)
    inc = ["-I", CR, "-I", os.path.join(RAST, "third_party", "glm"), "-I", KNN]
    # nvcc defaults (-O3 device code, -fmad=true), exactly what the reference's setup.py
    # ends up with: it passes no optimisation or math flags (RAST/setup.py:30).
    common = ["nvcc", "-std=c++17", "-O3", "-Xcompiler", "-fPIC", "-w", "-lineinfo"] + GENCODE + inc

    def compile_one(item):
        src, extra = item
        obj = os.path.join(outdir, os.path.basename(src) + ".o")
        _run(common + ["-include", "cstdint"] + extra + ["-c", src, "-o", obj])
        return obj

    with ThreadPoolExecutor(max_workers=5) as ex:
        objs = list(ex.map(compile_one, srcs))
    _run(["nvcc", "-shared", "-o", REF_SO] + objs + GENCODE)
    for o in objs:
        os.remove(o)
    return REF_SO


if __name__ == "__main__":
    force = "--force" in sys.argv
    print("oracle:", build_oracle(force))
    print("ref   :", build_ref(force))
