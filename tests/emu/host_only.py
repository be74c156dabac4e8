"""TEST INFRASTRUCTURE: run the product's Python wrappers (gs2d_b200.optim / loss / tracking) on HOST tensors against
the host build of their kernels (tests/emu/build_emu.py).

The product refuses CPU tensors and has no CPU path; that stays true.  For these tests only, a wrapper module's source
is loaded with (a) its `is_cuda` argument checks textually removed, (b) `torch.cuda.device / current_stream` replaced
by no-ops, (c) `_lib.load()` returning the emulated library.  Everything else in the wrapper -- shapes, strides,
scratch handling, ctypes marshalling, the autograd bridge -- runs unchanged, and below it the kernels' own source text.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import os
import re
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
import build_emu  # noqa: E402
from gs2d_b200 import _lib as L  # noqa: E402

_lib = None


def emulated_library():
    """libgs2d_emu.so with the prototypes of the product's ctypes mirror (for the entry points it has)."""
    global _lib
    if _lib is None:
        lib = C.CDLL(build_emu.build())
        for name, (restype, argtypes) in L.PROTOTYPES.items():
            if hasattr(lib, name):
                fn = getattr(lib, name)
                fn.restype, fn.argtypes = restype, argtypes
        _lib = lib
    return _lib


def load(module: str):
    """gs2d_b200.<module> as a host-only module bound to the emulated library."""
    src = open(os.path.join(ROOT, "gaus-slam_b200", "gs2d_b200", module + ".py")).read()
    src, n1 = re.subn(r"not (\w+)\.is_cuda or ", "", src)
    src, n2 = re.subn(r" or not (\w+)\.is_cuda", "", src)
    assert n1 + n2 >= 1 and "is_cuda" not in src, f"{module}: unexpected CUDA checks left"
    src = src.replace("from . import _lib as L", "L = None").replace("from .mapping import FLAT_WIDTH", "FLAT_WIDTH = 13")
    assert "from ." not in src, f"{module}: relative import not handled"
    mod = types.ModuleType(f"gs2d_b200_host_only_{module}")
    exec(compile(src, f"<host-only {module}.py>", "exec"), mod.__dict__)
    ns = types.SimpleNamespace(**{k: getattr(L, k) for k in dir(L) if not k.startswith("__")})
    ns.load = emulated_library
    ns.raw_stream = lambda torch_, device: 0
    mod.L = ns
    fake_torch = types.SimpleNamespace(**{k: getattr(torch, k) for k in dir(torch) if not k.startswith("__")})
    fake_torch.cuda = types.SimpleNamespace(device=lambda d: contextlib.nullcontext(),
                                            current_stream=lambda d=None: types.SimpleNamespace(cuda_stream=0))
    mod.torch = fake_torch
    return mod


class _FakeEvent:
    def record(self, stream=None):
        pass

    def synchronize(self):
        pass

    def query(self):
        return True


def _fake_torch():
    fake = types.SimpleNamespace(**{k: getattr(torch, k) for k in dir(torch) if not k.startswith("__")})
    fake.cuda = types.SimpleNamespace(device=lambda d: contextlib.nullcontext(), Event=_FakeEvent,
                                      current_stream=lambda d=None: types.SimpleNamespace(cuda_stream=0))
    return fake


def load_rasterizer():
    """The drop-in package gaus_2dgs_rasterization (its _C glue and its autograd bridge) as host-only modules over
    libgs2d_emu_full.so: `is_cuda` checks and `.pin_memory()` removed textually, torch.cuda replaced by no-ops."""
    sys.path.insert(0, HERE)
    import emu_run
    pkg = os.path.join(ROOT, "gaus-slam_b200", "gaus_2dgs_rasterization")
    src = open(os.path.join(pkg, "_C.py")).read()
    src, n = re.subn(r"    if not t\.is_cuda:\n        raise RuntimeError\([^\n]*\)\n", "", src)
    assert n == 1 and "is_cuda" not in src
    src = src.replace(".pin_memory()", "")
    src = src.replace("from gs2d_b200 import _lib as L", "L = None")
    glue = types.ModuleType("gaus_2dgs_rasterization_host_only._C")
    exec(compile(src, "<host-only _C.py>", "exec"), glue.__dict__)
    ns = types.SimpleNamespace(**{k: getattr(L, k) for k in dir(L) if not k.startswith("__")})
    ns.load = emu_run.library
    ns.raw_stream = lambda torch_, device: 0
    glue.L, glue.torch = ns, _fake_torch()
    src = open(os.path.join(pkg, "__init__.py")).read()
    assert "from . import _C" in src
    src = src.replace("from . import _C", "_C = None")
    mod = types.ModuleType("gaus_2dgs_rasterization_host_only")
    mod.__file__ = os.path.join(pkg, "__init__.py")
    exec(compile(src, "<host-only gaus_2dgs_rasterization/__init__.py>", "exec"), mod.__dict__)
    mod._C = glue
    return mod


def load_simple_knn():
    """simple_knn._C (distCUDA2) as a host-only module over the host build of the whole library."""
    sys.path.insert(0, HERE)
    import emu_run
    src = open(os.path.join(ROOT, "gaus-slam_b200", "simple_knn", "_C.py")).read()
    src, n = re.subn(r"    if not points\.is_cuda:\n        raise RuntimeError\([^\n]*\)\n", "", src)
    assert n == 1 and "is_cuda" not in src
    src = src.replace("from gs2d_b200 import _lib as L", "L = None")
    mod = types.ModuleType("simple_knn_host_only._C")
    mod.__file__ = os.path.join(ROOT, "gaus-slam_b200", "simple_knn", "_C.py")
    exec(compile(src, "<host-only simple_knn/_C.py>", "exec"), mod.__dict__)
    ns = types.SimpleNamespace(**{k: getattr(L, k) for k in dir(L) if not k.startswith("__")})
    ns.load = emu_run.library
    mod.L, mod.torch = ns, _fake_torch()
    return mod
