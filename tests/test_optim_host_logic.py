"""CPU: the HOST logic of gs2d_b200.optim.MapOptimizer (buffer ownership, re-binding after prune / extend / replace,
ctypes marshalling of gs2d_map_state / gs2d_adam_params) driven through the golden densification sequence of
tests/golden/optim_surgery.npz, with the library replaced by numpy stand-ins that follow the C ABI (pointers -> numpy
views, arithmetic = oracle/optim_ref.py).  The product refuses CPU tensors, so the module's source is loaded with its
`is_cuda` checks textually removed -- this checks the Python around the kernels, not the kernels (tests/test_optim.py,
tests/test_zz_map_compaction.py and tests/test_kernel_emulation.py do that)."""
import contextlib
import ctypes as C
import os
import types

import numpy as np
import torch

import optim_ref as R
import test_zz_map_compaction as T
from gs2d_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def view(ptr, n, dtype=np.float32):
    if isinstance(ptr, C.c_void_p):
        ptr = ptr.value
    if not n:
        return np.zeros(0, dtype)
    return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_byte * (n * np.dtype(dtype).itemsize))).contents).view(dtype)


class FakeLib:
    """What libgs2d.so's map entry points do, on host memory."""

    def gs2d_map_compact_scratch_bytes(self, P):
        return 4 * (1 + (P + 255) // 256)

    def gs2d_map_compact_plan(self, P, keep, scratch, st):
        view(scratch, 1, np.int32)[0] = int((view(keep, P, np.uint8) != 0).sum())
        return 0

    def gs2d_map_compact_move(self, P, cols, keep, src, dst, scratch, st):
        k = view(keep, P, np.uint8) != 0
        n, src, dst = int(k.sum()), src._obj, dst._obj
        for i, (name, w) in enumerate(zip(R.GROUPS, (3, 1, cols, 4, 3))):
            for sp, dp in ((getattr(src, name), getattr(dst, name)), (src.exp_avg[i], dst.exp_avg[i]),
                           (src.exp_avg_sq[i], dst.exp_avg_sq[i])):
                view(dp, n * w).reshape(n, w)[:] = view(sp, P * w).reshape(P, w)[k]
        return 0

    @staticmethod
    def _write_activations(act, P, o, s, r):
        view(o, P)[:] = act["opacities"].ravel()
        view(s, 2 * P)[:] = act["scales"].ravel()
        view(r, 4 * P)[:] = act["rotations"].ravel()

    def gs2d_map_activate(self, P, cols, ms, o, s, r, st):
        ms = ms._obj
        raw = dict(opacity=view(ms.opacity, P).reshape(P, 1), scaling=view(ms.scaling, P * cols).reshape(P, cols),
                   rotation=view(ms.rotation, 4 * P).reshape(P, 4))
        self._write_activations(R.activate_numpy(raw), P, o, s, r)
        return 0

    def gs2d_map_adam_step(self, P, cols, ms, g, ap, o, s, r, st):
        ms, ap = ms._obj, ap._obj
        widths = (3, 1, cols, 4, 3)
        raw = {k: view(getattr(ms, k), P * w).reshape(P, w) for k, w in zip(R.GROUPS, widths)}
        m = {k: view(ms.exp_avg[i], P * w).reshape(P, w) for i, (k, w) in enumerate(zip(R.GROUPS, widths))}
        v = {k: view(ms.exp_avg_sq[i], P * w).reshape(P, w) for i, (k, w) in enumerate(zip(R.GROUPS, widths))}
        x, mm, vv, act = R.map_step_numpy(raw, m, v, view(g, 13 * P).copy(), list(ap.lr), ap.step, (ap.beta1, ap.beta2), ap.eps)
        for k in R.GROUPS:
            raw[k][:], m[k][:], v[k][:] = x[k], mm[k], vv[k]
        self._write_activations(act, P, o, s, r)
        return 0


def _host_only_optim_module():
    src = open(os.path.join(ROOT, "gaus-slam_b200", "gs2d_b200", "optim.py")).read()
    for check in ("not t.is_cuda or ", " or not keep.is_cuda", " or not g.is_cuda", "not tensor.is_cuda or ",
                  " or not flat_grad.is_cuda"):
        assert check in src, check
        src = src.replace(check, "")
    assert "is_cuda" not in src
    src = src.replace("from . import _lib as L", "L = None").replace("from .mapping import FLAT_WIDTH", "FLAT_WIDTH = 13")
    mod = types.ModuleType("optim_host_only")
    exec(compile(src, "optim_host_only", "exec"), mod.__dict__)
    mod.L = types.SimpleNamespace(MapState=L.MapState, AdamParams=L.AdamParams, load=lambda: FakeLib(), check=L.check)
    mod.torch = types.SimpleNamespace(**{k: getattr(torch, k) for k in dir(torch) if not k.startswith("__")})
    mod.torch.cuda = types.SimpleNamespace(device=lambda d: contextlib.nullcontext(),
                                           current_stream=lambda d: types.SimpleNamespace(cuda_stream=0))
    return mod


def test_map_optimizer_host_logic_through_the_golden_surgery():
    mod = _host_only_optim_module()
    g, raw, args = T._golden()
    opt = mod.MapOptimizer(*[torch.tensor(raw[k]) for k in R.GROUPS], args)

    def step(tag):
        opt.step(torch.tensor(g[f"in_grad_{tag}"]))

    step("1"); T._check(opt, g, "1", False)
    step("2"); T._check(opt, g, "2", False)
    keep = torch.tensor(~g["in_remove_mask"])
    before = {k: (opt.raw[k].clone(), opt.exp_avg[k].clone(), opt.exp_avg_sq[k].clone()) for k in R.GROUPS}
    n = opt.prune(keep)
    assert n == int(keep.sum()) == opt.P and opt.flat_grad.numel() == 13 * n and opt.step_count == 2
    for k in R.GROUPS:
        for got, old in zip((opt.raw[k], opt.exp_avg[k], opt.exp_avg_sq[k]), before[k]):
            assert torch.equal(got, old[keep]), k
    T._check(opt, g, "pruned", False)
    step("3"); T._check(opt, g, "3", False)
    P0 = opt.P
    opt.extend(*[torch.tensor(g["in_new_" + k]) for k in R.GROUPS])
    assert opt.P == P0 + g["in_new_xyz"].shape[0] and opt.step_count == 3
    T._check(opt, g, "extended", False)
    step("4"); T._check(opt, g, "4", False)
    opt.replace("opacity", torch.tensor(g["ref_opacity_reset"]))
    T._check(opt, g, "reset", False)
    step("5"); T._check(opt, g, "5", False)
    rp = opt.render_params()
    assert rp["means3D"].data_ptr() == opt.raw["xyz"].data_ptr() and rp["scales"].shape == (opt.P, 2)
    n = opt.prune(torch.zeros(opt.P, dtype=torch.bool))                   # nothing kept
    assert n == 0 and opt.P == 0 and opt.scales.shape == (0, 2) and opt.flat_grad.numel() == 0
