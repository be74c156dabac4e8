"""Fused pose transform: the host side of gs2d_pose_transform_forward / _backward.

What Renderer_tracking / Renderer_BA do around the rasterizer (render/__init__.py:35-36, :97-98):

    means3D'   = (w2c[:3,:3] @ means3D.T + w2c[:3,3:]).T
    rotations' = quaternion_multiply(matrix_to_quaternion(w2c[None,:3,:3]), rotations).detach()

as one kernel each way.  The pose is either the tracker's own parameters (scene/Frame.py:66-67: `cam_rot` [4]
unnormalised quaternion, real first, and `cam_trans` [3]; w2c is then Transform.get_transform_matrix, :87-96) or a
4x4 `w2c` tensor (fix_w2c, LocalMap.get_frame_w2c).  Gradients flow to the pose (7 scalars, or the 4x4) and, when
`means3D` requires grad (bundle adjustment), to means3D; `rotations'` carries none, like the reference's `.detach()`.

    means_c, rots_c = pose_transform(xyz, rotations, cam_rot=q, cam_trans=t)      # tracking
    means_c, rots_c = pose_transform(xyz, rotations, w2c=w2c)                     # BA with a composed pose

CUDA only: there is no CPU path.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib as L

_scratch = {}


def _scratch_for(dev):
    key = (str(dev), L.raw_stream(torch, dev))
    if key not in _scratch:      # zero-initialised once; the backward kernel resets its ticket itself
        _scratch[key] = torch.zeros(L.load().gs2d_pose_transform_scratch_bytes(), dtype=torch.uint8, device=dev)
    return _scratch[key]


def _f32(t, name, shape):
    if not t.is_cuda or t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be a float32 CUDA tensor (there is no CPU path)")
    if tuple(t.shape) != shape:
        raise RuntimeError(f"{name} has shape {tuple(t.shape)}, expected {shape}")
    return t.contiguous()


def _ptr(t):
    return None if t is None else t.data_ptr()


class _PoseTransform(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xyz, rotations, cam_rot, cam_trans, w2c):
        P = int(xyz.shape[0])
        dev = xyz.device
        xyz = _f32(xyz, "means3D", (P, 3))
        rotations = _f32(rotations, "rotations", (P, 4))
        if w2c is None:
            cam_rot, cam_trans = _f32(cam_rot, "cam_rot", (4,)), _f32(cam_trans, "cam_trans", (3,))
        else:
            w2c = _f32(w2c, "w2c", (4, 4))
        means_out = torch.empty_like(xyz)
        rot_out = torch.empty_like(rotations)
        with torch.cuda.device(dev):
            st = C.c_void_p(L.raw_stream(torch, dev))
            L.check(L.load().gs2d_pose_transform_forward(P, xyz.data_ptr(), rotations.data_ptr(), _ptr(cam_rot),
                                                         _ptr(cam_trans), _ptr(w2c), means_out.data_ptr(),
                                                         rot_out.data_ptr(), st), "gs2d_pose_transform_forward")
        ctx.save_for_backward(xyz, cam_rot, cam_trans, w2c)
        ctx.need_xyz = ctx.needs_input_grad[0]
        ctx.mark_non_differentiable(rot_out)
        return means_out, rot_out

    @staticmethod
    def backward(ctx, g_means, _g_rot):
        xyz, cam_rot, cam_trans, w2c = ctx.saved_tensors
        P = int(xyz.shape[0])
        dev = xyz.device
        g = g_means.to(torch.float32).contiguous()
        d_rot = d_trans = d_w2c = None
        if w2c is None:
            d_rot, d_trans = torch.empty(4, device=dev), torch.empty(3, device=dev)
        else:
            d_w2c = torch.empty((4, 4), device=dev)
        d_xyz = torch.empty_like(xyz) if ctx.need_xyz else None
        with torch.cuda.device(dev):
            st = C.c_void_p(L.raw_stream(torch, dev))
            L.check(L.load().gs2d_pose_transform_backward(P, xyz.data_ptr(), _ptr(cam_rot), _ptr(cam_trans), _ptr(w2c),
                                                          g.data_ptr(), _scratch_for(dev).data_ptr(), _ptr(d_rot),
                                                          _ptr(d_trans), _ptr(d_w2c), _ptr(d_xyz), st),
                    "gs2d_pose_transform_backward")
        return d_xyz, None, d_rot, d_trans, d_w2c


def pose_transform(means3D, rotations, cam_rot=None, cam_trans=None, w2c=None):
    """-> (means3D', rotations') in the camera frame; exactly one of (cam_rot, cam_trans) / w2c."""
    if (w2c is None) == (cam_rot is None or cam_trans is None) or (w2c is not None and (cam_rot is not None or cam_trans is not None)):
        raise Exception("Please provide exactly one of either the (cam_rot, cam_trans) pair or w2c!")
    return _PoseTransform.apply(means3D, rotations, cam_rot, cam_trans, w2c)


class PoseOptimizer:
    """The tracker's pose optimiser (scene/Frame.py:66-73: torch.optim.Adam over `cam_rot` [4] and `cam_trans` [3] with
    one learning rate each, betas (0.7, 0.99), eps 1e-8; stepped once per tracking iteration, slam/Frontend.py:92) as
    one kernel launch per step: torch's foreach Adam issues ~25 launches for these two tiny parameters, which is most
    of the host time of a tracking iteration once the rasterizer is fast.

        opt = PoseOptimizer(cam_rot, cam_trans, lr_rot=4e-4, lr_trans=2e-3)
        opt.zero_grad(); loss.backward(); opt.step()

    Same results as torch.optim.Adam (operation order of torch/optim/adam.py); a parameter whose .grad is None is
    skipped together with its moments and its step counter, like torch."""

    def __init__(self, cam_rot, cam_trans, lr_rot, lr_trans, betas=(0.7, 0.99), eps=1e-8):
        for t, name, n in ((cam_rot, "cam_rot", 4), (cam_trans, "cam_trans", 3)):
            if not t.is_cuda or t.dtype != torch.float32 or tuple(t.shape) != (n,) or not t.is_contiguous():
                raise RuntimeError(f"{name} must be a contiguous float32 CUDA tensor of shape ({n},) (there is no CPU path)")
        self.cam_rot, self.cam_trans = cam_rot, cam_trans
        self.lr = [float(lr_rot), float(lr_trans)]
        self.betas, self.eps = (float(betas[0]), float(betas[1])), float(eps)
        self.state = torch.zeros(14, dtype=torch.float32, device=cam_rot.device)
        self.steps = [0, 0]

    def zero_grad(self, set_to_none=True):
        for p in (self.cam_rot, self.cam_trans):
            if set_to_none or p.grad is None:
                p.grad = None
            else:
                p.grad.zero_()

    def step(self):
        g = [self.cam_rot.grad, self.cam_trans.grad]
        for k in range(2):
            gk = g[k]
            if gk is not None:
                if not gk.is_cuda or gk.dtype != torch.float32:
                    raise RuntimeError("pose gradients must be float32 CUDA tensors")
                g[k] = gk.contiguous()
                self.steps[k] += 1
        ap = L.PoseAdamParams(lr=(C.c_double * 2)(*self.lr), beta1=self.betas[0], beta2=self.betas[1], eps=self.eps,
                              step=(C.c_int64 * 2)(*self.steps))
        dev = self.cam_rot.device
        with torch.cuda.device(dev):
            st = C.c_void_p(L.raw_stream(torch, dev))
            L.check(L.load().gs2d_pose_adam_step(self.cam_rot.data_ptr(), self.cam_trans.data_ptr(), _ptr(g[0]), _ptr(g[1]),
                                                 self.state.data_ptr(), C.byref(ap), st), "gs2d_pose_adam_step")
