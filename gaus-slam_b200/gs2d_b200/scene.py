"""Deterministic synthetic scenes shared by the oracle, the tests and bench.py.

"Replica-room0-shaped": an axis-aligned 6 x 3 x 4 m box room seen from inside, surfels lying on the
walls with a little jitter, intrinsics of configs/data/replica.yaml (1200x680, fx = fy = 600).
Everything is numpy float32 and depends only on (seed, sizes), so the same scene can be rebuilt on
the GPU box, in this container, and inside the C oracle's tests.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

REPLICA = dict(W=1200, H=680, fx=600.0, fy=600.0, cx=599.5, cy=339.5)
SCANNETPP = dict(W=1752, H=1168, fx=1150.0, fy=1150.0, cx=875.5, cy=583.5)


@dataclass
class Camera:
    W: int
    H: int
    tanfovx: float
    tanfovy: float
    viewmatrix: np.ndarray   # [4,4] float32, = w2c.T (what the reference hands to the kernel)
    projmatrix: np.ndarray   # [4,4] float32, = (P @ w2c).T
    campos: np.ndarray       # [3]
    w2c: np.ndarray          # [4,4]


def make_camera(w2c: np.ndarray, W: int, H: int, fx: float, fy: float, cx: float, cy: float,
                near: float = 0.01, far: float = 100.0) -> Camera:
    """Same matrices as render/render_2dgs.py:6-31 (setup_camera)."""
    w2c = np.asarray(w2c, dtype=np.float32)
    proj = np.array([[2 * fx / W, 0.0, -(W - 2 * cx) / W, 0.0],
                     [0.0, 2 * fy / H, -(H - 2 * cy) / H, 0.0],
                     [0.0, 0.0, far / (far - near), -(far * near) / (far - near)],
                     [0.0, 0.0, 1.0, 0.0]], dtype=np.float32)
    view_t = np.ascontiguousarray(w2c.T)
    full = np.ascontiguousarray((view_t @ proj.T).astype(np.float32))
    campos = np.linalg.inv(w2c.astype(np.float64))[:3, 3].astype(np.float32)
    return Camera(W=W, H=H, tanfovx=W / (2 * fx), tanfovy=H / (2 * fy), viewmatrix=view_t,
                  projmatrix=full, campos=campos, w2c=w2c)


def yaw_pose(yaw_deg: float, t=(0.0, 0.0, 0.0)) -> np.ndarray:
    a = np.deg2rad(yaw_deg)
    c, s = np.cos(a), np.sin(a)
    w2c = np.eye(4, dtype=np.float32)
    w2c[:3, :3] = np.array([[c, 0, -s], [0, 1, 0], [s, 0, c]], dtype=np.float32)
    w2c[:3, 3] = np.asarray(t, dtype=np.float32)
    return w2c


def view_sweep(K: int, intr=REPLICA, seed: int = 0):
    """K cameras on a +-30 degree yaw sweep with +-0.5 m translation (multi-view configs)."""
    rng = np.random.default_rng(seed + 1000)
    cams = []
    for k in range(K):
        yaw = 0.0 if K == 1 else -30.0 + 60.0 * k / (K - 1)
        t = rng.uniform(-0.5, 0.5, size=3) * np.array([1.0, 0.3, 1.0])
        if k == 0:
            yaw, t = 0.0, np.zeros(3)
        cams.append(make_camera(yaw_pose(yaw, t), **intr))
    return cams


def _normal_to_quat(n: np.ndarray, spin: np.ndarray) -> np.ndarray:
    """Unit quaternions (w,x,y,z) of rotations whose third column is n, spun by `spin` in-plane."""
    a = np.tile(np.array([0.0, 1.0, 0.0]), (n.shape[0], 1))
    par = np.abs(n[:, 1]) > 0.9
    a[par] = np.array([1.0, 0.0, 0.0])
    t1 = np.cross(a, n)
    t1 /= np.linalg.norm(t1, axis=1, keepdims=True)
    t2 = np.cross(n, t1)
    c, s = np.cos(spin)[:, None], np.sin(spin)[:, None]
    u = c * t1 + s * t2
    v = -s * t1 + c * t2
    R = np.stack([u, v, n], axis=2)  # columns
    # matrix -> quaternion (robust branch-free form)
    m00, m11, m22 = R[:, 0, 0], R[:, 1, 1], R[:, 2, 2]
    qw = np.sqrt(np.maximum(0.0, 1 + m00 + m11 + m22)) / 2
    qx = np.sqrt(np.maximum(0.0, 1 + m00 - m11 - m22)) / 2
    qy = np.sqrt(np.maximum(0.0, 1 - m00 + m11 - m22)) / 2
    qz = np.sqrt(np.maximum(0.0, 1 - m00 - m11 + m22)) / 2
    qx = np.copysign(qx, R[:, 2, 1] - R[:, 1, 2])
    qy = np.copysign(qy, R[:, 0, 2] - R[:, 2, 0])
    qz = np.copysign(qz, R[:, 1, 0] - R[:, 0, 1])
    q = np.stack([qw, qx, qy, qz], axis=1)
    return q / np.linalg.norm(q, axis=1, keepdims=True)


def room_scene(P: int, intr=REPLICA, seed: int = 0, scale_mult: float = 1.0):
    """P surfels on the walls of a 6 x 3 x 4 m room, generated from the identity-pose camera.

    Returns a dict of float32 arrays: means3D [P,3], scales [P,2], rotations [P,4] (w,x,y,z, unit),
    opacities [P,1], colors [P,3].
    """
    rng = np.random.default_rng(seed)
    W, H, fx, fy, cx, cy = intr["W"], intr["H"], intr["fx"], intr["fy"], intr["cx"], intr["cy"]
    u = rng.uniform(-0.02 * W, 1.02 * W, size=P)
    v = rng.uniform(-0.02 * H, 1.02 * H, size=P)
    d = np.stack([(u - cx) / fx, (v - cy) / fy, np.ones(P)], axis=1)   # ray, z = 1
    # room (camera frame of the identity pose, y down): x in [-2.5,3.5], y in [-1.2,1.8], z in [-1,3];
    # off-centre so that the front wall, the left wall and the ceiling are all in view
    lo, hi = np.array([-2.5, -1.2, -1.0]), np.array([3.5, 1.8, 3.0])
    with np.errstate(divide="ignore", invalid="ignore"):
        t_axis = np.where(d > 0, hi / d, np.where(d < 0, lo / d, np.inf))
    axis = np.argmin(t_axis, axis=1)
    z = t_axis[np.arange(P), axis]                                   # view depth of the wall hit
    z = z + rng.normal(0.0, 0.01, size=P)
    pts = d * z[:, None]
    n = np.zeros((P, 3))
    n[np.arange(P), axis] = -np.sign(d[np.arange(P), axis])          # inward wall normal
    n = n + rng.normal(0.0, 0.05, size=(P, 3))
    n /= np.linalg.norm(n, axis=1, keepdims=True)
    q = _normal_to_quat(n, rng.uniform(0, 2 * np.pi, size=P))
    s = (z / ((fx + fy) / 2))[:, None] * rng.uniform(0.7, 1.5, size=(P, 2)) * scale_mult
    opa = 1.0 / (1.0 + np.exp(-rng.normal(0.0, 1.5, size=(P, 1))))
    rgb = rng.uniform(0.0, 1.0, size=(P, 3))
    f32 = np.float32
    return dict(means3D=pts.astype(f32), scales=s.astype(f32), rotations=q.astype(f32),
                opacities=opa.astype(f32), colors=rgb.astype(f32))


def image_grads(W: int, H: int, seed: int = 0):
    """Upstream gradients dL/dcolor [3,H,W], dL/dallmap [7,H,W] ~ N(0,1)/(W*H)."""
    rng = np.random.default_rng(seed + 77)
    px = W * H
    g_color = (rng.normal(size=(3, H, W)) / px).astype(np.float32)
    g_others = (rng.normal(size=(7, H, W)) / px).astype(np.float32)
    return g_color, g_others
