"""TEST INFRASTRUCTURE ONLY (never imported by the product): the reference's pose transform, restated.

`pose_transform_torch` follows Renderer_tracking / Renderer_BA (render/__init__.py:35-36, :97-98) and
Transform.get_transform_matrix (scene/Frame.py:87-96) line by line with torch, so autograd gives the gradients
the reference gets.  The reference takes its quaternion algebra from pytorch3d (requirements.txt / README: version
unpinned, "stable"), which is NOT in the reference checkout and not installed here: `quaternion_to_matrix`,
`matrix_to_quaternion`, `quaternion_raw_multiply`, `standardize_quaternion`, `quaternion_multiply` below restate
pytorch3d/transforms/rotation_conversions.py (v0.7.x) from its published algorithm.  Parity pin: the matrix algebra
and its gradients are pinned by torch autograd; the pytorch3d restatement is pinned only by its own invariants
(round trip, R(q) orthonormal, sign convention) -- "parity unpinned" for the sign of rotations', which cannot
change a rendered image (q and -q are the same rotation).
`pose_backward_numpy` is the float64 hand-derived chain rule the CUDA kernel implements.
"""
import numpy as np


# ---- pytorch3d.transforms restated (torch) ------------------------------------------------------------
def quaternion_to_matrix(quaternions):
    import torch
    r, i, j, k = torch.unbind(quaternions, -1)
    two_s = 2.0 / (quaternions * quaternions).sum(-1)
    o = torch.stack((1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                     two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                     two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return o.reshape(quaternions.shape[:-1] + (3, 3))


def _sqrt_positive_part(x):
    import torch
    ret = torch.zeros_like(x)
    positive_mask = x > 0
    ret[positive_mask] = torch.sqrt(x[positive_mask])
    return ret


def standardize_quaternion(quaternions):
    import torch
    return torch.where(quaternions[..., 0:1] < 0, -quaternions, quaternions)


def matrix_to_quaternion(matrix):
    import torch
    batch_dim = matrix.shape[:-2]
    m00, m01, m02, m10, m11, m12, m20, m21, m22 = torch.unbind(matrix.reshape(batch_dim + (9,)), dim=-1)
    q_abs = _sqrt_positive_part(torch.stack([1.0 + m00 + m11 + m22, 1.0 + m00 - m11 - m22, 1.0 - m00 + m11 - m22,
                                             1.0 - m00 - m11 + m22], dim=-1))
    quat_by_rijk = torch.stack([
        torch.stack([q_abs[..., 0] ** 2, m21 - m12, m02 - m20, m10 - m01], dim=-1),
        torch.stack([m21 - m12, q_abs[..., 1] ** 2, m10 + m01, m02 + m20], dim=-1),
        torch.stack([m02 - m20, m10 + m01, q_abs[..., 2] ** 2, m12 + m21], dim=-1),
        torch.stack([m10 - m01, m20 + m02, m21 + m12, q_abs[..., 3] ** 2], dim=-1)], dim=-2)
    flr = torch.tensor(0.1).to(dtype=q_abs.dtype, device=q_abs.device)
    quat_candidates = quat_by_rijk / (2.0 * q_abs[..., None].max(flr))
    out = quat_candidates[torch.nn.functional.one_hot(q_abs.argmax(dim=-1), num_classes=4) > 0.5, :].reshape(batch_dim + (4,))
    return standardize_quaternion(out)


def quaternion_raw_multiply(a, b):
    import torch
    aw, ax, ay, az = torch.unbind(a, -1)
    bw, bx, by, bz = torch.unbind(b, -1)
    ow = aw * bw - ax * bx - ay * by - az * bz
    ox = aw * bx + ax * bw + ay * bz - az * by
    oy = aw * by - ax * bz + ay * bw + az * bx
    oz = aw * bz + ax * by - ay * bx + az * bw
    return torch.stack((ow, ox, oy, oz), -1)


def quaternion_multiply(a, b):
    return standardize_quaternion(quaternion_raw_multiply(a, b))


# ---- the reference's code path ---------------------------------------------------------------------------
def get_transform_matrix(cam_rot, cam_trans):
    """scene/Frame.py:87-96 (build_rotation = pytorch3d quaternion_to_matrix, utils/common_utils.py:44-45)."""
    import torch
    q = torch.nn.functional.normalize(cam_rot[None, ...])[0]
    w2c = torch.eye(4, dtype=cam_rot.dtype, device=cam_rot.device)
    w2c[:3, :3] = quaternion_to_matrix(q[None, ...])[0]
    w2c[:3, 3] = cam_trans
    return w2c


def pose_transform_torch(means3D, rotations, w2c):
    """render/__init__.py:35-36 / :97-98."""
    means = (w2c[:3, :3] @ means3D.T + w2c[:3, 3:]).T
    rots = quaternion_multiply(matrix_to_quaternion(w2c[None, :3, :3]), rotations).detach()
    return means, rots


# ---- float64 hand-derived backward (what csrc/pose.cu implements) ----------------------------------------
def pose_backward_numpy(xyz, g, cam_rot=None, w2c=None):
    """g = dL/dmeans3D' [P,3].  -> dict(dR [3,3], dt [3], dxyz [P,3], and dq [4] when cam_rot is given)."""
    x = xyz.astype(np.float64); g = g.astype(np.float64)
    G = g.T @ x
    out = dict(dR=G, dt=g.sum(0))
    if cam_rot is None:
        R = w2c[:3, :3].astype(np.float64)
    else:
        q = cam_rot.astype(np.float64)
        n = max(np.sqrt((q * q).sum()), 1e-12)
        r, i, j, k = q / n
        s = r * r + i * i + j * j + k * k
        two_s = 2.0 / s
        N = np.array([[-(j * j + k * k), i * j - k * r, i * k + j * r],
                      [i * j + k * r, -(i * i + k * k), j * k - i * r],
                      [i * k - j * r, j * k + i * r, -(i * i + j * j)]])
        R = np.eye(3) + two_s * N
        dN = np.array([
            [[0, -k, j], [k, 0, -i], [-j, i, 0]],
            [[0, j, k], [j, -2 * i, -r], [k, r, -2 * i]],
            [[-2 * j, i, r], [i, 0, k], [-r, k, -2 * j]],
            [[-2 * k, -r, i], [r, -2 * k, j], [i, j, 0]]])
        qh = np.array([r, i, j, k])
        dqh = two_s * (dN * G[None]).sum((1, 2)) - 4.0 * (G * N).sum() / (s * s) * qh
        out["dq"] = (dqh - qh * (qh * dqh).sum()) / n
    out["dxyz"] = g @ R
    return out
