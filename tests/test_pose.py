"""Fused pose transform + pose-gradient reduction.  CPU: the restated pytorch3d algebra keeps
its invariants and the hand-derived chain rule (float64 numpy) equals torch autograd through the reference's code
path.  GPU: gs2d_pose_transform_forward / _backward through the C ABI (gs2d_b200.tracking.pose_transform) agree with
the torch transliteration: forward 1e-6 relative, pose gradients 5e-4 relative (bar: 1e-3) (fp32 sums over P in another order)."""
import numpy as np
import pytest
import torch

import pose_ref as R


def _scene(P, seed):
    rng = np.random.default_rng(seed)
    xyz = rng.uniform(-3, 3, (P, 3)).astype(np.float32)
    rot = rng.normal(0, 1, (P, 4)).astype(np.float32)
    rot /= np.linalg.norm(rot, axis=1, keepdims=True)
    g = (rng.normal(0, 1, (P, 3)) * 10.0 ** rng.uniform(-5, -2, (P, 1))).astype(np.float32)
    return xyz, rot, g


POSES = [np.array([1.0, 0.0, 0.0, 0.0]), np.array([0.9, 0.1, -0.3, 0.2]), np.array([-0.2, 1.4, 0.3, -0.8]),
         np.array([1e-3, 0.0, 1.0, 0.0]), np.array([0.0, 0.0, 0.0, -2.0])]


def _rel(a, b):
    return float(np.abs(np.asarray(a, np.float64) - b).max() / max(np.abs(b).max(), 1e-30))


def _rot_err(a, b):
    """Max abs difference of two quaternion arrays; rows whose real part is ~0 may legitimately differ in sign
    (standardize_quaternion flips on w < 0, and w = 0 +- rounding is decided differently in fp32 and fp64)."""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    d = np.abs(a - b).max(1)
    flip = np.abs(a + b).max(1)
    free = np.abs(b[:, 0]) < 1e-5
    return float(np.where(free, np.minimum(d, flip), d).max())


def _torch_path(xyz, rot, g, q=None, t=None, w2c=None, dev="cpu", dtype=torch.float64, xyz_grad=True):
    x = torch.tensor(xyz, device=dev, dtype=dtype, requires_grad=xyz_grad)
    r = torch.tensor(rot, device=dev, dtype=dtype)
    leaves = {}
    if w2c is None:
        leaves["q"] = torch.tensor(q, device=dev, dtype=dtype, requires_grad=True)
        leaves["t"] = torch.tensor(t, device=dev, dtype=dtype, requires_grad=True)
        m = R.get_transform_matrix(leaves["q"], leaves["t"])
    else:
        leaves["w2c"] = torch.tensor(w2c, device=dev, dtype=dtype, requires_grad=True)
        m = leaves["w2c"]
    means, rots = R.pose_transform_torch(x, r, m)
    assert not rots.requires_grad
    means.backward(torch.tensor(g, device=dev, dtype=dtype))
    out = dict(means=means.detach().cpu().numpy(), rots=rots.cpu().numpy(), w2c=m.detach().cpu().numpy())
    out.update({"d" + k: v.grad.cpu().numpy() for k, v in leaves.items()})
    if xyz_grad:
        out["dxyz"] = x.grad.cpu().numpy()
    return out


def test_restated_quaternion_algebra_invariants():
    rng = np.random.default_rng(0)
    q = torch.tensor(rng.normal(0, 1, (500, 4)))
    q = torch.nn.functional.normalize(q)
    Rm = R.quaternion_to_matrix(q)
    eye = torch.eye(3, dtype=Rm.dtype).expand(500, 3, 3)
    assert torch.allclose(Rm @ Rm.transpose(1, 2), eye, atol=1e-12) and torch.allclose(torch.linalg.det(Rm), torch.ones(500, dtype=Rm.dtype))
    back = R.matrix_to_quaternion(Rm)
    assert torch.all(back[:, 0] >= 0) and torch.allclose(back, R.standardize_quaternion(q), atol=1e-9)
    # composition: R(a * b) = R(a) R(b)
    a, b = q[:250], q[250:]
    assert torch.allclose(R.quaternion_to_matrix(R.quaternion_multiply(a, b)), Rm[:250] @ Rm[250:], atol=1e-12)
    # 180-degree rotations (w = 0) take the other candidates of matrix_to_quaternion
    for axis in range(3):
        h = torch.zeros(1, 4, dtype=torch.float64)
        h[0, 1 + axis] = 1.0
        assert torch.allclose(R.quaternion_to_matrix(R.matrix_to_quaternion(R.quaternion_to_matrix(h))), R.quaternion_to_matrix(h), atol=1e-12)


@pytest.mark.parametrize("pose", range(len(POSES)))
def test_hand_derived_chain_rule_equals_autograd(pose):
    xyz, rot, g = _scene(300, pose)
    q, t = POSES[pose], np.array([0.3, -0.2, 0.5])
    ref = _torch_path(xyz, rot, g, q=q, t=t)
    mine = R.pose_backward_numpy(xyz, g, cam_rot=q)
    assert _rel(mine["dq"], ref["dq"]) <= 1e-10 and _rel(mine["dt"], ref["dt"]) <= 1e-12
    assert _rel(mine["dxyz"], ref["dxyz"]) <= 1e-12
    # matrix form
    ref2 = _torch_path(xyz, rot, g, w2c=ref["w2c"])
    mine2 = R.pose_backward_numpy(xyz, g, w2c=ref["w2c"])
    assert _rel(mine2["dR"], ref2["dw2c"][:3, :3]) <= 1e-12 and _rel(mine2["dt"], ref2["dw2c"][:3, 3]) <= 1e-12
    assert np.all(ref2["dw2c"][3] == 0)


def _ours(xyz, rot, g, q=None, t=None, w2c=None, xyz_grad=True):
    from gs2d_b200 import tracking
    dev = torch.device("cuda:0")
    x = torch.tensor(xyz, device=dev, requires_grad=xyz_grad)
    r = torch.tensor(rot, device=dev)
    kw, leaves = {}, {}
    if w2c is None:
        leaves["q"] = torch.tensor(q, device=dev, dtype=torch.float32, requires_grad=True)
        leaves["t"] = torch.tensor(t, device=dev, dtype=torch.float32, requires_grad=True)
        kw = dict(cam_rot=leaves["q"], cam_trans=leaves["t"])
    else:
        leaves["w2c"] = torch.tensor(w2c, device=dev, dtype=torch.float32, requires_grad=True)
        kw = dict(w2c=leaves["w2c"])
    means, rots = tracking.pose_transform(x, r, **kw)
    assert not rots.requires_grad
    means.backward(torch.tensor(g, device=dev))
    out = dict(means=means.detach().cpu().numpy(), rots=rots.cpu().numpy())
    out.update({"d" + k: v.grad.cpu().numpy() for k, v in leaves.items()})
    if xyz_grad:
        out["dxyz"] = x.grad.cpu().numpy()
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("P,pose", [(300, 0), (300, 1), (5000, 2), (1, 3), (300, 4), (300001, 1)])
def test_fused_pose_transform_equals_the_reference_path(P, pose):
    xyz, rot, g = _scene(P, 10 + pose)
    q, t = POSES[pose], np.array([0.3, -0.2, 0.5])
    ref = _torch_path(xyz, rot, g, q=q, t=t)                    # float64 torch on the CPU
    for rep in range(2):                                         # twice: the ticket must be left ready
        got = _ours(xyz, rot, g, q=q, t=t, xyz_grad=(rep == 0))
        assert _rel(got["means"], ref["means"]) <= 1e-6
        assert _rot_err(got["rots"], ref["rots"]) <= 2e-6
        assert _rel(got["dt"], ref["dt"]) <= 5e-4 and _rel(got["dq"], ref["dq"]) <= 5e-4
        if rep == 0:
            assert _rel(got["dxyz"], ref["dxyz"]) <= 1e-6
    # matrix form (fix_w2c / LocalMap.get_frame_w2c)
    got2 = _ours(xyz, rot, g, w2c=ref["w2c"])
    ref2 = _torch_path(xyz, rot, g, w2c=ref["w2c"])
    assert _rel(got2["means"], ref2["means"]) <= 1e-6 and _rot_err(got2["rots"], ref2["rots"]) <= 2e-6
    assert _rel(got2["dw2c"], ref2["dw2c"]) <= 5e-4 and np.all(got2["dw2c"][3] == 0)


@pytest.mark.gpu
def test_fused_pose_transform_vs_fp32_torch_on_the_gpu():
    """Same thing against the transliteration run in fp32 on the GPU (what the reference actually executes)."""
    xyz, rot, g = _scene(20000, 42)
    q, t = POSES[1], np.array([0.1, 0.0, -0.4])
    ref = _torch_path(xyz, rot, g, q=q, t=t, dev="cuda:0", dtype=torch.float32)
    got = _ours(xyz, rot, g, q=q, t=t)
    assert _rel(got["means"], ref["means"]) <= 2e-6 and _rot_err(got["rots"], ref["rots"]) <= 2e-6
    assert _rel(got["dq"], ref["dq"]) <= 1e-3 and _rel(got["dt"], ref["dt"]) <= 1e-3


@pytest.mark.gpu
def test_pose_transform_argument_errors():
    from gs2d_b200 import tracking
    dev = torch.device("cuda:0")
    x, r = torch.zeros((4, 3), device=dev), torch.zeros((4, 4), device=dev)
    q, t = torch.tensor([1.0, 0, 0, 0], device=dev), torch.zeros(3, device=dev)
    with pytest.raises(Exception, match="exactly one"):
        tracking.pose_transform(x, r)
    with pytest.raises(Exception, match="exactly one"):
        tracking.pose_transform(x, r, cam_rot=q, cam_trans=t, w2c=torch.eye(4, device=dev))
    with pytest.raises(RuntimeError):
        tracking.pose_transform(x.cpu(), r.cpu(), cam_rot=q.cpu(), cam_trans=t.cpu())


def pose_adam_sequence(make_optimizer, device, steps=6):
    """Drives an optimiser over cam_rot / cam_trans with a fixed gradient sequence (step 3: cam_trans has no gradient,
    which torch skips together with its moments and step counter) -> the parameters after every step."""
    rng = np.random.default_rng(11)
    q = torch.tensor([0.9995, 0.01, -0.02, 0.015], dtype=torch.float32, device=device, requires_grad=True)
    t = torch.tensor([0.03, -0.02, 0.05], dtype=torch.float32, device=device, requires_grad=True)
    opt = make_optimizer(q, t)
    out = []
    for i in range(steps):
        opt.zero_grad(set_to_none=True)
        q.grad = torch.tensor(rng.normal(0, 10.0 ** rng.integers(-3, 2), 4), dtype=torch.float32, device=device)
        if i != 3:
            t.grad = torch.tensor(rng.normal(0, 10.0 ** rng.integers(-3, 2), 3), dtype=torch.float32, device=device)
        opt.step()
        out.append((q.detach().cpu().numpy().copy(), t.detach().cpu().numpy().copy()))
    return out


def torch_pose_adam(q, t):
    # scene/Frame.py:70-73
    return torch.optim.Adam([{"params": [q], "lr": 4e-4, "name": "cam_rots"}, {"params": [t], "lr": 2e-3, "name": "cam_trans"}],
                            lr=0.0, eps=1e-8, betas=(0.7, 0.99))


def check_pose_adam(got, want):
    for i, ((q1, t1), (q0, t0)) in enumerate(zip(got, want)):
        assert np.allclose(q1, q0, rtol=2e-6, atol=1e-9), (i, q1, q0)
        assert np.allclose(t1, t0, rtol=2e-6, atol=1e-9), (i, t1, t0)


@pytest.mark.gpu
def test_fused_pose_adam_equals_torch_adam():
    """gs2d_b200.tracking.PoseOptimizer (one launch per step) against the reference's optimiser for the seven pose
    scalars, torch.optim.Adam with Frame.py's settings, including a step in which one parameter has no gradient."""
    from gs2d_b200 import tracking
    dev = torch.device("cuda:0")
    want = pose_adam_sequence(torch_pose_adam, dev)
    got = pose_adam_sequence(lambda q, t: tracking.PoseOptimizer(q, t, lr_rot=4e-4, lr_trans=2e-3), dev)
    check_pose_adam(got, want)
    with pytest.raises(RuntimeError):
        tracking.PoseOptimizer(torch.zeros(4), torch.zeros(3), 1e-3, 1e-3)
