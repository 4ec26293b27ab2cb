"""Independent ground truth for the restated Isaac Gym helpers (oracle/isaac_math.py).

The reference star-imports `quat_mul`, `quat_rotate`, `quat_rotate_inverse`, `get_euler_xyz`, `compute_heading_and_up`,
`compute_rot`, `scale`, `unscale`, `normalize` from `isaacgym.torch_utils` / `isaacgymenvs.utils.torch_jit_utils`
(reference fly.py:3-4; call sites fly.py:129, 629, 789-795), packages that are neither vendored nor installable here.
The golden obs vectors were generated THROUGH the restatement, so nothing reference-held pins it; this
file pins it against `scipy.spatial.transform.Rotation` on random unit quaternions and against the closed forms of
SURVEY.md §8c (identity quaternion => up_proj = heading_proj = 1, roll = pitch = yaw = 0).
"""
import math

import numpy as np
import pytest
import torch
from scipy.spatial.transform import Rotation as R

from oracle import isaac_math as im

N = 4096
TWO_PI = 2.0 * math.pi


def _unit_quats(seed, n=N):
    g = torch.Generator().manual_seed(seed)
    q = torch.randn(n, 4, generator=g, dtype=torch.float64)
    return (q / q.norm(dim=-1, keepdim=True))


def _ang_diff(a, b):
    """smallest distance between angles on the circle"""
    d = (a - b) % TWO_PI
    return np.minimum(d, TWO_PI - d)


def test_quat_mul_matches_scipy_composition():
    a, b = _unit_quats(1), _unit_quats(2)
    got = im.quat_mul(a.float(), b.float()).double().numpy()
    want = (R.from_quat(a.numpy()) * R.from_quat(b.numpy())).as_quat()      # scipy: (x, y, z, w), Hamilton product
    sign = np.sign((got * want).sum(-1, keepdims=True))                        # q and -q are the same rotation
    np.testing.assert_allclose(got, want * sign, atol=2e-6)
    # identity on either side, and q * conj(q) = 1
    one = torch.tensor([[0.0, 0.0, 0.0, 1.0]]).repeat(N, 1)
    np.testing.assert_allclose(im.quat_mul(a.float(), one).numpy(), a.float().numpy(), atol=1e-6)
    np.testing.assert_allclose(im.quat_mul(one, a.float()).numpy(), a.float().numpy(), atol=1e-6)
    np.testing.assert_allclose(im.quat_mul(a.float(), im.quat_conjugate(a.float())).numpy(), one.numpy(), atol=1e-6)


def test_quat_rotate_and_inverse_match_scipy():
    q = _unit_quats(3)
    g = torch.Generator().manual_seed(4)
    v = torch.randn(N, 3, generator=g, dtype=torch.float64) * 3.0
    rot = R.from_quat(q.numpy())
    np.testing.assert_allclose(im.quat_rotate(q.float(), v.float()).double().numpy(), rot.apply(v.numpy()), atol=1e-5)
    np.testing.assert_allclose(im.quat_rotate_inverse(q.float(), v.float()).double().numpy(),
                               rot.inv().apply(v.numpy()), atol=1e-5)
    back = im.quat_rotate_inverse(q.float(), im.quat_rotate(q.float(), v.float()))
    np.testing.assert_allclose(back.numpy(), v.float().numpy(), atol=2e-5)
    np.testing.assert_allclose(im.get_basis_vector(q.float(), v.float()).numpy(),
                               im.quat_rotate(q.float(), v.float()).numpy())


def test_get_euler_xyz_matches_scipy_extrinsic_xyz():
    q = _unit_quats(5)
    sinp = 2.0 * (q[:, 3] * q[:, 1] - q[:, 2] * q[:, 0])
    keep = sinp.abs() < 0.999                                                   # away from the gimbal-lock pole
    q = q[keep]
    roll, pitch, yaw = im.get_euler_xyz(q.float())
    for t in (roll, pitch, yaw):                                                # `% 2 pi`: every angle lands in [0, 2 pi)
        assert float(t.min()) >= 0.0 and float(t.max()) < TWO_PI + 1e-6
    want = R.from_quat(q.numpy()).as_euler("xyz")                               # extrinsic x-y-z == yaw(z) pitch(y) roll(x)
    assert _ang_diff(roll.double().numpy(), want[:, 0]).max() < 5e-4
    assert _ang_diff(pitch.double().numpy(), want[:, 1]).max() < 5e-4
    assert _ang_diff(yaw.double().numpy(), want[:, 2]).max() < 5e-4
    # and the angles rebuild the rotation
    rebuilt = R.from_euler("xyz", np.stack([roll.numpy(), pitch.numpy(), yaw.numpy()], -1).astype(np.float64))
    err = (rebuilt * R.from_quat(q.numpy()).inv()).magnitude()
    assert err.max() < 1e-3


def test_get_euler_xyz_pole_is_clamped():
    s = math.sqrt(0.5)
    q = torch.tensor([[0.0, s, 0.0, s], [0.0, -s, 0.0, s]])                    # pitch = +-90 degrees: |sinp| >= 1
    _, pitch, _ = im.get_euler_xyz(q)
    np.testing.assert_allclose(pitch.numpy(), [math.pi / 2, (-math.pi / 2) % TWO_PI], atol=1e-3)


def test_closed_forms_identity_quaternion():
    n = 8
    ident = torch.tensor([[0.0, 0.0, 0.0, 1.0]]).repeat(n, 1)
    pos = torch.zeros(n, 3)
    targets = torch.tensor([[1000.0, 0.0, 0.0]]).repeat(n, 1)
    b0 = torch.tensor([[1.0, 0.0, 0.0]]).repeat(n, 1)
    b1 = torch.tensor([[0.0, 0.0, 1.0]]).repeat(n, 1)
    tq, up_proj, heading_proj, up_vec, heading_vec = im.compute_heading_and_up(ident, ident, targets - pos, b0, b1, 2)
    assert torch.equal(tq, ident)
    assert torch.equal(up_proj, torch.ones(n)) and torch.equal(heading_proj, torch.ones(n))
    assert torch.equal(up_vec, b1) and torch.equal(heading_vec, b0)
    vel = torch.randn(n, 3)
    ang = torch.randn(n, 3)
    vel_loc, angvel_loc, roll, pitch, yaw, angle = im.compute_rot(tq, vel, ang, targets, pos)
    assert torch.equal(vel_loc, vel) and torch.equal(angvel_loc, ang)
    for t in (roll, pitch, yaw, angle):
        assert torch.equal(t, torch.zeros(n))


def test_heading_and_up_against_scipy():
    q = _unit_quats(6)
    g = torch.Generator().manual_seed(7)
    pos = torch.randn(N, 3, generator=g, dtype=torch.float64) * 4.0
    targets = torch.tensor([[1000.0, 0.0, 0.0]], dtype=torch.float64).repeat(N, 1)
    to_target = targets - pos
    to_target[:, 2] = 0.0                                                       # fly.py:787-788
    ident = torch.tensor([[0.0, 0.0, 0.0, 1.0]]).repeat(N, 1)
    b0 = torch.tensor([[1.0, 0.0, 0.0]]).repeat(N, 1)
    b1 = torch.tensor([[0.0, 0.0, 1.0]]).repeat(N, 1)
    tq, up_proj, heading_proj, up_vec, heading_vec = im.compute_heading_and_up(
        q.float(), ident, to_target.float(), b0, b1, 2)
    rot = R.from_quat(q.numpy())
    np.testing.assert_allclose(up_vec.double().numpy(), rot.apply([0.0, 0.0, 1.0]), atol=2e-6)
    np.testing.assert_allclose(heading_vec.double().numpy(), rot.apply([1.0, 0.0, 0.0]), atol=2e-6)
    dirs = (to_target / to_target.norm(dim=-1, keepdim=True)).numpy()
    np.testing.assert_allclose(heading_proj.double().numpy(), (rot.apply([1.0, 0.0, 0.0]) * dirs).sum(-1), atol=3e-6)
    np.testing.assert_allclose(up_proj.double().numpy(), rot.apply([0.0, 0.0, 1.0])[:, 2], atol=2e-6)
    # compute_rot: world -> body velocities; the walk-target angle uses z, not y (as upstream)
    vel = torch.randn(N, 3, generator=g, dtype=torch.float64)
    vel_loc, _, _, _, yaw, angle = im.compute_rot(tq, vel.float(), vel.float(), targets.float(), pos.float())
    np.testing.assert_allclose(vel_loc.double().numpy(), rot.inv().apply(vel.numpy()), atol=1e-5)
    want_angle = np.arctan2((targets[:, 2] - pos[:, 2]).numpy(), (targets[:, 0] - pos[:, 0]).numpy()) - yaw.double().numpy()
    np.testing.assert_allclose(angle.double().numpy(), want_angle, atol=1e-5)


def test_scale_unscale_normalize():
    g = torch.Generator().manual_seed(8)
    lo = -torch.rand(18, generator=g) * 3 - 0.1
    hi = torch.rand(18, generator=g) * 3 + 0.1
    x = torch.rand(64, 18, generator=g) * 2 - 1
    s = im.scale(x, lo, hi)
    assert bool(((s >= lo - 1e-6) & (s <= hi + 1e-6)).all())
    np.testing.assert_allclose(im.unscale(s, lo, hi).numpy(), x.numpy(), atol=2e-6)
    np.testing.assert_allclose(im.scale(torch.tensor([-1.0, 1.0, 0.0]), torch.tensor(-2.0), torch.tensor(4.0)).numpy(),
                               [-2.0, 4.0, 1.0])
    v = torch.randn(32, 3, generator=g)
    np.testing.assert_allclose(im.normalize(v).norm(dim=-1).numpy(), np.ones(32), atol=1e-6)
    assert torch.equal(im.normalize(torch.zeros(2, 3)), torch.zeros(2, 3))     # clamp(min=1e-9): no NaN at zero
    assert im.get_axis_params(1.0, 2) == [0.0, 0.0, 1.0]                       # fly.py:125
    np.testing.assert_allclose(im.quat_conjugate(torch.tensor([[1.0, 2.0, 3.0, 4.0]])).numpy(), [[-1, -2, -3, 4]])
