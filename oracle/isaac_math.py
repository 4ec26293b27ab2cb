"""Restatement of the un-vendored Isaac Gym tensor helpers used on the hot path.

TEST INFRASTRUCTURE (see oracle/__init__.py).

The reference star-imports these from ``isaacgym.torch_utils`` and
``isaacgymenvs.utils.torch_jit_utils`` (reference fly.py:3-4); neither package
is vendored or version-pinned by the reference (no requirements file), and
neither is installable offline.  The formulas below follow the published
Isaac Gym Preview-4 / IsaacGymEnvs sources as summarised in SURVEY.md
Appendix A; call sites in the reference: ``scale`` fly.py:629, ``unscale``
fly.py:795, ``compute_heading_and_up`` fly.py:789-790, ``compute_rot``
fly.py:792-793, ``quat_conjugate`` fly.py:129, ``to_torch`` fly.py:121-135,
``get_axis_params`` fly.py:125.

All functions are TorchScript so that the reference's own
``@torch.jit.script`` functions can call them when ``ref_harness`` installs
this module under the Isaac Gym module names.  Quaternions are (x, y, z, w).
"""
from typing import List, Tuple

import numpy as np
import torch
from torch import Tensor


def to_torch(x, dtype=torch.float, device="cpu", requires_grad=False):
    return torch.tensor(x, dtype=dtype, device=device, requires_grad=requires_grad)


def get_axis_params(value, axis_idx, x_value=0.0, dtype=float, n_dims=3):
    zs = np.zeros((n_dims,))
    assert axis_idx < n_dims
    zs[axis_idx] = 1.0
    params = np.where(zs == 1.0, value, zs)
    params[0] = x_value
    return list(params.astype(dtype))


@torch.jit.script
def scale(x: Tensor, lower: Tensor, upper: Tensor) -> Tensor:
    return 0.5 * (x + 1.0) * (upper - lower) + lower


@torch.jit.script
def unscale(x: Tensor, lower: Tensor, upper: Tensor) -> Tensor:
    return (2.0 * x - upper - lower) / (upper - lower)


@torch.jit.script
def normalize(x: Tensor, eps: float = 1e-9) -> Tensor:
    return x / x.norm(p=2, dim=-1).clamp(min=eps, max=None).unsqueeze(-1)


@torch.jit.script
def copysign(a: float, b: Tensor) -> Tensor:
    a_t = torch.tensor(a, device=b.device, dtype=torch.float).repeat(b.shape[0])
    return torch.abs(a_t) * torch.sign(b)


@torch.jit.script
def quat_mul(a: Tensor, b: Tensor) -> Tensor:
    assert a.shape == b.shape
    shape = a.shape
    a = a.reshape(-1, 4)
    b = b.reshape(-1, 4)
    x1, y1, z1, w1 = a[:, 0], a[:, 1], a[:, 2], a[:, 3]
    x2, y2, z2, w2 = b[:, 0], b[:, 1], b[:, 2], b[:, 3]
    ww = (z1 + x1) * (x2 + y2)
    yy = (w1 - y1) * (w2 + z2)
    zz = (w1 + y1) * (w2 - z2)
    xx = ww + yy + zz
    qq = 0.5 * (xx + (z1 - x1) * (x2 - y2))
    w = qq - ww + (z1 - y1) * (y2 - z2)
    x = qq - xx + (x1 + w1) * (x2 + w2)
    y = qq - yy + (w1 - x1) * (y2 + z2)
    z = qq - zz + (z1 + y1) * (w2 - x2)
    return torch.stack([x, y, z, w], dim=-1).view(shape)


@torch.jit.script
def quat_conjugate(a: Tensor) -> Tensor:
    shape = a.shape
    a = a.reshape(-1, 4)
    return torch.cat((-a[:, :3], a[:, -1:]), dim=-1).view(shape)


@torch.jit.script
def quat_rotate(q: Tensor, v: Tensor) -> Tensor:
    shape = q.shape
    q_w = q[:, -1]
    q_vec = q[:, :3]
    a = v * (2.0 * q_w ** 2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * torch.bmm(q_vec.view(shape[0], 1, 3), v.view(shape[0], 3, 1)).squeeze(-1) * 2.0
    return a + b + c


@torch.jit.script
def quat_rotate_inverse(q: Tensor, v: Tensor) -> Tensor:
    shape = q.shape
    q_w = q[:, -1]
    q_vec = q[:, :3]
    a = v * (2.0 * q_w ** 2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * torch.bmm(q_vec.view(shape[0], 1, 3), v.view(shape[0], 3, 1)).squeeze(-1) * 2.0
    return a - b + c


@torch.jit.script
def get_basis_vector(q: Tensor, v: Tensor) -> Tensor:
    return quat_rotate(q, v)


@torch.jit.script
def get_euler_xyz(q: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    qx, qy, qz, qw = 0, 1, 2, 3
    sinr_cosp = 2.0 * (q[:, qw] * q[:, qx] + q[:, qy] * q[:, qz])
    cosr_cosp = q[:, qw] * q[:, qw] - q[:, qx] * q[:, qx] - q[:, qy] * q[:, qy] + q[:, qz] * q[:, qz]
    roll = torch.atan2(sinr_cosp, cosr_cosp)

    sinp = 2.0 * (q[:, qw] * q[:, qy] - q[:, qz] * q[:, qx])
    pitch = torch.where(torch.abs(sinp) >= 1, copysign(np.pi / 2.0, sinp), torch.asin(sinp))

    siny_cosp = 2.0 * (q[:, qw] * q[:, qz] + q[:, qx] * q[:, qy])
    cosy_cosp = q[:, qw] * q[:, qw] + q[:, qx] * q[:, qx] - q[:, qy] * q[:, qy] - q[:, qz] * q[:, qz]
    yaw = torch.atan2(siny_cosp, cosy_cosp)

    return roll % (2 * np.pi), pitch % (2 * np.pi), yaw % (2 * np.pi)


@torch.jit.script
def compute_heading_and_up(torso_rotation: Tensor, inv_start_rot: Tensor, to_target: Tensor,
                           vec0: Tensor, vec1: Tensor, up_idx: int
                           ) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    num_envs = torso_rotation.shape[0]
    target_dirs = normalize(to_target)

    torso_quat = quat_mul(torso_rotation, inv_start_rot)
    up_vec = get_basis_vector(torso_quat, vec1).view(num_envs, 3)
    heading_vec = get_basis_vector(torso_quat, vec0).view(num_envs, 3)
    up_proj = up_vec[:, up_idx]
    heading_proj = torch.bmm(heading_vec.view(num_envs, 1, 3), target_dirs.view(num_envs, 3, 1)).view(num_envs)

    return torso_quat, up_proj, heading_proj, up_vec, heading_vec


@torch.jit.script
def compute_rot(torso_quat: Tensor, velocity: Tensor, ang_velocity: Tensor, targets: Tensor,
                torso_positions: Tensor) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor, Tensor]:
    vel_loc = quat_rotate_inverse(torso_quat, velocity)
    angvel_loc = quat_rotate_inverse(torso_quat, ang_velocity)

    roll, pitch, yaw = get_euler_xyz(torso_quat)

    walk_target_angle = torch.atan2(targets[:, 2] - torso_positions[:, 2],
                                    targets[:, 0] - torso_positions[:, 0])
    angle_to_target = walk_target_angle - yaw

    return vel_loc, angvel_loc, roll, pitch, yaw, angle_to_target


EXPORTED = [
    "to_torch", "get_axis_params", "scale", "unscale", "normalize", "copysign", "quat_mul",
    "quat_conjugate", "quat_rotate", "quat_rotate_inverse", "get_basis_vector", "get_euler_xyz",
    "compute_heading_and_up", "compute_rot",
]
