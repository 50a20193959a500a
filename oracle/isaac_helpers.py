"""RESTATEMENT of ``isaacgym.torch_utils`` helpers used on the hot path.  TEST INFRASTRUCTURE.

Provenance: NVIDIA Isaac Gym Preview, ``python/isaacgym/torch_utils.py`` (binary
distribution, un-vendored; the reference pins no version and the module is not
under ``/root/reference``).  Written from the published definitions summarised
in SURVEY.md Appendix A -- NOT copied from a file.  **Parity unpinned** for this
module: nothing in the reference's tree fixes these results.  What constrains
each helper is its call site in the reference:

=====================  =======================================================
helper                 reference call site that constrains it
=====================  =======================================================
compute_heading_and_up isaacgymenvs/tasks/humanoid.py:397-398
compute_rot            isaacgymenvs/tasks/humanoid.py:400-401
normalize_angle        isaacgymenvs/tasks/humanoid.py:403-405
unscale                isaacgymenvs/tasks/humanoid.py:406
quat_conjugate         isaacgymenvs/tasks/humanoid.py:108
torch_rand_float       isaacgymenvs/tasks/humanoid.py:258-259
tensor_clamp           isaacgymenvs/tasks/humanoid.py:261
to_torch               isaacgymenvs/tasks/humanoid.py:106-116
get_axis_params        isaacgymenvs/tasks/humanoid.py:106,179
get_euler_xyz          isaacgymenvs/tasks/gogoro_new.py:392,696
quat_rotate_inverse    isaacgymenvs/tasks/gogoro_new.py:698-699
quat_apply / normalize isaacgymenvs/tasks/gogoro_new.py:798-803
=====================  =======================================================

Cross-check available in-tree (tests/test_oracle_helpers.py, golden vectors made
by tests/golden/make_golden.py): ``quat_mul``, ``quat_conjugate``, ``quat_rotate``
against ``isaacgymenvs/tasks/amp/poselib/poselib/core/rotation3d.py:36-49,81-86,229-235``.

All quaternions are xyzw.  Every function is TorchScript-compatible so that the
reference's lifted ``@torch.jit.script`` functions can resolve these names.
The order of floating-point operations is the upstream one (it matters for the
last ulp); do not "simplify" expressions here.
"""
from typing import List, Tuple

import numpy as np
import torch
from torch import Tensor


def to_torch(x, dtype=torch.float, device='cpu', requires_grad=False):
    return torch.tensor(x, dtype=dtype, device=device, requires_grad=requires_grad)


def get_axis_params(value, axis_idx, x_value=0., dtype=float, n_dims=3):
    """[0]*n_dims with [axis_idx] = value, then [0] = x_value (so (1., 2) -> [0, 0, 1])."""
    out = [0.] * n_dims
    out[axis_idx] = value
    out[0] = x_value  # unconditional upstream; the reference only passes axis_idx 1 or 2
    return [dtype(v) for v in out]


def torch_rand_float(lower: float, upper: float, shape: Tuple[int, int], device: str):
    return (upper - lower) * torch.rand(shape[0], shape[1], device=device) + lower


def rand_float_from_unit(lower: float, upper: float, unit: Tensor) -> Tensor:
    """torch_rand_float with the U(0,1) draw supplied (same arithmetic: (hi-lo)*u + lo)."""
    return (upper - lower) * unit + lower


def tensor_clamp(t: Tensor, min_t: Tensor, max_t: Tensor) -> Tensor:
    return torch.max(torch.min(t, max_t), min_t)


def scale(x: Tensor, lower: Tensor, upper: Tensor) -> Tensor:
    return 0.5 * (x + 1.0) * (upper - lower) + lower


def unscale(x: Tensor, lower: Tensor, upper: Tensor) -> Tensor:
    return (2.0 * x - upper - lower) / (upper - lower)


def normalize(x: Tensor, eps: float = 1e-9) -> Tensor:
    return x / x.norm(p=2, dim=-1).clamp(min=eps, max=None).unsqueeze(-1)


def normalize_angle(x: Tensor) -> Tensor:
    return torch.atan2(torch.sin(x), torch.cos(x))


def copysign(a: float, b: Tensor) -> Tensor:
    mag = torch.zeros_like(b) + a
    return torch.abs(mag) * torch.sign(b)


def quat_conjugate(a: Tensor) -> Tensor:
    shape = a.shape
    flat = a.reshape(-1, 4)
    return torch.cat((-flat[:, :3], flat[:, -1:]), dim=-1).view(shape)


def quat_mul(a: Tensor, b: Tensor) -> Tensor:
    """Hamilton product a (x) b in the 9-multiply factorised form (Appendix A)."""
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


def quat_apply(a: Tensor, b: Tensor) -> Tensor:
    shape = b.shape
    a = a.reshape(-1, 4)
    b = b.reshape(-1, 3)
    xyz = a[:, :3]
    t = xyz.cross(b, dim=-1) * 2
    return (b + a[:, 3:] * t + xyz.cross(t, dim=-1)).view(shape)


def quat_rotate(q: Tensor, v: Tensor) -> Tensor:
    shape = q.shape
    q_w = q[:, -1]
    q_vec = q[:, :3]
    a = v * (2.0 * q_w ** 2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * torch.bmm(q_vec.view(shape[0], 1, 3), v.view(shape[0], 3, 1)).squeeze(-1) * 2.0
    return a + b + c


def quat_rotate_inverse(q: Tensor, v: Tensor) -> Tensor:
    shape = q.shape
    q_w = q[:, -1]
    q_vec = q[:, :3]
    a = v * (2.0 * q_w ** 2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * torch.bmm(q_vec.view(shape[0], 1, 3), v.view(shape[0], 3, 1)).squeeze(-1) * 2.0
    return a - b + c


def get_basis_vector(q: Tensor, v: Tensor) -> Tensor:
    return quat_rotate(q, v)


def get_euler_xyz(q: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    """Each angle is returned ``% (2*pi)`` (torch floor-mod), i.e. in [0, 2*pi]."""
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


def compute_rot(torso_quat: Tensor, velocity: Tensor, ang_velocity: Tensor, targets: Tensor,
                torso_positions: Tensor) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor, Tensor]:
    vel_loc = quat_rotate_inverse(torso_quat, velocity)
    angvel_loc = quat_rotate_inverse(torso_quat, ang_velocity)

    roll, pitch, yaw = get_euler_xyz(torso_quat)

    # indices 2 and 0 (z and x) -- upstream definition, kept as is
    walk_target_angle = torch.atan2(targets[:, 2] - torso_positions[:, 2],
                                    targets[:, 0] - torso_positions[:, 0])
    angle_to_target = walk_target_angle - yaw

    return vel_loc, angvel_loc, roll, pitch, yaw, angle_to_target


HELPER_NAMES: List[str] = [
    "to_torch", "get_axis_params", "torch_rand_float", "tensor_clamp", "scale", "unscale",
    "normalize", "normalize_angle", "copysign", "quat_conjugate", "quat_mul", "quat_apply",
    "quat_rotate", "quat_rotate_inverse", "get_basis_vector", "get_euler_xyz",
    "compute_heading_and_up", "compute_rot",
]
