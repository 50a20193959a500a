"""CPU restatement of the Anymal hot path (second half).  TEST INFRASTRUCTURE.

Follows ``/root/reference/isaacgymenvs/tasks/anymal.py`` (this fork's copy: 61 observations through
``compute_anymal_observations_V2`` -- raw local velocities, joint torques, per-body contact norms, no action
block; file:line cited per statement; constants from ``cfg/task/Anymal.yaml``).  Anymal is the canonical
``targets = action_scale * actions + default_dof_pos`` PD-target write (:229-232) and the contact-force
termination (:362-366).  The JIT-level functions are pinned bit-for-bit against the reference's own lifted
TorchScript functions (tests/test_oracle_vs_reference.py, tests/golden/anymal_jit.npz); helpers from
``oracle.isaac_helpers`` are a restated un-vendored module (parity unpinned there).

Random draws are inputs (U(0,1) "unit" tensors); the affine map is upstream ``torch_rand_float``
(``(hi - lo) * rand + lo``).
"""
from dataclasses import dataclass, field
from typing import Dict, List, Tuple

import torch
from torch import Tensor

from . import isaac_helpers as ih

NUM_DOF = 12          # :84
NUM_OBS = 61          # :83
NUM_BODIES = 13       # base + 4 x (HIP, THIGH, SHANK) with collapse_fixed_joints (:192,206-213): 61 = 48 + 13
RESET_CMD_COLS = 3


@dataclass
class AnymalConsts:
    """anymal.py:46-97, cfg/task/Anymal.yaml:10-64."""
    lin_vel_scale: float = 2.0
    ang_vel_scale: float = 0.25
    dof_pos_scale: float = 1.0
    dof_vel_scale: float = 0.05
    action_scale: float = 0.5
    rew_lin_vel_xy: float = 1.0
    rew_ang_vel_z: float = 0.5
    rew_torque: float = -0.000025
    command_x_range: List[float] = field(default_factory=lambda: [-2.0, 2.0])
    command_y_range: List[float] = field(default_factory=lambda: [-1.0, 1.0])
    command_yaw_range: List[float] = field(default_factory=lambda: [-1.0, 1.0])
    dt: float = 0.02
    episode_length_s: float = 50.0
    base_index: int = 0
    knee_indices: Tuple[int, ...] = (2, 5, 8, 11)      # the four THIGH bodies
    clip_actions: float = 1.0
    clip_obs: float = 5.0

    @property
    def max_episode_length(self) -> int:
        return int(self.episode_length_s / self.dt + 0.5)                       # :92

    def rew_scales(self) -> Dict[str, float]:
        """:52-55 scaled by dt at :96-97 (python doubles)."""
        return {"lin_vel_xy": self.rew_lin_vel_xy * self.dt, "ang_vel_z": self.rew_ang_vel_z * self.dt,
                "torque": self.rew_torque * self.dt}

    @staticmethod
    def from_cfg(cfg: dict) -> "AnymalConsts":
        e, l, c = cfg["env"], cfg["env"]["learn"], cfg["env"]["control"]
        r = e["randomCommandVelocityRanges"]
        return AnymalConsts(
            lin_vel_scale=l["linearVelocityScale"], ang_vel_scale=l["angularVelocityScale"],
            dof_pos_scale=l["dofPositionScale"], dof_vel_scale=l["dofVelocityScale"], action_scale=c["actionScale"],
            rew_lin_vel_xy=l["linearVelocityXYRewardScale"], rew_ang_vel_z=l["angularVelocityZRewardScale"],
            rew_torque=l["torqueRewardScale"], command_x_range=list(r["linear_x"]), command_y_range=list(r["linear_y"]),
            command_yaw_range=list(r["yaw"]), dt=cfg["sim"]["dt"], episode_length_s=l["episodeLength_s"],
            base_index=e.get("baseIndex", 0), knee_indices=tuple(e.get("kneeIndices", (2, 5, 8, 11))),
            clip_actions=e.get("clipActions", float("inf")), clip_obs=e.get("clipObservations", float("inf")))


# --------------------------------------------------------------------------------------
# JIT-level functions
# --------------------------------------------------------------------------------------

def anymal_reward(root_states: Tensor, commands: Tensor, torques: Tensor, contact_forces: Tensor, knee_indices: Tensor,
                  episode_lengths: Tensor, rew_scales: Dict[str, float], base_index: int, max_episode_length: int
                  ) -> Tuple[Tensor, Tensor]:
    """anymal.py:323-372 ``compute_anymal_reward``."""
    base_quat = root_states[:, 3:7]
    base_lin_vel = ih.quat_rotate_inverse(base_quat, root_states[:, 7:10])                      # :342
    base_ang_vel = ih.quat_rotate_inverse(base_quat, root_states[:, 10:13])                     # :343
    lin_vel_error = torch.sum(torch.square(commands[:, :2] - base_lin_vel[:, :2]), dim=1)       # :346
    ang_vel_error = torch.square(commands[:, 2] - base_ang_vel[:, 2])                           # :348
    rew_lin_vel_xy = torch.exp(-lin_vel_error / 0.25) * rew_scales["lin_vel_xy"]                # :349
    rew_ang_vel_z = torch.exp(-ang_vel_error / 0.25) * rew_scales["ang_vel_z"]                  # :350
    rew_torque = torch.sum(torch.square(torques), dim=1) * rew_scales["torque"]                 # :352
    total_reward = rew_lin_vel_xy + rew_ang_vel_z + rew_torque                                  # :356
    total_reward = torch.clip(total_reward, 0., None)                                           # :357
    reset = torch.norm(contact_forces[:, base_index, :], dim=1) > 1.                            # :359
    reset = reset | torch.any(torch.norm(contact_forces[:, knee_indices, :], dim=2) > 1., dim=1)   # :360
    time_out = episode_lengths >= max_episode_length - 1                                        # :361
    reset = reset | time_out                                                                    # :362
    return total_reward.detach(), reset


def anymal_observations(root_states: Tensor, commands: Tensor, torques: Tensor, contact_forces: Tensor, dof_pos: Tensor,
                        default_dof_pos: Tensor, dof_vel: Tensor, gravity_vec: Tensor, dof_pos_scale: float,
                        dof_vel_scale: float) -> Tensor:
    """anymal.py:393-431 ``compute_anymal_observations_V2`` (actions and the two velocity scales are unused there)."""
    base_quat = root_states[:, 3:7]
    base_lin_vel = ih.quat_rotate_inverse(base_quat, root_states[:, 7:10])                      # :411
    base_ang_vel = ih.quat_rotate_inverse(base_quat, root_states[:, 10:13])                     # :412
    projected_gravity = ih.quat_rotate(base_quat, gravity_vec)                                  # :413
    dof_pos_scaled = (dof_pos - default_dof_pos) * dof_pos_scale                                # :414
    contact = torch.norm(contact_forces, dim=-1)                                                # :417
    return torch.cat((base_lin_vel, base_ang_vel, torques, projected_gravity, contact, dof_pos_scaled,
                      dof_vel * dof_vel_scale, commands), dim=-1)                               # :419-427


def anymal_wm_reward(base_lin_vel: Tensor, base_ang_vel: Tensor, torques: Tensor, commands: Tensor,
                     rew_scales: Dict[str, float]) -> Tensor:
    """anymal.py:375-396, the fork's second ``compute_reward`` over [B,T,*] tensors."""
    lin_vel_error = torch.sum(torch.square(commands[..., :2] - base_lin_vel[..., :2]), dim=2)
    ang_vel_error = torch.square(commands[..., 2] - base_ang_vel[..., 2])
    rew_lin_vel_xy = torch.exp(-lin_vel_error / 0.25) * rew_scales["lin_vel_xy"]
    rew_ang_vel_z = torch.exp(-ang_vel_error / 0.25) * rew_scales["ang_vel_z"]
    rew_torque = torch.sum(torch.square(torques), dim=2) * rew_scales["torque"]
    return torch.clip(rew_lin_vel_xy + rew_ang_vel_z + rew_torque, 0., None)


def _rand_float(lo: float, hi: float, unit: Tensor) -> Tensor:
    return (hi - lo) * unit + lo                                                                # torch_rand_float


class AnymalOracle:
    """State + method bodies of ``Anymal`` (anymal.py:40-147, 229-321) without isaacgym, fake-sim semantics."""

    def __init__(self, consts: AnymalConsts, root_states: Tensor, dof_state: Tensor, torques: Tensor, contact_forces: Tensor,
                 default_dof_pos_row: Tensor, base_init_state: Tensor):
        self.c = consts
        n, d = root_states.shape[0], default_dof_pos_row.shape[0]
        self.num_envs, self.num_dof = n, d
        self.root_states, self.dof_state = root_states, dof_state
        self.dof_pos = dof_state.view(n, d, 2)[..., 0]                                          # :122
        self.dof_vel = dof_state.view(n, d, 2)[..., 1]                                          # :123
        self.contact_forces = contact_forces.view(n, -1, 3)                                     # :124
        self.torques = torques.view(n, d)                                                       # :125
        self.commands = torch.zeros(n, 3)                                                       # :127
        self.default_dof_pos = default_dof_pos_row.repeat(n, 1)                                 # :131-136
        self.initial_root_states = base_init_state.repeat(n, 1)                                 # :140-141
        self.gravity_vec = torch.tensor([0.0, 0.0, -1.0]).repeat(n, 1)                          # :142
        self.actions = torch.zeros(n, d)                                                        # :143
        self.knee_indices = torch.tensor(consts.knee_indices, dtype=torch.long)
        self.max_episode_length = consts.max_episode_length
        self.rew_scales = consts.rew_scales()
        self.obs_buf = torch.zeros(n, NUM_OBS)
        self.rew_buf = torch.zeros(n)
        self.reset_buf = torch.ones(n, dtype=torch.long)
        self.progress_buf = torch.zeros(n, dtype=torch.long)
        self.timeout_buf = torch.zeros(n, dtype=torch.long)
        self.targets = torch.zeros(n, d)

    # :229-232 (+ clamp vec_task.py:327)
    def pre_physics_step(self, actions: Tensor, clip: bool = False):
        if clip:
            actions = torch.clamp(actions, -self.c.clip_actions, self.c.clip_actions)
        self.actions = actions.clone()                                                          # :230
        self.targets = self.c.action_scale * self.actions + self.default_dof_pos               # :231

    # :247-266
    def compute_observations(self):
        self.obs_buf[:] = anymal_observations(self.root_states, self.commands, self.torques, self.contact_forces,
                                              self.dof_pos, self.default_dof_pos, self.dof_vel, self.gravity_vec,
                                              self.c.dof_pos_scale, self.c.dof_vel_scale)

    # :244-259 (compute_reward)
    def compute_reward(self):
        self.rew_buf[:], self.reset_buf[:] = anymal_reward(
            self.root_states, self.commands, self.torques, self.contact_forces, self.knee_indices, self.progress_buf,
            self.rew_scales, self.c.base_index, self.max_episode_length)

    # :283-311; unit_pos/unit_vel [K,D], unit_cmd [K,3] (x, y, yaw in draw order :305-307)
    def reset_idx(self, env_ids: Tensor, unit_pos: Tensor, unit_vel: Tensor, unit_cmd: Tensor):
        c = self.c
        positions_offset = _rand_float(0.5, 1.5, unit_pos)                                      # :288
        velocities = _rand_float(-0.1, 0.1, unit_vel)                                           # :289
        self.dof_pos[env_ids] = self.default_dof_pos[env_ids] * positions_offset                # :291
        self.dof_vel[env_ids] = velocities                                                      # :292
        self.root_states[env_ids] = self.initial_root_states[env_ids]                           # :296-298 (fake sim row copy)
        self.commands[env_ids, 0] = _rand_float(c.command_x_range[0], c.command_x_range[1], unit_cmd[:, 0])       # :305
        self.commands[env_ids, 1] = _rand_float(c.command_y_range[0], c.command_y_range[1], unit_cmd[:, 1])       # :306
        self.commands[env_ids, 2] = _rand_float(c.command_yaw_range[0], c.command_yaw_range[1], unit_cmd[:, 2])   # :307
        self.progress_buf[env_ids] = 0                                                          # :309
        self.reset_buf[env_ids] = 1                                                             # :310 (sic; overwritten by compute_reward)

    # :234-242
    def post_physics_step(self, unit_pos: Tensor = None, unit_vel: Tensor = None, unit_cmd: Tensor = None):
        self.progress_buf += 1                                                                  # :235
        env_ids = self.reset_buf.nonzero(as_tuple=False).squeeze(-1)                            # :237
        if len(env_ids) > 0:
            k = len(env_ids)
            self.reset_idx(env_ids, unit_pos[:k], unit_vel[:k], unit_cmd[:k])                   # :239
        self.compute_observations()                                                             # :241
        self.compute_reward()                                                                   # :242
        self.timeout_buf = (self.progress_buf >= self.max_episode_length - 1) & (self.reset_buf != 0)   # vec_task.py:345
        return env_ids
