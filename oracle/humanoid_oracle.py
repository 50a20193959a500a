"""CPU restatement of the Humanoid hot path.  TEST INFRASTRUCTURE (checker + timed CPU baseline).

Follows ``/root/reference/isaacgymenvs/tasks/humanoid.py`` (file:line cited per
function).  The JIT-level functions are pinned bit-for-bit against the
reference's own lifted TorchScript functions (tests/golden/humanoid_*.npz,
tests/test_oracle_vs_reference.py).  The helpers they call come from
``oracle.isaac_helpers`` (restated upstream module: parity unpinned there).

Everything is torch fp32 on CPU, same operators in the same order as the
reference so that the results are identical, not merely close.
"""
from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import torch
from torch import Tensor

from . import isaac_helpers as ih

NUM_DOF = 21          # humanoid.py:68 (numActions = 21 = DOFs of nv_humanoid)
NUM_OBS = 110         # humanoid.py:67
NUM_SENSOR = 12       # humanoid.py:82-83  (2 foot sensors x 6)


@dataclass
class HumanoidConsts:
    """Scalars the task reads from cfg (humanoid.py:46-68,115; cfg/task/Humanoid.yaml:7-29,43)."""
    dof_vel_scale: float = 0.1
    angular_velocity_scale: float = 0.25
    contact_force_scale: float = 0.01
    power_scale: float = 1.0
    heading_weight: float = 0.5
    up_weight: float = 0.1
    actions_cost_scale: float = 0.01
    energy_cost_scale: float = 0.05
    joints_at_limit_cost_scale: float = 0.25
    death_cost: float = -1.0
    termination_height: float = 0.8
    max_episode_length: float = 1000
    dt: float = 0.0166
    clip_actions: float = 1.0
    clip_obs: float = float("inf")

    @staticmethod
    def from_cfg(cfg: dict) -> "HumanoidConsts":
        e = cfg["env"]
        return HumanoidConsts(
            dof_vel_scale=e["dofVelocityScale"],
            angular_velocity_scale=e.get("angularVelocityScale", 0.1),   # humanoid.py:49
            contact_force_scale=e["contactForceScale"],
            power_scale=e["powerScale"],
            heading_weight=e["headingWeight"],
            up_weight=e["upWeight"],
            actions_cost_scale=e["actionsCost"],
            energy_cost_scale=e["energyCost"],
            joints_at_limit_cost_scale=e["jointsAtLimitCost"],
            death_cost=e["deathCost"],
            termination_height=e["terminationHeight"],
            max_episode_length=e["episodeLength"],
            dt=cfg["sim"]["dt"],
            clip_actions=e.get("clipActions", float("inf")),            # vec_task.py:108
            clip_obs=e.get("clipObservations", float("inf")),           # vec_task.py:107
        )


# --------------------------------------------------------------------------------------
# JIT-level functions
# --------------------------------------------------------------------------------------

def humanoid_observations(root_states: Tensor, targets: Tensor, potentials: Tensor,
                          inv_start_rot: Tensor, dof_pos: Tensor, dof_vel: Tensor,
                          dof_force: Tensor, dof_limits_lower: Tensor, dof_limits_upper: Tensor,
                          dof_vel_scale: float, sensor_force_torques: Tensor, actions: Tensor,
                          dt: float, contact_force_scale: float, angular_velocity_scale: float,
                          basis_vec0: Tensor, basis_vec1: Tensor
                          ) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """humanoid.py:379-417 ``compute_humanoid_observations`` (obs_buf argument dropped: unused there).

    Returns (obs[N,110], potentials', prev_potentials', up_vec, heading_vec).
    """
    pos = root_states[:, 0:3]
    quat = root_states[:, 3:7]
    lin = root_states[:, 7:10]
    ang = root_states[:, 10:13]

    # humanoid.py:391-395
    to_target = targets - pos
    to_target[:, 2] = 0
    prev_potentials_new = potentials.clone()
    potentials_new = -torch.norm(to_target, p=2, dim=-1) / dt

    # humanoid.py:397-401
    torso_quat, up_proj, heading_proj, up_vec, heading_vec = ih.compute_heading_and_up(
        quat, inv_start_rot, to_target, basis_vec0, basis_vec1, 2)
    vel_loc, angvel_loc, roll, pitch, yaw, angle_to_target = ih.compute_rot(
        torso_quat, lin, ang, targets, pos)

    # humanoid.py:403-406
    roll_n = ih.normalize_angle(roll)
    yaw_n = ih.normalize_angle(yaw)
    att_n = ih.normalize_angle(angle_to_target)
    dof_pos_scaled = ih.unscale(dof_pos, dof_limits_lower, dof_limits_upper)

    # humanoid.py:409-414 -- column layout of the 110-wide row
    pieces = [
        pos[:, 2].view(-1, 1),                                  # 0
        vel_loc,                                                # 1:4
        angvel_loc * angular_velocity_scale,                    # 4:7
        yaw_n.unsqueeze(-1),                                    # 7
        roll_n.unsqueeze(-1),                                   # 8
        att_n.unsqueeze(-1),                                    # 9
        up_proj.unsqueeze(-1),                                  # 10
        heading_proj.unsqueeze(-1),                             # 11
        dof_pos_scaled,                                         # 12:33
        dof_vel * dof_vel_scale,                                # 33:54
        dof_force * contact_force_scale,                        # 54:75
        sensor_force_torques.view(-1, 12) * contact_force_scale,  # 75:87
        actions,                                                # 87:108
        potentials_new.unsqueeze(-1) / 60000,                   # 108
        prev_potentials_new.unsqueeze(-1) / 60000,              # 109
    ]
    obs = torch.cat(pieces, dim=-1)
    return obs, potentials_new, prev_potentials_new, up_vec, heading_vec


def humanoid_reward_terms(obs_buf: Tensor, actions: Tensor, up_weight: float, heading_weight: float,
                          potentials: Tensor, prev_potentials: Tensor,
                          joints_at_limit_cost_scale: float, max_motor_effort: float,
                          motor_efforts: Tensor) -> Dict[str, Tensor]:
    """All reward terms of humanoid.py:344-363, including the ones the fork leaves dead."""
    hp = obs_buf[:, 11]
    heading_weight_tensor = torch.ones_like(hp) * heading_weight
    heading_reward = torch.where(hp > 0.8, heading_weight_tensor, heading_weight * hp / 0.8)   # :345-346

    up_reward = torch.zeros_like(heading_reward)
    up_reward = torch.where(obs_buf[:, 10] > 0.93, up_reward + up_weight, up_reward)           # :349-350

    actions_cost = torch.sum(actions ** 2, dim=-1)                                             # :352

    motor_effort_ratio = motor_efforts / max_motor_effort                                      # :355
    p_abs = torch.abs(obs_buf[:, 12:33])
    scaled_cost = joints_at_limit_cost_scale * (p_abs - 0.98) / 0.02                           # :356
    dof_at_limit_cost = torch.sum((p_abs > 0.98) * scaled_cost * motor_effort_ratio.unsqueeze(0), dim=-1)  # :357
    electricity_cost = torch.sum(torch.abs(actions * obs_buf[:, 33:54]) * motor_effort_ratio.unsqueeze(0), dim=-1)  # :359

    alive_reward = torch.ones_like(potentials) * 2.0                                           # :362
    progress_reward = potentials - prev_potentials                                             # :363
    return dict(heading_reward=heading_reward, up_reward=up_reward, actions_cost=actions_cost,
                dof_at_limit_cost=dof_at_limit_cost, electricity_cost=electricity_cost,
                alive_reward=alive_reward, progress_reward=progress_reward)


def humanoid_reward(obs_buf: Tensor, reset_buf: Tensor, progress_buf: Tensor, actions: Tensor,
                    up_weight: float, heading_weight: float, potentials: Tensor, prev_potentials: Tensor,
                    actions_cost_scale: float, energy_cost_scale: float, joints_at_limit_cost_scale: float,
                    max_motor_effort: float, motor_efforts: Tensor, termination_height: float,
                    death_cost: float, max_episode_length: float) -> Tuple[Tensor, Tensor]:
    """humanoid.py:323-376 ``compute_humanoid_reward``: live formula = alive(2.0) + up."""
    t = humanoid_reward_terms(obs_buf, actions, up_weight, heading_weight, potentials, prev_potentials,
                              joints_at_limit_cost_scale, max_motor_effort, motor_efforts)
    total_reward = t["alive_reward"] + t["up_reward"]                                          # :365
    fallen = obs_buf[:, 0] < termination_height
    total_reward = torch.where(fallen, torch.ones_like(total_reward) * death_cost, total_reward)  # :370
    reset = torch.where(fallen, torch.ones_like(reset_buf), reset_buf)                         # :373
    reset = torch.where(progress_buf >= max_episode_length - 1, torch.ones_like(reset_buf), reset)  # :374
    return total_reward, reset


def humanoid_wm_reward(obs_buf: Tensor, actions: Tensor, up_weight: float, heading_weight: float,
                       actions_cost_scale: float, energy_cost_scale: float,
                       joints_at_limit_cost_scale: float, max_motor_effort: float, motor_efforts: Tensor,
                       termination_height: float, death_cost: float) -> Tensor:
    """humanoid.py:425-472 ``compute_reward`` over obs[B,T,110]: progress + alive + up + heading."""
    potentials = obs_buf[:, :, 108]
    prev_potentials = obs_buf[:, :, 109]
    hp = obs_buf[:, :, 11]
    heading_weight_tensor = torch.ones_like(hp) * heading_weight
    heading_reward = torch.where(hp > 0.8, heading_weight_tensor, heading_weight * hp / 0.8)
    up_reward = torch.zeros_like(heading_reward)
    up_reward = torch.where(obs_buf[:, :, 10] > 0.93, up_reward + up_weight, up_reward)
    alive_reward = torch.ones_like(potentials) * 2.0
    progress_reward = potentials - prev_potentials
    total_reward = progress_reward + alive_reward + up_reward + heading_reward                 # :467
    total_reward = torch.where(obs_buf[:, :, 0] < termination_height,
                               torch.ones_like(total_reward) * death_cost, total_reward)       # :470
    return total_reward


_SCRIPTED = {}


def scripted(name: str):
    """TorchScript-compiled variant (legacy executor, like vec_task.py:198-200) for the timed baseline."""
    if name not in _SCRIPTED:
        torch._C._jit_set_profiling_mode(False)
        torch._C._jit_set_profiling_executor(False)
        _SCRIPTED[name] = torch.jit.script(globals()[name])
    return _SCRIPTED[name]


# --------------------------------------------------------------------------------------
# Task-level restatement over plain tensors (fake-sim semantics, noise passed in)
# --------------------------------------------------------------------------------------

class HumanoidOracle:
    """State + method bodies of ``Humanoid`` (humanoid.py:43-117, 219-296) without isaacgym.

    Fake-sim semantics: ``refresh_*`` are no-ops on tensors that are
    already current; ``set_actor_root_state_tensor_indexed(src, ids)`` is the row copy
    ``root_states[ids] = src[ids]``; ``set_dof_state_tensor_indexed`` is a no-op because
    the writes went through the aliasing ``dof_pos``/``dof_vel`` views already;
    ``set_dof_actuation_force_tensor`` stores the tensor in ``self.forces``.
    """

    def __init__(self, consts: HumanoidConsts, root_states: Tensor, dof_state: Tensor, dof_force: Tensor,
                 sensors: Tensor, dof_limits_lower: Tensor, dof_limits_upper: Tensor, motor_efforts: Tensor,
                 start_rotation: Optional[Tensor] = None, jit: bool = False):
        self.c = consts
        n = root_states.shape[0]
        d = dof_limits_lower.shape[0]
        self.num_envs, self.num_dof = n, d
        self.root_states = root_states
        self.dof_state = dof_state                                  # [N*D, 2]
        self.dof_pos = dof_state.view(n, d, 2)[..., 0]              # humanoid.py:97
        self.dof_vel = dof_state.view(n, d, 2)[..., 1]              # humanoid.py:98
        self.dof_force_tensor = dof_force
        self.vec_sensor_tensor = sensors
        self.dof_limits_lower = dof_limits_lower
        self.dof_limits_upper = dof_limits_upper
        self.motor_efforts = motor_efforts
        self.max_motor_effort = float(motor_efforts.max())          # humanoid.py:170
        # humanoid.py:92-93
        self.initial_root_states = root_states.clone()
        self.initial_root_states[:, 7:13] = 0
        # humanoid.py:99-103
        zero = torch.tensor([0.0])
        self.initial_dof_pos = torch.zeros_like(self.dof_pos, dtype=torch.float)
        self.initial_dof_pos = torch.where(dof_limits_lower > zero, dof_limits_lower,
                                           torch.where(dof_limits_upper < zero, dof_limits_upper,
                                                       self.initial_dof_pos))
        self.initial_dof_vel = torch.zeros_like(self.dof_vel, dtype=torch.float)
        # humanoid.py:106-117
        self.up_vec = ih.to_torch(ih.get_axis_params(1., 2)).repeat((n, 1))
        self.heading_vec = ih.to_torch([1, 0, 0]).repeat((n, 1))
        if start_rotation is None:
            start_rotation = torch.tensor([0.0, 0.0, 0.0, 1.0])    # humanoid.py:180-182
        self.inv_start_rot = ih.quat_conjugate(start_rotation).repeat((n, 1))
        self.basis_vec0 = self.heading_vec.clone()
        self.basis_vec1 = self.up_vec.clone()
        self.targets = ih.to_torch([1000, 0, 0]).repeat((n, 1))
        self.potentials = ih.to_torch([-1000. / consts.dt]).repeat(n)
        self.prev_potentials = self.potentials.clone()
        # vec_task.py:263-277
        self.obs_buf = torch.zeros((n, NUM_OBS), dtype=torch.float)
        self.rew_buf = torch.zeros(n, dtype=torch.float)
        self.reset_buf = torch.ones(n, dtype=torch.long)
        self.timeout_buf = torch.zeros(n, dtype=torch.long)
        self.progress_buf = torch.zeros(n, dtype=torch.long)
        self.randomize_buf = torch.zeros(n, dtype=torch.long)
        self.actions = torch.zeros((n, d), dtype=torch.float)
        self.forces = torch.zeros((n, d), dtype=torch.float)
        self._obs_fn = scripted("humanoid_observations") if jit else humanoid_observations
        self._rew_fn = scripted("humanoid_reward") if jit else humanoid_reward

    # humanoid.py:281-285 (+ the clamp of vec_task.py:327 when ``clip`` is True)
    def pre_physics_step(self, actions: Tensor, clip: bool = False):
        if clip:
            actions = torch.clamp(actions, -self.c.clip_actions, self.c.clip_actions)
        self.actions = actions.clone()
        self.forces = self.actions * self.motor_efforts.unsqueeze(0) * self.c.power_scale

    # humanoid.py:239-251
    def compute_observations(self):
        c = self.c
        (self.obs_buf[:], self.potentials[:], self.prev_potentials[:],
         self.up_vec[:], self.heading_vec[:]) = self._obs_fn(
            self.root_states, self.targets, self.potentials, self.inv_start_rot,
            self.dof_pos, self.dof_vel, self.dof_force_tensor,
            self.dof_limits_lower, self.dof_limits_upper, c.dof_vel_scale,
            self.vec_sensor_tensor, self.actions, c.dt, c.contact_force_scale,
            c.angular_velocity_scale, self.basis_vec0, self.basis_vec1)

    # humanoid.py:219-237
    def compute_reward(self):
        c = self.c
        self.rew_buf[:], self.reset_buf = self._rew_fn(
            self.obs_buf, self.reset_buf, self.progress_buf, self.actions,
            c.up_weight, c.heading_weight, self.potentials, self.prev_potentials,
            c.actions_cost_scale, c.energy_cost_scale, c.joints_at_limit_cost_scale,
            self.max_motor_effort, self.motor_efforts, c.termination_height, c.death_cost,
            float(c.max_episode_length))

    # humanoid.py:253-279; unit_pos/unit_vel are the U(0,1) draws torch_rand_float would make,
    # row k belonging to env_ids[k]
    def reset_idx(self, env_ids: Tensor, unit_pos: Tensor, unit_vel: Tensor):
        positions = ih.rand_float_from_unit(-0.2, 0.2, unit_pos)                     # :258
        velocities = ih.rand_float_from_unit(-0.1, 0.1, unit_vel)                    # :259
        self.dof_pos[env_ids] = ih.tensor_clamp(self.initial_dof_pos[env_ids] + positions,
                                                self.dof_limits_lower, self.dof_limits_upper)   # :261
        self.dof_vel[env_ids] = velocities                                           # :262
        self.root_states[env_ids] = self.initial_root_states[env_ids]                # :265-267 (fake sim)
        to_target = self.targets[env_ids] - self.initial_root_states[env_ids, 0:3]   # :273
        to_target[:, 2] = 0                                                          # :274
        self.prev_potentials[env_ids] = -torch.norm(to_target, p=2, dim=-1) / self.c.dt   # :275
        self.potentials[env_ids] = self.prev_potentials[env_ids].clone()             # :276
        self.progress_buf[env_ids] = 0                                               # :278
        self.reset_buf[env_ids] = 0                                                  # :279

    # humanoid.py:287-296; unit_pos/unit_vel: [K, D] rows in ascending env-id order of the set reset bits
    def post_physics_step(self, unit_pos: Optional[Tensor] = None, unit_vel: Optional[Tensor] = None):
        self.progress_buf += 1
        self.randomize_buf += 1
        env_ids = self.reset_buf.nonzero(as_tuple=False).flatten()
        if len(env_ids) > 0:
            k = len(env_ids)
            self.reset_idx(env_ids, unit_pos[:k], unit_vel[:k])
        self.compute_observations()
        self.compute_reward()
        return env_ids

    # vec_task.py:313-359 (no DR lambdas; physics = identity / caller-provided callback)
    def step(self, actions: Tensor, unit_pos=None, unit_vel=None, physics=None):
        self.pre_physics_step(actions, clip=True)
        if physics is not None:
            physics(self)
        self.post_physics_step(unit_pos, unit_vel)
        self.timeout_buf = (self.progress_buf >= self.c.max_episode_length - 1) & (self.reset_buf != 0)  # :345
        obs = torch.clamp(self.obs_buf, -self.c.clip_obs, self.c.clip_obs)           # :353
        return obs, self.rew_buf, self.reset_buf, {"time_outs": self.timeout_buf}
