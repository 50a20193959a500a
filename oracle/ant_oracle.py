"""CPU restatement of the Ant hot path.  TEST INFRASTRUCTURE.

Follows ``/root/reference/isaacgymenvs/tasks/ant.py`` (file:line cited per function).  Ant is the
Humanoid path with D = 8, 60 observations, 24 sensor floats, no DOF-force block, RAW Euler angles
in the obs row, and -- unlike this fork's Humanoid -- EVERY reward term live.  The JIT-level
functions are pinned bit-for-bit against the reference's own lifted TorchScript functions
(tests/golden/ant_jit.npz).  Helpers from ``oracle.isaac_helpers`` (parity unpinned there).
"""
from dataclasses import dataclass
from typing import Optional, Tuple

import torch
from torch import Tensor

from . import isaac_helpers as ih

NUM_DOF = 8           # ant.py:68
NUM_OBS = 60          # ant.py:67
NUM_SENSOR = 24       # ant.py:82-83 (4 foot sensors x 6)


@dataclass
class AntConsts:
    """ant.py:46-61,115; cfg/task/Ant.yaml:10-29,43."""
    dof_vel_scale: float = 0.2
    contact_force_scale: float = 0.1
    power_scale: float = 1.0
    heading_weight: float = 0.5
    up_weight: float = 0.1
    actions_cost_scale: float = 0.005
    energy_cost_scale: float = 0.05
    joints_at_limit_cost_scale: float = 0.1
    death_cost: float = -2.0
    termination_height: float = 0.31
    max_episode_length: float = 1000
    dt: float = 0.0166
    clip_actions: float = 1.0
    clip_obs: float = float("inf")

    @staticmethod
    def from_cfg(cfg: dict) -> "AntConsts":
        e = cfg["env"]
        return AntConsts(
            dof_vel_scale=e["dofVelocityScale"], contact_force_scale=e["contactForceScale"],
            power_scale=e["powerScale"], heading_weight=e["headingWeight"], up_weight=e["upWeight"],
            actions_cost_scale=e["actionsCost"], energy_cost_scale=e["energyCost"],
            joints_at_limit_cost_scale=e["jointsAtLimitCost"], death_cost=e["deathCost"],
            termination_height=e["terminationHeight"], max_episode_length=e["episodeLength"],
            dt=cfg["sim"]["dt"], clip_actions=e.get("clipActions", float("inf")),
            clip_obs=e.get("clipObservations", float("inf")))


def ant_observations(root_states: Tensor, targets: Tensor, potentials: Tensor, inv_start_rot: Tensor,
                     dof_pos: Tensor, dof_vel: Tensor, dof_limits_lower: Tensor, dof_limits_upper: Tensor,
                     dof_vel_scale: float, sensor_force_torques: Tensor, actions: Tensor, dt: float,
                     contact_force_scale: float, basis_vec0: Tensor, basis_vec1: Tensor, up_axis_idx: int
                     ) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """ant.py:367-408 ``compute_ant_observations`` (obs_buf argument dropped: unused there)."""
    pos = root_states[:, 0:3]
    to_target = targets - pos                                                  # :377
    to_target[:, 2] = 0.0                                                      # :378
    prev_potentials_new = potentials.clone()                                   # :380
    potentials_new = -torch.norm(to_target, p=2, dim=-1) / dt                  # :381
    torso_quat, up_proj, heading_proj, up_vec, heading_vec = ih.compute_heading_and_up(
        root_states[:, 3:7], inv_start_rot, to_target, basis_vec0, basis_vec1, 2)       # :383-384
    vel_loc, angvel_loc, roll, pitch, yaw, angle_to_target = ih.compute_rot(
        torso_quat, root_states[:, 7:10], root_states[:, 10:13], targets, pos)          # :386-387
    dof_pos_scaled = ih.unscale(dof_pos, dof_limits_lower, dof_limits_upper)   # :389
    pieces = [
        pos[:, up_axis_idx].view(-1, 1),                                       # 0
        vel_loc, angvel_loc,                                                   # 1:4, 4:7 (no angular-velocity scale)
        yaw.unsqueeze(-1), roll.unsqueeze(-1), angle_to_target.unsqueeze(-1),  # 7, 8, 9: RAW angles ([0, 2pi], not normalized)
        up_proj.unsqueeze(-1), heading_proj.unsqueeze(-1),                     # 10, 11
        dof_pos_scaled,                                                        # 12:20
        dof_vel * dof_vel_scale,                                               # 20:28
        sensor_force_torques.view(-1, 24) * contact_force_scale,               # 28:52
        actions,                                                               # 52:60
    ]
    return torch.cat(pieces, dim=-1), potentials_new, prev_potentials_new, up_vec, heading_vec


def ant_reward(obs_buf: Tensor, reset_buf: Tensor, progress_buf: Tensor, actions: Tensor, up_weight: float,
               heading_weight: float, potentials: Tensor, prev_potentials: Tensor, actions_cost_scale: float,
               energy_cost_scale: float, joints_at_limit_cost_scale: float, termination_height: float,
               death_cost: float, max_episode_length: float) -> Tuple[Tensor, Tensor]:
    """ant.py:318-364 ``compute_ant_reward``: every term live."""
    hp = obs_buf[:, 11]
    heading_weight_tensor = torch.ones_like(hp) * heading_weight
    heading_reward = torch.where(hp > 0.8, heading_weight_tensor, heading_weight * hp / 0.8)     # :337-338
    up_reward = torch.zeros_like(heading_reward)
    up_reward = torch.where(obs_buf[:, 10] > 0.93, up_reward + up_weight, up_reward)             # :341-342
    actions_cost = torch.sum(actions ** 2, dim=-1)                                               # :345
    electricity_cost = torch.sum(torch.abs(actions * obs_buf[:, 20:28]), dim=-1)                 # :346
    dof_at_limit_cost = torch.sum(obs_buf[:, 12:20] > 0.99, dim=-1)                              # :347 (int64 count)
    alive_reward = torch.ones_like(potentials) * 0.5                                             # :350
    progress_reward = potentials - prev_potentials                                               # :351
    total_reward = progress_reward + alive_reward + up_reward + heading_reward - \
        actions_cost_scale * actions_cost - energy_cost_scale * electricity_cost - \
        dof_at_limit_cost * joints_at_limit_cost_scale                                           # :353-354
    fallen = obs_buf[:, 0] < termination_height
    total_reward = torch.where(fallen, torch.ones_like(total_reward) * death_cost, total_reward)  # :357
    reset = torch.where(fallen, torch.ones_like(reset_buf), reset_buf)                           # :360
    reset = torch.where(progress_buf >= max_episode_length - 1, torch.ones_like(reset_buf), reset)  # :361
    return total_reward, reset


class AntOracle:
    """State + method bodies of ``Ant`` (ant.py:42-117, 213-297) without isaacgym; fake-sim semantics as in
    ``humanoid_oracle.HumanoidOracle``."""

    def __init__(self, consts: AntConsts, root_states: Tensor, dof_state: Tensor, sensors: Tensor,
                 dof_limits_lower: Tensor, dof_limits_upper: Tensor, joint_gears: Tensor):
        self.c = consts
        n, d = root_states.shape[0], dof_limits_lower.shape[0]
        self.num_envs, self.num_dof = n, d
        self.root_states, self.dof_state = root_states, dof_state
        self.dof_pos = dof_state.view(n, d, 2)[..., 0]                         # ant.py:97
        self.dof_vel = dof_state.view(n, d, 2)[..., 1]                         # ant.py:98
        self.vec_sensor_tensor = sensors
        self.dof_limits_lower, self.dof_limits_upper, self.joint_gears = dof_limits_lower, dof_limits_upper, joint_gears
        self.initial_root_states = root_states.clone()                         # ant.py:92-93
        self.initial_root_states[:, 7:13] = 0
        zero = torch.tensor([0.0])
        self.initial_dof_pos = torch.zeros_like(self.dof_pos, dtype=torch.float)
        self.initial_dof_pos = torch.where(dof_limits_lower > zero, dof_limits_lower,
                                           torch.where(dof_limits_upper < zero, dof_limits_upper, self.initial_dof_pos))
        self.up_vec = ih.to_torch(ih.get_axis_params(1., 2)).repeat((n, 1))    # ant.py:106-117
        self.heading_vec = ih.to_torch([1, 0, 0]).repeat((n, 1))
        self.inv_start_rot = ih.quat_conjugate(torch.tensor([0.0, 0.0, 0.0, 1.0])).repeat((n, 1))
        self.basis_vec0, self.basis_vec1 = self.heading_vec.clone(), self.up_vec.clone()
        self.targets = ih.to_torch([1000, 0, 0]).repeat((n, 1))
        self.potentials = ih.to_torch([-1000. / consts.dt]).repeat(n)
        self.prev_potentials = self.potentials.clone()
        self.obs_buf = torch.zeros((n, NUM_OBS), dtype=torch.float)
        self.rew_buf = torch.zeros(n, dtype=torch.float)
        self.reset_buf = torch.ones(n, dtype=torch.long)
        self.progress_buf = torch.zeros(n, dtype=torch.long)
        self.timeout_buf = torch.zeros(n, dtype=torch.long)
        self.randomize_buf = torch.zeros(n, dtype=torch.long)
        self.actions = torch.zeros((n, d), dtype=torch.float)
        self.forces = torch.zeros((n, d), dtype=torch.float)

    def pre_physics_step(self, actions: Tensor, clip: bool = False):           # ant.py:273-277
        if clip:
            actions = torch.clamp(actions, -self.c.clip_actions, self.c.clip_actions)
        self.actions = actions.clone()
        self.forces = self.actions * self.joint_gears * self.c.power_scale

    def compute_observations(self):                                            # ant.py:232-245
        c = self.c
        (self.obs_buf[:], self.potentials[:], self.prev_potentials[:], self.up_vec[:], self.heading_vec[:]) = \
            ant_observations(self.root_states, self.targets, self.potentials, self.inv_start_rot, self.dof_pos,
                             self.dof_vel, self.dof_limits_lower, self.dof_limits_upper, c.dof_vel_scale,
                             self.vec_sensor_tensor, self.actions, c.dt, c.contact_force_scale,
                             self.basis_vec0, self.basis_vec1, 2)

    def compute_reward(self):                                                  # ant.py:213-230
        c = self.c
        self.rew_buf[:], self.reset_buf[:] = ant_reward(
            self.obs_buf, self.reset_buf, self.progress_buf, self.actions, c.up_weight, c.heading_weight,
            self.potentials, self.prev_potentials, c.actions_cost_scale, c.energy_cost_scale,
            c.joints_at_limit_cost_scale, c.termination_height, c.death_cost, float(c.max_episode_length))

    def reset_idx(self, env_ids: Tensor, unit_pos: Tensor, unit_vel: Tensor):  # ant.py:247-271
        positions = ih.rand_float_from_unit(-0.2, 0.2, unit_pos)
        velocities = ih.rand_float_from_unit(-0.1, 0.1, unit_vel)
        self.dof_pos[env_ids] = ih.tensor_clamp(self.initial_dof_pos[env_ids] + positions,
                                                self.dof_limits_lower, self.dof_limits_upper)
        self.dof_vel[env_ids] = velocities
        self.root_states[env_ids] = self.initial_root_states[env_ids]
        to_target = self.targets[env_ids] - self.initial_root_states[env_ids, 0:3]
        to_target[:, 2] = 0.0
        self.prev_potentials[env_ids] = -torch.norm(to_target, p=2, dim=-1) / self.c.dt
        self.potentials[env_ids] = self.prev_potentials[env_ids].clone()
        self.progress_buf[env_ids] = 0
        self.reset_buf[env_ids] = 0

    def post_physics_step(self, unit_pos: Optional[Tensor] = None, unit_vel: Optional[Tensor] = None):   # ant.py:279-287
        self.progress_buf += 1
        self.randomize_buf += 1                                                # ant.py:281
        env_ids = self.reset_buf.nonzero(as_tuple=False).flatten()
        if len(env_ids) > 0:
            self.reset_idx(env_ids, unit_pos[:len(env_ids)], unit_vel[:len(env_ids)])
        self.compute_observations()
        self.compute_reward()
        return env_ids

    def step(self, actions: Tensor, unit_pos=None, unit_vel=None, physics=None):   # vec_task.py:313-359
        self.pre_physics_step(actions, clip=True)
        if physics is not None:
            physics(self)
        self.post_physics_step(unit_pos, unit_vel)
        self.timeout_buf = (self.progress_buf >= self.c.max_episode_length - 1) & (self.reset_buf != 0)
        return self.obs_buf, self.rew_buf, self.reset_buf, {"time_outs": self.timeout_buf}
