"""CPU restatement of the MA_OP3 hot path.  TEST INFRASTRUCTURE.

Follows ``/root/reference/isaacgymenvs/tasks/MA_OP3.py`` (two ROBOTIS OP3 humanoids carrying a table; file:line cited
per statement; constants from ``cfg/task/MA_OP3.yaml``).  The three JIT functions ``compute_op3_observations``
(:662-708), ``compute_op3_reward`` (:550-658) and ``compute_objective_reward`` (:496-547) are pinned bit-for-bit
against the reference's own lifted TorchScript functions (tests/test_oracle_vs_reference.py,
tests/golden/ma_op3_jit.npz).  The method bodies ``compute_observations`` (:411-438), ``compute_reward`` (:360-409),
the five contact-force terms (:319-358) and ``reset_idx`` (:440-489) are restated line by line.

What the reference leaves undefined, and what is assumed here (and in the CUDA path) -- each marked ASSUMPTION below:

* ``pre_physics_step`` is ``pass`` and ``post_physics_step`` zeroes the buffers (:309-317).  The un-stubbed step used
  here is the ordering of the task this file derives from (anymal.py:229-242): ``actions`` stored and
  ``targets = action_scale * actions + default_dof_pos`` before physics; ``progress_buf += 1``, reset of the envs whose
  ``reset_buf`` is set, ``compute_observations``, ``compute_reward`` after it.
* ``self.prev_torques`` is read at :373 and assigned nowhere; ``self.last_actions`` is only ever zeroed (:486).  Both
  are task state here, refreshed at the END of the step (``prev_torques = torques``, ``last_actions = actions``).
* ``numObservations`` / ``numActions`` are forced to 1 (:62-63), which makes ``self.actions[:, i]`` one-dimensional and
  the ``torch.cat`` of :697 fail.  Here actions are ``[N, agents, 22]`` (what ``default_dof_pos[:, :, i]`` :126-129 and
  ``MA_VecTask.zero_actions`` multi_vec_task.py:215 imply) and an agent's observation row has 88 entries.

``_reward_gripper_contact_hold`` (:346-358) reduces over ALL envs (``left_contacts.sum() == 1``): the term is a batch
-global 0/1 added to every env's reward.  It is restated as such; ``grip_counts`` exposes the two counts so that a
sharded run can sum them over ranks first.

Random draws are inputs (U(0,1) "unit" tensors); the affine map is upstream ``torch_rand_float``.
"""
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import torch
from torch import Tensor

from . import isaac_helpers as ih

NUM_AGENTS = 2
NUM_DOF = 22                      # cfg/task/MA_OP3.yaml defaultJointAngles: 22 named joints
NUM_OBS = 88                      # 3 + 3 + 3 + 22 + 22 + 22 + 7 + 3 + 3 (:697-706)
ACTORS_PER_ENV = 3                # two OP3 + the table (:78-79)
LEG_IDS = [2, 3, 4, 5, 6, 7, 11, 12, 13, 14, 15, 16, 17, 21]     # :622-623
OP3_DOF_NAMES = ["head_pan", "head_tilt", "l_hip_yaw", "l_hip_roll", "l_hip_pitch", "l_knee", "l_ank_pitch", "l_ank_roll",
                 "l_sho_pitch", "l_sho_roll", "l_el", "l_gr", "r_hip_yaw", "r_hip_roll", "r_hip_pitch", "r_knee",
                 "r_ank_pitch", "r_ank_roll", "r_sho_pitch", "r_sho_roll", "r_el", "r_gr"]


@dataclass
class OP3Consts:
    """MA_OP3.py:16-73, cfg/task/MA_OP3.yaml."""
    lin_vel_scale: float = 3.0
    ang_vel_scale: float = 0.25
    dof_pos_scale: float = 1.0
    dof_vel_scale: float = 0.01
    action_scale: float = 1.0
    scales: Dict[str, float] = field(default_factory=lambda: {
        "lin_vel_xy": 1.0, "ang_vel_z": 0.5, "torque": -0.000025, "heading_scale": 1.0, "up_scale": 0.1, "air_time": 0.5,
        "syns_hip": -0.00025, "no_fly": 0.5, "stand_scale": -0.0, "action_rate": -0.01})
    command_x_range: List[float] = field(default_factory=lambda: [0.0, 0.0])
    command_y_range: List[float] = field(default_factory=lambda: [0.0, 10.0])
    command_yaw_range: List[float] = field(default_factory=lambda: [0.0, 0.0])
    dt: float = 0.0166
    episode_length_s: float = 50.0
    clip_actions: float = 1.0
    clip_obs: float = 5.0

    @property
    def max_episode_length(self) -> int:
        return int(self.episode_length_s / self.dt + 0.5)                       # :69

    def rew_scales(self) -> Dict[str, float]:
        """:25-35 scaled by dt at :80-81 (python doubles)."""
        return {k: v * self.dt for k, v in self.scales.items()}

    @staticmethod
    def from_cfg(cfg: dict) -> "OP3Consts":
        e, l, c = cfg["env"], cfg["env"]["learn"], cfg["env"]["control"]
        r = e["randomCommandVelocityRanges"]
        scales = {"lin_vel_xy": l["linearVelocityXYRewardScale"], "ang_vel_z": l["angularVelocityZRewardScale"],
                  "torque": l["torqueRewardScale"], "heading_scale": l["headingScale"], "up_scale": l["upScale"],
                  "air_time": l["feetAirTimeRewardScale"], "syns_hip": l["syncronizeHipRewardScale"],
                  "no_fly": l["noflyRewardScale"], "stand_scale": l["standRewardScale"],
                  "action_rate": l["actionRateRewardScale"]}
        return OP3Consts(
            lin_vel_scale=l["linearVelocityScale"], ang_vel_scale=l["angularVelocityScale"],
            dof_pos_scale=l["dofPositionScale"], dof_vel_scale=l["dofVelocityScale"], action_scale=c["actionScale"],
            scales=scales, command_x_range=list(r["linear_x"]), command_y_range=list(r["linear_y"]),
            command_yaw_range=list(r["yaw"]), dt=cfg["sim"]["dt"], episode_length_s=l["episodeLength_s"],
            clip_actions=e.get("clipActions", float("inf")), clip_obs=e.get("clipObservations", float("inf")))


# --------------------------------------------------------------------------------------
# JIT-level functions
# --------------------------------------------------------------------------------------

def op3_observations(root_states: Tensor, target: Tensor, dof_pos: Tensor, default_dof_pos: Tensor, dof_vel: Tensor,
                     gravity_vec: Tensor, actions: Tensor, object_state: Tensor, lin_vel_scale: float, ang_vel_scale: float,
                     dof_pos_scale: float, dof_vel_scale: float, dt: float, potentials: Tensor, goal_pos: Tensor
                     ) -> Tuple[Tensor, Tensor, Tensor]:
    """MA_OP3.py:662-708 ``compute_op3_observations`` -> (obs [N,88], potentials, prev_potentials)."""
    torso_position = root_states[:, 0:3]
    to_target = goal_pos - torso_position                                                       # :680
    to_target[:, 2] = 0.0                                                                       # :682
    object_pose = object_state[:, 0:7]
    object_linvel = object_state[:, 7:10]
    prev_potentials_new = potentials.clone()                                                    # :688
    potentials = -torch.norm(to_target, p=2, dim=-1) / dt                                       # :689
    base_quat = root_states[:, 3:7]
    base_lin_vel = ih.quat_rotate_inverse(base_quat, root_states[:, 7:10]) * lin_vel_scale      # :692
    base_ang_vel = ih.quat_rotate_inverse(base_quat, root_states[:, 10:13]) * ang_vel_scale     # :693
    projected_gravity = ih.quat_rotate(base_quat, gravity_vec)                                  # :694
    dof_pos_scaled = (dof_pos - default_dof_pos) * dof_pos_scale                                # :695
    obs = torch.cat((base_lin_vel, base_ang_vel, projected_gravity, dof_pos_scaled, dof_vel * dof_vel_scale, actions,
                     object_pose, object_linvel, target), dim=-1)                               # :697-706
    return obs, potentials, prev_potentials_new


def op3_reward(root_states: Tensor, commands: Tensor, prev_torques: Tensor, torques: Tensor, episode_lengths: Tensor,
               rew_scales: Dict[str, float], max_episode_length: int, targets: Tensor, vec0: Tensor, vec1: Tensor,
               inv_start_rot: Tensor, reset_buf: Tensor, actions: Tensor, last_actions: Tensor, dof_pos: Tensor,
               default_dof_pos: Tensor, potentials: Tensor, prev_potentials: Tensor, rew_airtime: Tensor, rew_step: Tensor,
               rew_no_fly: Tensor, rew_gripper: Tensor, rew_gripper_hold: Tensor, gravity_vec: Tensor, root_table: Tensor
               ) -> Tuple[Tensor, Tensor]:
    """MA_OP3.py:550-658 ``compute_op3_reward``.  ``reset_buf`` is modified IN PLACE at :653 like the reference's."""
    quat = root_states[:, 3:7]
    ang_vel = ih.quat_rotate_inverse(quat, root_states[:, 10:13])                               # :579
    position = root_states[:, 0:3]
    rotation = root_states[:, 3:7]
    num_envs = position.shape[0]
    torso_quat = ih.quat_mul(rotation, inv_start_rot)                                           # :593
    up_vec = ih.get_basis_vector(torso_quat, vec1).view(num_envs, 3)                            # :601 (heading terms :595-600 are dead)
    up_proj = up_vec[:, 2]
    rew_up = torch.zeros_like(up_proj)
    rew_up = torch.where(up_proj > 0.95, rew_up + rew_scales['up_scale'], rew_up)               # :605-606
    alive_reward = torch.ones_like(potentials) * 2                                              # :611
    progress_reward = (potentials - prev_potentials) * 5                                        # :612
    projected_gravity = ih.quat_rotate_inverse(quat, gravity_vec)                               # :616
    rew_orient = torch.sum(torch.square(projected_gravity[:, :2]), dim=1) * -0                  # :616
    rew_torque = torch.sum(torch.abs(prev_torques - torques), dim=1) * rew_scales["torque"]     # :619
    leg_position_pinalty = torch.sum(torch.abs(dof_pos[:, LEG_IDS] - default_dof_pos[:, LEG_IDS]), dim=1)   # :622-623
    rew_syns = leg_position_pinalty * rew_scales["syns_hip"]                                    # :624
    rew_action_rate = torch.sum(torch.square(last_actions - actions), dim=1) * rew_scales["action_rate"]   # :627
    rew_stand_still = torch.sum(torch.abs(dof_pos - default_dof_pos), dim=1) * \
        (torch.norm(commands[:, :2], dim=1) < 0.1) * rew_scales["stand_scale"]                  # :630
    ang_vel_error = torch.square(targets[:, 2] - ang_vel[:, 2])                                 # :633
    rew_ang_vel_z = torch.exp(-ang_vel_error / 0.1) * 0.1                                       # :633
    max_distance = 0.32
    scale_reward = 0.5
    agent_distance_to_table = torch.norm(position - root_table[:, :3], p=2, dim=-1)             # :645
    rew_distobject = torch.exp(-agent_distance_to_table / max_distance) * scale_reward          # :645
    total_reward = progress_reward + alive_reward + rew_torque + rew_up + rew_airtime + rew_ang_vel_z + rew_step + \
        rew_no_fly + rew_stand_still + rew_action_rate + rew_syns + rew_orient + rew_distobject + rew_gripper_hold   # :648-649
    total_reward = torch.clip(total_reward, 0., None)                                           # :650
    reset_buf[episode_lengths >= max_episode_length] = 1                                        # :653 (in place)
    reset_buf = torch.where(up_proj < 0.90, torch.ones_like(reset_buf), reset_buf)              # :656
    reset_buf = torch.where(agent_distance_to_table > 0.40, torch.ones_like(reset_buf), reset_buf)   # :656
    reset_buf[episode_lengths == 0] = 0                                                         # :656
    return total_reward.detach(), reset_buf


def objective_reward(reward_agent: Tensor, reset_buf: Tensor, target_pos: Tensor, object_pos: Tensor, table_root: Tensor,
                     initial_pos: Tensor, potentials: Tensor, prev_potentials: Tensor, vec0: Tensor, vec1: Tensor,
                     num_envs: int) -> Tuple[Tensor, Tensor]:
    """MA_OP3.py:496-547 ``compute_objective_reward`` (goal distances :510-511 and the heading reward :517-526 are dead)."""
    table_position = table_root[:, 0:3]
    table_quat = table_root[:, 3:7]
    table_height = table_position[:, 2]
    up_vec = ih.get_basis_vector(table_quat, vec1).view(num_envs, 3)                            # :528
    up_proj = up_vec[:, 2]
    rew_up = torch.zeros_like(up_proj)
    rew_up = torch.where(up_proj > 0.98, rew_up + 0.1, rew_up)                                  # :532-533
    rew_up = torch.cat((rew_up.unsqueeze(1), rew_up.unsqueeze(1)), dim=1)
    rew_base_height = torch.square(table_position[:, 2] - 0.29) * -0.001                        # :536
    rew_base_height = torch.cat((rew_base_height.unsqueeze(1), rew_base_height.unsqueeze(1)), dim=1)
    dist_rew = (potentials - prev_potentials) * 5                                               # :540
    dist_rew = torch.cat((dist_rew.unsqueeze(1), dist_rew.unsqueeze(1)), dim=1)
    rew_buf = reward_agent + dist_rew + rew_up + rew_base_height                                # :542
    reset_buf = torch.where(up_proj < 0.90, torch.ones_like(reset_buf), reset_buf)              # :544
    reset_buf = torch.where(table_height < 0.25, torch.ones_like(reset_buf), reset_buf)         # :545
    return rew_buf, reset_buf


def grip_counts(contact_forces: Tensor, gripper_indices: Tensor) -> Tensor:
    """int64 [agents, 2]: how many envs have an x contact force above 0.1 on the left / right gripper (:351-352)."""
    a = gripper_indices.shape[0]
    out = torch.zeros((a, 2), dtype=torch.long)
    for i in range(a):
        out[i, 0] = (contact_forces[:, gripper_indices[i, 0], 0] > 0.1).sum()
        out[i, 1] = (contact_forces[:, gripper_indices[i, 1], 0] > 0.1).sum()
    return out


def _rand_float(lo: float, hi: float, unit: Tensor) -> Tensor:
    return (hi - lo) * unit + lo                                                                # torch_rand_float


class MAOP3Oracle:
    """State + method bodies of ``MA_OP3`` (MA_OP3.py:14-148, 309-489) without isaacgym, fake-sim semantics."""

    def __init__(self, consts: OP3Consts, root_states: Tensor, dof_state: Tensor, torques: Tensor, contact_forces: Tensor,
                 default_dof_pos_row: Tensor, feet_indices: Tensor, gripper_indices: Tensor,
                 initial_root_states: Optional[Tensor] = None, start_rotation: Optional[Tensor] = None):
        self.c = consts
        a, d = NUM_AGENTS, NUM_DOF
        n = root_states.numel() // (ACTORS_PER_ENV * 13)
        self.num_envs, self.num_agents, self.num_dof, self.dt = n, a, d, consts.dt
        self.root_states = root_states.view(n, ACTORS_PER_ENV, 13)                              # :100
        if initial_root_states is None:
            initial_root_states = self.root_states.clone()                                      # :99
            initial_root_states[:, :, 7:13] = 0                                                 # :100
        self.initial_root_states = initial_root_states
        self.goal_pos = torch.tensor([0.0, -10.0, 0.0]).repeat((n, 1))                          # :100
        self.table_state = self.root_states[:, 2]                                               # :102
        self.table_positions = self.table_state[:, 0:3]                                         # :103
        self.table_potentials = torch.tensor([-1000. / self.dt], dtype=torch.float).repeat(n)   # :106
        self.prev_table_potentials = self.table_potentials.clone()                              # :107
        self.dof_state = dof_state
        self.dof_pos = dof_state.view(n, a, d, 2)[..., 0]                                       # :110
        self.dof_vel = dof_state.view(n, a, d, 2)[..., 1]                                       # :111
        self.default_dof_pos = default_dof_pos_row.view(1, 1, d).repeat(n, a, 1)                # :112,126-129
        self.contact_forces = contact_forces.view(n, -1, 3)                                     # :114
        self.torques = torques.view(n, a, d)                                                    # :115
        self.prev_torques = torch.zeros(n, a, d)                                                # ASSUMPTION (see header)
        self.commands = torch.zeros(n, 3)                                                       # :117
        self.gravity_vec = torch.tensor([0.0, 0.0, -1.0]).repeat((n, a, 1))                     # :132
        self.actions = torch.zeros(n, a, d)                                                     # :133 with [N, agents, 22]
        self.feet_indices, self.gripper_indices = feet_indices, gripper_indices                 # :238-246,294-306
        f = feet_indices.shape[1]
        self.feet_air_time = torch.zeros(n, a, f)                                               # :134
        self.last_contacts = torch.zeros(n, a, f, dtype=torch.bool)                             # :135
        self.last_actions = self.actions.clone()                                                # :137
        self.up_vec = torch.tensor([0.0, 0.0, 1.0]).repeat((n, a + 1, 1))                       # :139
        self.heading_vec = torch.tensor([1.0, 0.0, 0.0]).repeat((n, a + 1, 1))                  # :140
        self.basis_vec0 = self.heading_vec.clone()
        self.basis_vec1 = self.up_vec.clone()
        if start_rotation is None:
            start_rotation = torch.tensor([0.0, 0.0, 0.0, 1.0])                                 # :228-231, euler (0,0,0)
        self.inv_start_rot = ih.quat_conjugate(start_rotation).repeat((n, a, 1))                # :143
        self.targets = torch.tensor([0.0, 10.0, 0.0]).repeat((n, 1))                            # :144
        self.potentials = torch.tensor([-1000. / self.dt], dtype=torch.float).repeat(n, a)      # :145
        self.prev_potentials = self.potentials.clone()                                          # :146
        self.max_episode_length = consts.max_episode_length
        self.rew_scales = consts.rew_scales()
        self.obs_buf = torch.zeros(n, a, NUM_OBS)                                               # multi_vec_task.py:111-112
        self.rew_buf = torch.zeros(n, a)                                                        # multi_vec_task.py:115-116
        self.reset_buf = torch.ones(n, dtype=torch.long)
        self.progress_buf = torch.zeros(n, dtype=torch.long)
        self.timeout_buf = torch.zeros(n, dtype=torch.long)
        self.pd_targets = torch.zeros(n, a, d)
        self.global_grip_counts: Optional[Tensor] = None      # set by a sharded run: counts summed over all shards

    # ASSUMPTION (see header): anymal.py:229-232 (+ clamp vec_task.py:327)
    def pre_physics_step(self, actions: Tensor, clip: bool = False):
        if clip:
            actions = torch.clamp(actions, -self.c.clip_actions, self.c.clip_actions)
        self.actions = actions.clone().view(self.num_envs, self.num_agents, self.num_dof)
        self.pd_targets = self.c.action_scale * self.actions + self.default_dof_pos

    # :319-329
    def _reward_air_time(self, agent_idx: int) -> Tensor:
        feet_agent = self.feet_indices[agent_idx]
        contact = self.contact_forces[:, feet_agent, 2] > 1.1
        contact_filt = torch.logical_or(contact, self.last_contacts[:, agent_idx])
        self.last_contacts[:, agent_idx] = contact
        first_contact = (self.feet_air_time[:, agent_idx] > 0.) * contact_filt
        self.feet_air_time[:, agent_idx] += self.dt
        rew_airTime = torch.sum((self.feet_air_time[:, agent_idx] - 0.5) * first_contact, dim=1)
        rew_airTime *= (torch.norm(self.commands[:, :2], dim=1) > 0.1)
        self.feet_air_time[:, agent_idx] *= ~contact_filt
        return rew_airTime

    # :331-334
    def _reward_feet_contact_forces(self, agent_idx: int) -> Tensor:
        feet_agent = self.feet_indices[agent_idx]
        return torch.sum((torch.norm(self.contact_forces[:, feet_agent, :], dim=-1) - 450).clip(min=0.), dim=1)

    # :336-340
    def _reward_no_fly(self, agent_idx: int) -> Tensor:
        feet_agent = self.feet_indices[agent_idx]
        contacts = self.contact_forces[:, feet_agent, 2] > 0.1
        single_contact = torch.sum(1. * contacts, dim=1) == 1
        return 1. * single_contact

    # :342-344
    def _reward_gripper_contact_forces(self, agent_idx: int) -> Tensor:
        gripper_agent = self.gripper_indices[agent_idx]
        return torch.sum((torch.norm(self.contact_forces[:, gripper_agent, :], dim=-1) - 400).clip(min=0.), dim=1)

    # :346-358 -- a reduction over ALL envs; returns a 0-dim bool tensor like the reference
    def _reward_gripper_contact_hold(self, agent_idx: int) -> Tensor:
        if self.global_grip_counts is not None:
            cnt = self.global_grip_counts[agent_idx]
        else:
            cnt = grip_counts(self.contact_forces, self.gripper_indices)[agent_idx]
        successful_left_grip = (cnt[0] == 1)
        successful_right_grip = (cnt[1] == 1)
        return successful_left_grip and successful_right_grip

    # :360-409
    def compute_reward(self, actions=None):
        for i in range(self.num_agents):
            reward_airtime = self._reward_air_time(i) * self.rew_scales["air_time"]
            reward_step = self._reward_feet_contact_forces(i)
            reward_no_fly = self._reward_no_fly(i) * self.rew_scales["no_fly"]
            reward_gripper = self._reward_gripper_contact_forces(i) * 0.5
            reward_gripper_hold = self._reward_gripper_contact_hold(i)
            self.rew_buf[:, i], self.reset_buf[:] = op3_reward(
                self.root_states[:, i], self.commands, self.prev_torques[:, i], self.torques[:, i], self.progress_buf,
                self.rew_scales, self.max_episode_length, self.targets, self.heading_vec[:, i], self.up_vec[:, i],
                self.inv_start_rot[:, i], self.reset_buf, self.actions[:, i], self.last_actions[:, i], self.dof_pos[:, i],
                self.default_dof_pos[:, i], self.potentials[:, i], self.prev_potentials[:, i], reward_airtime, reward_step,
                reward_no_fly, reward_gripper, reward_gripper_hold, self.gravity_vec[:, i], self.table_state)
        self.rew_buf[:], self.reset_buf[:] = objective_reward(
            self.rew_buf[:], self.reset_buf[:], self.targets, self.table_positions, self.root_states[:, self.num_agents],
            self.initial_root_states[:, self.num_agents, 0:3], self.table_potentials, self.prev_table_potentials,
            self.basis_vec0[:, 2], self.basis_vec1[:, 2], self.num_envs)

    # :411-438
    def compute_observations(self):
        to_target_goal = self.targets - self.table_positions                                    # :418
        self.prev_table_potentials = -torch.norm(to_target_goal, p=2, dim=-1) / self.dt         # :419
        for i in range(self.num_agents):
            self.obs_buf[:, i], self.potentials[:, i], self.prev_potentials[:, i] = op3_observations(
                self.root_states[:, i], self.targets, self.dof_pos[:, i], self.default_dof_pos[:, i], self.dof_vel[:, i],
                self.gravity_vec[:, i], self.actions[:, i], self.table_state, self.c.lin_vel_scale, self.c.ang_vel_scale,
                self.c.dof_pos_scale, self.c.dof_vel_scale, self.dt, self.potentials[:, i], self.goal_pos)

    # :440-489; unit_cmd [K,3] = the draws of :472-474 in order (x, y, yaw)
    def reset_idx(self, env_ids: Tensor, unit_cmd: Tensor):
        c = self.c
        self.dof_pos[env_ids] = self.default_dof_pos[env_ids] + 0.0                             # :445-456 (offsets are zeros)
        self.dof_vel[env_ids] = 0.0                                                             # :457
        self.root_states[env_ids] = self.initial_root_states[env_ids]                           # :460-470 (fake sim row copy)
        self.commands[env_ids, 0] = _rand_float(c.command_x_range[0], c.command_x_range[1], unit_cmd[:, 0])       # :472
        self.commands[env_ids, 1] = _rand_float(c.command_y_range[0], c.command_y_range[1], unit_cmd[:, 1])       # :473
        self.commands[env_ids, 2] = _rand_float(c.command_yaw_range[0], c.command_yaw_range[1], unit_cmd[:, 2])   # :474
        to_target = self.goal_pos[env_ids].unsqueeze(1) - self.initial_root_states[env_ids, 0:self.num_agents, 0:3]   # :476
        to_target[:, :, 2] = 0.0
        self.prev_potentials[env_ids] = -torch.norm(to_target, p=2, dim=-1) / self.dt           # :478
        self.potentials[env_ids] = self.prev_potentials[env_ids].clone()                        # :479
        to_target_goal = self.targets[env_ids] - self.initial_root_states[env_ids, 2, 0:3]      # :481
        to_target_goal[:, 2] = 0.0
        self.prev_table_potentials[env_ids] = -torch.norm(to_target_goal, p=2, dim=1) / self.dt  # :484
        self.table_potentials[env_ids] = self.prev_table_potentials[env_ids].clone()            # :484
        self.last_actions[env_ids] = 0.                                                         # :486
        self.feet_air_time[env_ids] = 0                                                         # :489
        self.progress_buf[env_ids] = 0                                                          # :489
        self.reset_buf[env_ids] = 0                                                             # :489

    # ASSUMPTION (see header): anymal.py:234-242 ordering + end-of-step refresh of prev_torques / last_actions
    def post_physics_step(self, unit_cmd: Optional[Tensor] = None):
        self.progress_buf += 1
        env_ids = self.reset_buf.nonzero(as_tuple=False).squeeze(-1)
        if len(env_ids) > 0:
            self.reset_idx(env_ids, unit_cmd[:len(env_ids)])
        self.compute_observations()
        self.compute_reward()
        self.prev_torques = self.torques.clone()
        self.last_actions = self.actions.clone()
        self.timeout_buf = (self.progress_buf >= self.max_episode_length - 1) & (self.reset_buf != 0)   # multi_vec_task.py:193
        return env_ids


def oracle_from_state(st: Dict[str, Tensor], consts: Optional[OP3Consts] = None, default_row: Optional[Tensor] = None,
                      feet: Optional[Tensor] = None, grip: Optional[Tensor] = None) -> MAOP3Oracle:
    """Build the oracle from a ``synthetic.ma_op3_state`` snapshot (all task state taken from it)."""
    consts = consts or OP3Consts()
    if default_row is None:
        default_row = torch.zeros(NUM_DOF)
        default_row[4], default_row[5], default_row[14], default_row[15] = -0.95, 1.49, 0.95, -1.49   # MA_OP3.yaml
    n = st["root_states"].numel() // (ACTORS_PER_ENV * 13)
    if feet is None or grip is None:
        nb = st["contact_forces"].numel() // (n * 3)
        per = (nb - 1) // NUM_AGENTS
        feet = torch.tensor([[a * per + b for b in (10, 11, 16, 17)] for a in range(NUM_AGENTS)], dtype=torch.long)
        grip = torch.tensor([[a * per + b for b in (20, 24)] for a in range(NUM_AGENTS)], dtype=torch.long)
    o = MAOP3Oracle(consts, st["root_states"].clone(), st["dof_state"].clone(), st["torques"].clone(),
                    st["contact_forces"].clone(), default_row, feet, grip,
                    initial_root_states=st.get("initial_root_states"))
    for name in ("prev_torques", "commands", "actions", "last_actions", "feet_air_time", "last_contacts", "potentials",
                 "prev_potentials", "table_potentials", "progress_buf", "reset_buf"):
        if name in st:
            setattr(o, name, st[name].clone())
    return o


def run_jit_level(o: MAOP3Oracle, fn_obs=op3_observations, fn_rew=op3_reward, fn_obj=objective_reward) -> Dict[str, Tensor]:
    """compute_observations + compute_reward of ``o`` with the three JIT-level functions swapped for ``fn_*`` (the
    reference's lifted TorchScript functions when pinning).  Mutates ``o`` like the methods do."""
    saved = globals()["op3_observations"], globals()["op3_reward"], globals()["objective_reward"]
    globals()["op3_observations"], globals()["op3_reward"], globals()["objective_reward"] = fn_obs, fn_rew, fn_obj
    try:
        o.compute_observations()
        o.compute_reward()
    finally:
        globals()["op3_observations"], globals()["op3_reward"], globals()["objective_reward"] = saved
    return dict(obs=o.obs_buf.clone(), potentials=o.potentials.clone(), prev_potentials=o.prev_potentials.clone(),
                prev_table_potentials=o.prev_table_potentials.clone(), rew=o.rew_buf.clone(), reset=o.reset_buf.clone(),
                feet_air_time=o.feet_air_time.clone(), last_contacts=o.last_contacts.clone())
