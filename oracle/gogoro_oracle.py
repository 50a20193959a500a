"""CPU restatement of the registered Gogoro scooter hot path.  TEST INFRASTRUCTURE.

Follows ``/root/reference/isaacgymenvs/tasks/gogoro_new.py`` (file:line cited per
function; constants from ``isaacgymenvs/cfg/task/Gogoro.yaml``).  The JIT-level
functions are pinned bit-for-bit against the reference's own lifted TorchScript
functions (tests/golden/gogoro_*.npz).  Helpers from ``oracle.isaac_helpers`` are a
restated un-vendored module (parity unpinned there).

All random draws are inputs ("unit" tensors: U(0,1) or N(0,1) samples); the
affine maps applied to them are the reference's (``get_randoms`` :466-467,
``get_randoms_norm`` :469-470).
"""
from dataclasses import dataclass, field
from typing import List, Optional, Tuple

import numpy as np
import torch
from torch import Tensor

from . import isaac_helpers as ih

NUM_DOF = 39          # assets/urdf/gogoro/urdf/scooter_V13.urdf: 34 revolute + 2 continuous + 3 prismatic
NUM_OBS = 6           # gogoro_new.py:101
NUM_ACT = 1           # gogoro_new.py:102
HISTORY = 5           # gogoro_new.py:92
RESET_NOISE_COLS = 10  # see GogoroOracle.reset_idx


@dataclass
class GogoroConsts:
    """cfg["env"], cfg["noises"], cfg["sim"] values read at gogoro_new.py:38-68,84-87,122."""
    max_steps: int = 1000
    dt: float = 0.03
    imu_filter_noise: List[float] = field(default_factory=lambda: [0, 0.001])
    imu_noise: List[float] = field(default_factory=lambda: [0, 0.001])
    speed_sensor_noise: List[float] = field(default_factory=lambda: [0, 0.3])
    steering_action_noise: List[float] = field(default_factory=lambda: [0, 0.03])
    speed_range: List[float] = field(default_factory=lambda: [4.0, 13.0])
    speed_freq_update: int = 300
    yaw_freq_update: int = 300
    seat_offset_x_range: List[float] = field(default_factory=lambda: [0, 0.02])
    seat_offset_y_range: List[float] = field(default_factory=lambda: [0, 0.02])
    seat_offset_z_range: List[float] = field(default_factory=lambda: [0, 0.02])
    seat_offset_xr_range: List[float] = field(default_factory=lambda: [0, 0.05])
    steering_offset: List[float] = field(default_factory=lambda: [0, 0.01])
    speed_sensor_offset: List[float] = field(default_factory=lambda: [-0.5, 0.5])
    max_steering: float = 0.5             # gogoro_new.py:86
    max_steering_change: float = 0.2      # gogoro_new.py:87
    steering_joint: int = 0               # dof_name_to_id['steering_joint'] (asset dependent)
    rear_wheel_joint: int = 1             # dof_name_to_id['rear_wheel_joint']

    @staticmethod
    def from_cfg(cfg: dict, steering_joint: int = 0, rear_wheel_joint: int = 1) -> "GogoroConsts":
        nz = cfg["noises"]
        return GogoroConsts(
            max_steps=cfg["env"]["max_steps"], dt=cfg["sim"]["dt"],
            imu_filter_noise=list(nz["imu_filter_noise"]), imu_noise=list(nz["imu_noise"]),
            speed_sensor_noise=list(nz["speed_sensor_noise"]),
            steering_action_noise=list(nz["steering_action_noise"]),
            speed_range=list(nz["speed_range"]), speed_freq_update=nz["speed_freq_update"],
            yaw_freq_update=nz["yaw_freq_update"],
            seat_offset_x_range=list(nz["seat_offset_x_range"]),
            seat_offset_y_range=list(nz["seat_offset_y_range"]),
            seat_offset_z_range=list(nz["seat_offset_z_range"]),
            seat_offset_xr_range=list(nz["seat_offset_xr_range"]),
            steering_offset=list(nz["steering_offset"]),
            speed_sensor_offset=list(nz["speed_sensor_offset"]),
            steering_joint=steering_joint, rear_wheel_joint=rear_wheel_joint)


# --------------------------------------------------------------------------------------
# JIT-level functions
# --------------------------------------------------------------------------------------

def shortest_angle_distance(yaw1: Tensor, yaw2: Tensor) -> Tensor:
    """gogoro_new.py:687-689 (torch ``%`` = floor-mod)."""
    return (yaw2 - yaw1 + torch.pi) % (2 * torch.pi) - torch.pi


def gogoro_observations(root_state: Tensor, desired_yaw: Tensor, last_command: Tensor) -> Tensor:
    """gogoro_new.py:692-723 ``compute_gogoro_observations`` -> [N,6] = roll, d_roll, d_yaw, speed, d_yaw_cmd, cmd."""
    quat = root_state[:, 3:7]
    roll, _, yaw = ih.get_euler_xyz(quat)                                   # :696
    lin_b = ih.quat_rotate_inverse(quat, root_state[:, 7:10])               # :698
    ang_b = ih.quat_rotate_inverse(quat, root_state[:, 10:13])              # :699

    roll = torch.where(roll > np.pi, roll - np.pi * 2, roll)                # :704
    roll = torch.where(roll < -np.pi, roll + np.pi * 2, roll)               # :705
    yaw = torch.where(yaw > np.pi, yaw - np.pi * 2, yaw)                    # :707
    yaw = torch.where(yaw < -np.pi, yaw + np.pi * 2, yaw)                   # :708
    delta_yaw = shortest_angle_distance(yaw, desired_yaw)                   # :710

    cols = [roll, ang_b[:, 0], ang_b[:, 2], lin_b[:, 0], delta_yaw, last_command]   # :714-721
    return torch.cat([c.unsqueeze(-1) for c in cols], dim=1)


def gogoro_reward(curent_obs: Tensor, progress_buf: Tensor, action_history: Tensor,
                  max_episode_length: int) -> Tuple[Tensor, Tensor]:
    """gogoro_new.py:645-684 ``compute_gogoro_reward`` (last frame of curent_obs[N,B,6])."""
    tilt = curent_obs[:, -1, 0]
    dtilt = curent_obs[:, -1, 1]
    yaw_err = curent_obs[:, -1, 4]

    max_tilt = 0.30
    tilt_err = torch.clamp(tilt / max_tilt, -1.0, 1.0)                      # :655-656
    yaw_err = torch.clamp(yaw_err / np.pi, -1.0, 1.0)                       # :658-659
    dtilt_err = torch.clamp(dtilt / 0.3, -1.0, 1.0)                         # :661-662

    reward1 = 1 / (1 + (yaw_err * 30) ** 2)                                 # :667
    reward2 = 1 - tilt_err ** 2                                             # :670
    reward4 = 1 - dtilt_err ** 2                                            # :671
    command_energy = torch.sum(1 - (action_history ** 2), dim=-1)           # :673
    reward = reward1 * 5.0 + reward2 * 0.2 + reward4 * 0.3 + command_energy * 0.5   # :675

    felt = torch.abs(tilt) >= max_tilt                                      # :677
    finished = progress_buf >= max_episode_length - 1                       # :678
    reset = torch.where(finished, 1, 0)                                     # :679
    reset = torch.where(felt, 1, reset)                                     # :680
    reward = torch.where(felt, -100.0, reward)                              # :682
    return reward, reset


def euler_to_quaternion(roll: Tensor, pitch: Tensor, yaw: Tensor) -> Tensor:
    """gogoro_new.py:496-502 (xyzw)."""
    sr, cr = torch.sin(roll / 2), torch.cos(roll / 2)
    sp, cp = torch.sin(pitch / 2), torch.cos(pitch / 2)
    sy, cy = torch.sin(yaw / 2), torch.cos(yaw / 2)
    qx = sr * cp * cy - cr * sp * sy
    qy = cr * sp * cy + sr * cp * sy
    qz = cr * cp * sy - sr * sp * cy
    qw = cr * cp * cy + sr * sp * sy
    return torch.stack((qx, qy, qz, qw), dim=-1)


def _uniform(unit: Tensor, bounds) -> Tensor:
    return bounds[0] + unit * (bounds[1] - bounds[0])                       # get_randoms :466-467


def _normal(unit: Tensor, mean_std) -> Tensor:
    return mean_std[0] + unit * mean_std[1]                                 # get_randoms_norm :469-470


# --------------------------------------------------------------------------------------
# Task-level restatement over plain tensors (fake-sim semantics, noise passed in)
# --------------------------------------------------------------------------------------

class GogoroOracle:
    """State + method bodies of ``Gogoro`` (gogoro_new.py:32-150, 347-389, 424-590) without isaacgym.

    OUT OF SCOPE and therefore absent: ``apply_randomizations`` (:476), the per-env
    ``set_actor_dof_properties`` loops incl. the damping draw (:560-578,595-601), the
    viewer debug lines (:392-420), the terrain branch (``USE_TERAIN`` False, :26).
    The constructor does NOT run the reference's initial ``reset_idx(all)`` (:150);
    callers do it explicitly so that the noise is visible.
    """

    def __init__(self, consts: GogoroConsts, root_tensor: Tensor, state_dof: Tensor, thormang_pose: Tensor,
                 curent_speed: Optional[Tensor] = None):
        self.c = consts
        n = root_tensor.shape[0]
        self.n_envs = n
        self.num_dof = thormang_pose.shape[1]
        self.root_tensor = root_tensor
        self.state_dof = state_dof
        self.dof_pos = state_dof.view(n, self.num_dof, 2)[..., 0]           # :137
        self.dof_vel = state_dof.view(n, self.num_dof, 2)[..., 1]           # :138
        self.thormang_pose = thormang_pose                                   # :234,262
        self.root_reset_tensor = root_tensor.clone().detach()                # :144
        self.root_reset_tensor[:, 7:13] = 0                                  # :145
        self.imu_offsets = torch.zeros(n)                                    # :70
        self.steer_offsets = torch.zeros(n)                                  # :71
        self.curent_speed_offset = torch.zeros(n)                            # :72
        self.config_vector = torch.zeros((n, 5))                             # :78
        self.yaw_command = torch.zeros(n)                                    # :81
        self.curent_speed = torch.zeros(n) if curent_speed is None else curent_speed   # :90
        self.curent_command = torch.zeros(n)                                 # :91
        self.action_history = torch.zeros((n, HISTORY))                      # :92
        self.buffer_obs = torch.zeros((n, 1, NUM_OBS))                       # :103-105
        self.max_episode_length = int(consts.max_steps)                      # :39
        self.obs_buf = torch.zeros((n, NUM_OBS))
        self.rew_buf = torch.zeros(n)
        self.reset_buf = torch.ones(n, dtype=torch.long)
        self.timeout_buf = torch.zeros(n, dtype=torch.long)
        self.progress_buf = torch.zeros(n, dtype=torch.long)
        self.dof_pos_command = torch.zeros((n, self.num_dof))
        self.dof_speed_command = torch.zeros((n, self.num_dof))

    # gogoro_new.py:347-369; steer_unit: N(0,1) draw [N]
    def pre_physics_step(self, actions: Tensor, steer_unit: Tensor):
        c = self.c
        self.action_history[:, :-1] = self.action_history[:, 1:].clone()     # :349
        self.action_history[:, -1] = actions[:, 0]                           # :350
        actions = actions * c.max_steering_change                            # :353
        actions = torch.clamp(actions, -c.max_steering_change, c.max_steering_change)   # :354
        self.curent_command += actions[:, 0]                                 # :355
        self.curent_command = torch.clamp(self.curent_command, -c.max_steering, c.max_steering)   # :359
        dof_pos_command = torch.zeros(self.n_envs, self.num_dof)             # :360
        noise = _normal(steer_unit, c.steering_action_noise)                 # :362
        dof_pos_command[:, c.steering_joint] = self.curent_command + self.steer_offsets + noise   # :363
        dof_speed_command = torch.zeros((self.n_envs, self.num_dof))         # :366
        dof_speed_command[:, c.rear_wheel_joint] = self.curent_speed         # :367
        self.dof_pos_command, self.dof_speed_command = dof_pos_command, dof_speed_command

    # gogoro_new.py:424-462; sensor_unit: N(0,1) draws [N,5], column order = order of the draws (c0,c1,c2,c3,c4)
    def compute_obs_rwd(self, sensor_unit: Tensor):
        c = self.c
        obs_ = gogoro_observations(self.root_tensor.clone(), self.yaw_command.clone(),
                                   self.curent_command.clone())             # :430-434
        self.buffer_obs[:, :-1] = self.buffer_obs[:, 1:].clone()             # :436
        self.buffer_obs[:, -1] = obs_.clone()                                # :437
        self.rew_buf[:], self.reset_buf[:] = gogoro_reward(
            self.buffer_obs, self.progress_buf, self.action_history, self.max_episode_length)   # :439-444
        r = self.buffer_obs.clone()                                          # :449
        r[:, -1, 0] += _normal(sensor_unit[:, 0], c.imu_filter_noise) + self.imu_offsets   # :451
        r[:, -1, 1] += _normal(sensor_unit[:, 1], c.imu_noise)               # :452
        r[:, -1, 2] += _normal(sensor_unit[:, 2], c.imu_noise)               # :453
        r[:, -1, 3] += _normal(sensor_unit[:, 3], c.speed_sensor_noise)      # :455
        r[:, -1, 3] += self.curent_speed_offset                              # :456
        r[:, -1, 3] = torch.clamp(r[:, -1, 4], 0, 5.0)                       # :457 (reads channel 4, sic)
        r[:, -1, 3] = torch.round(r[:, -1, 4])                               # :458 (reads channel 4, sic)
        r[:, -1, 4] += _normal(sensor_unit[:, 4], c.imu_filter_noise)        # :460
        self.obs_buf = r.flatten(start_dim=1)                                # :462

    # gogoro_new.py:505-590 (+ randomize :474-482, generate_spawn_r :485-492).
    # unit: [K, 10] draws, row k for env_ids[k]; columns in the order the reference draws them:
    #   0 U speed (:479)   1 N steer_offsets (:481, overwritten at :575)   2 U speed sensor offset (:482)
    #   3 U initial_target (:486)   4 U initial_rot offset (:487)   5..9 N config_vector[0..4] (:554-559)
    def reset_idx(self, env_ids: Tensor, unit: Tensor):
        c = self.c
        if len(env_ids) == 0:                                                # :507-509
            return
        self.curent_speed[env_ids] = _uniform(unit[:, 0], c.speed_range)     # :479
        self.steer_offsets[env_ids] = _normal(unit[:, 1], c.steering_offset)  # :481
        self.curent_speed_offset[env_ids] = _uniform(unit[:, 2], c.speed_sensor_offset)   # :482
        initial_target = (unit[:, 3] * 2 - 1) * np.pi                        # :486
        initial_rot = initial_target + _uniform(unit[:, 4], [-1.57, 1.57])   # :487
        zeros = torch.zeros(len(env_ids))
        r = euler_to_quaternion(zeros, zeros, initial_rot)                   # :488-490
        test = self.root_reset_tensor.clone()                                # :518
        test[env_ids, 3:7] = r                                               # :519
        test[env_ids, 7:13] = 0.0                                            # :520
        test[env_ids, 2] = 0.03                                              # :537
        self.root_tensor[env_ids] = test[env_ids]                            # :547 (fake sim row copy)
        self.dof_pos[env_ids] = self.thormang_pose[env_ids]                  # :549
        self.dof_vel[env_ids] = 0.0                                          # :550
        self.config_vector[env_ids, 0] = _normal(unit[:, 5], c.seat_offset_x_range)    # :554
        self.config_vector[env_ids, 1] = _normal(unit[:, 6], c.seat_offset_y_range)    # :555
        self.config_vector[env_ids, 2] = _normal(unit[:, 7], c.seat_offset_z_range)    # :556
        self.config_vector[env_ids, 3] = _normal(unit[:, 8], c.seat_offset_xr_range)   # :557
        self.config_vector[env_ids, 4] = _normal(unit[:, 9], c.steering_offset)        # :559
        self.imu_offsets[env_ids] = self.config_vector[env_ids, 3]           # :574
        self.steer_offsets[env_ids] = self.config_vector[env_ids, 4]         # :575
        self.progress_buf[env_ids] = 0                                       # :583
        self.reset_buf[env_ids] = 0                                          # :584
        self.obs_buf[env_ids] = 0                                            # :585
        self.buffer_obs[env_ids] = 0.0                                       # :586
        self.curent_command[env_ids] = 0.0                                   # :587
        self.yaw_command[env_ids] = initial_target.clone()                   # :589
        self.action_history[env_ids] = 0.0                                   # :590

    # gogoro_new.py:373-389. reset_unit [K,10] compact; sensor_unit [N,5]; speed_unit [K'] compact
    # (envs whose progress == speed_freq_update, ascending); yaw_unit [N] dense (:387 draws n_envs values)
    def post_physics_step(self, sensor_unit: Tensor, reset_unit: Optional[Tensor] = None,
                          speed_unit: Optional[Tensor] = None, yaw_unit: Optional[Tensor] = None):
        c = self.c
        self.progress_buf += 1                                               # :375
        env_ids = self.reset_buf.nonzero(as_tuple=False).squeeze(-1)         # :377
        if len(env_ids) > 0:
            self.reset_idx(env_ids, reset_unit[:len(env_ids)])               # :379
        self.compute_obs_rwd(sensor_unit)                                    # :382
        speed_change = (self.progress_buf == c.speed_freq_update).nonzero(as_tuple=False).flatten()   # :384
        yaw_change = (self.progress_buf == c.yaw_freq_update).nonzero(as_tuple=False).flatten()       # :385
        if len(speed_change) > 0:
            self.curent_speed[speed_change] = _uniform(speed_unit[:len(speed_change)], c.speed_range)  # :386
        if len(yaw_change) > 0:
            self.yaw_command[yaw_change] = _uniform(yaw_unit, [-np.pi, np.pi])[yaw_change]            # :387
        self.yaw_command = torch.where(self.yaw_command > np.pi, self.yaw_command - np.pi * 2, self.yaw_command)    # :388
        self.yaw_command = torch.where(self.yaw_command < -np.pi, self.yaw_command + np.pi * 2, self.yaw_command)   # :389
        return env_ids, speed_change, yaw_change

    # vec_task.py:313-359; clipActions/clipObservations default to inf for Gogoro (Gogoro.yaml has neither key)
    def step(self, actions: Tensor, steer_unit: Tensor, sensor_unit: Tensor, reset_unit=None,
             speed_unit=None, yaw_unit=None, physics=None):
        self.pre_physics_step(actions, steer_unit)
        if physics is not None:
            physics(self)
        self.post_physics_step(sensor_unit, reset_unit, speed_unit, yaw_unit)
        self.timeout_buf = (self.progress_buf >= self.max_episode_length - 1) & (self.reset_buf != 0)
        return self.obs_buf, self.rew_buf, self.reset_buf, {"time_outs": self.timeout_buf}
