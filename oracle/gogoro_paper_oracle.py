"""CPU restatement of the Gogoro 8-channel x 20-frame history variant.  TEST INFRASTRUCTURE.

Follows ``/root/reference/isaacgymenvs/tasks/gogoro_realistic_turning_sim_paper.py`` (file:line cited
per statement; constants from ``isaacgymenvs/cfg/task/Gogoro_paper.yaml``) -- the configuration used
for the paper runs.  The JIT-level functions are pinned bit-for-bit against the reference's own lifted
TorchScript functions (tests/test_oracle_vs_reference.py, tests/golden/gogoro_paper_jit.npz); helpers
from ``oracle.isaac_helpers`` are a restated un-vendored module (parity unpinned there).

All random draws are inputs (U(0,1) "unit" tensors); the affine map applied to them is the reference's
``get_randoms`` (:549-550).  Module flags of the reference file as saved (:22-32): DEBUG = True, hence
USE_STEER_DELAY = False (fixed delay line tap ``-3``, :357), PUSH_ROBOT = True, CENTER_ROBOT = True,
RANDOM_DAMPING = False, DEBUG_START_SPEED = True, DEBUGUSETERRAIN = False.  ``use_steer_delay`` restores the
per-env tap (:355) for the non-debug setting.

OUT OF SCOPE and absent: the PUSH_ROBOT block (:441-486: a PhysX rigid-body force tensor, hard-coded
``[:2048]``), the viewer lines (:415-438), the per-env ``set_actor_dof_properties`` calls (:657-675), the
terrain branch.

Reference quirk kept as its evident intent: ``speed_command_change[speed_command_change]`` (:397-400) indexes
the hit list with itself; it is the identity when the hit set is a prefix ``0..k-1`` of the env ids and raises
an IndexError otherwise.  Here the hit set itself is used.
"""
from dataclasses import dataclass, field
from typing import List, Optional, Tuple

import numpy as np
import torch
from torch import Tensor

from . import isaac_helpers as ih

NUM_DOF = 39
NUM_CH = 8            # :113 num_obs
BUFF = 20             # :116 buff_size
NUM_OBS = NUM_CH * BUFF
RESET_NOISE_COLS = 5  # see GogoroPaperOracle.reset_idx
SENSOR_NOISE_COLS = 6


@dataclass
class GogoroPaperConsts:
    """cfg["env"], cfg["noises"], cfg["sim"] values read at :47-84 (Gogoro_paper.yaml:7-58)."""
    max_steps: int = 3600
    dt: float = 0.03
    imu_filter_noise: List[float] = field(default_factory=lambda: [-0.003, 0.003])
    imu_noise: List[float] = field(default_factory=lambda: [-0.003, 0.003])
    speed_sensor_noise: List[float] = field(default_factory=lambda: [0, 0.3])
    speed_sensor_offset: List[float] = field(default_factory=lambda: [-0.3, 0.3])
    imu_x_offset: List[float] = field(default_factory=lambda: [-0.02, 0.02])
    steering_offset: List[float] = field(default_factory=lambda: [-0.05, 0.05])
    command_delay: List[int] = field(default_factory=lambda: [0, 5])
    speed_range: List[float] = field(default_factory=lambda: [5.0, 20.0])
    speed_freq_update: int = 300
    yaw_freq_update: int = 300
    max_steering: float = 0.5             # :86
    use_steer_delay: bool = False         # USE_STEER_DELAY = not DEBUG (:31)
    start_speed: float = 1.3              # DEBUG_START_SPEED (:26, :638-641)
    steering_joint: int = 0
    rear_wheel_joint: int = 1

    @staticmethod
    def from_cfg(cfg: dict, steering_joint: int = 0, rear_wheel_joint: int = 1) -> "GogoroPaperConsts":
        nz = cfg["noises"]
        return GogoroPaperConsts(
            max_steps=cfg["env"]["max_steps"], dt=cfg["sim"]["dt"],
            imu_filter_noise=list(nz["imu_filter_noise"]), imu_noise=list(nz["imu_noise"]),
            speed_sensor_noise=list(nz["speed_sensor_noise"]), speed_sensor_offset=list(nz["speed_sensor_offset"]),
            imu_x_offset=list(nz["imu_x_offset"]), steering_offset=list(nz["steering_offset"]),
            command_delay=list(nz["command_delay"]), speed_range=list(nz["speed_range"]),
            speed_freq_update=nz["speed_freq_update"], yaw_freq_update=nz["yaw_freq_update"],
            use_steer_delay=bool(cfg["env"].get("useSteerDelay", False)),
            start_speed=float(cfg["env"].get("startSpeed", 1.3)),
            steering_joint=steering_joint, rear_wheel_joint=rear_wheel_joint)


# --------------------------------------------------------------------------------------
# JIT-level functions
# --------------------------------------------------------------------------------------

def shortest_angle_distance(yaw1: Tensor, yaw2: Tensor) -> Tensor:
    """:765-767 (torch ``%`` = floor-mod)."""
    return (yaw2 - yaw1 + torch.pi) % (2 * torch.pi) - torch.pi


def paper_observations(root_state: Tensor, desired_yaw: Tensor, curent_command: Tensor, steer_delay: Tensor) -> Tensor:
    """:770-807 ``compute_gogoro_observations`` -> [N,8] = roll, yaw, d_roll, d_yaw, speed, delta_yaw, command, delay."""
    quat = root_state[:, 3:7]
    roll, _, yaw = ih.get_euler_xyz(quat)                                   # :775
    lin_b = ih.quat_rotate_inverse(quat, root_state[:, 7:10])               # :777
    ang_b = ih.quat_rotate_inverse(quat, root_state[:, 10:13])              # :778
    d_roll = ang_b[:, 0]                                                    # :780
    d_yaw = ang_b[:, 2]                                                     # :781
    roll = torch.where(roll > np.pi, roll - np.pi * 2, roll)                # :783
    roll = torch.where(roll < -np.pi, roll + np.pi * 2, roll)               # :784
    yaw = torch.where(yaw > np.pi, yaw - np.pi * 2, yaw)                    # :786
    yaw = torch.where(yaw < -np.pi, yaw + np.pi * 2, yaw)                   # :787
    delta_yaw = shortest_angle_distance(yaw, desired_yaw)                   # :789
    speed = lin_b[:, 0]                                                     # :791
    cols = [roll, yaw, d_roll, d_yaw, speed, delta_yaw, curent_command, steer_delay]   # :793-804
    return torch.cat([c.unsqueeze(-1) for c in cols], dim=1)


def paper_reward(act_buff: Tensor, curent_obs: Tensor, progress_buf: Tensor, max_episode_length: int
                 ) -> Tuple[Tensor, Tensor]:
    """:713-762 ``compute_gogoro_reward``.  ``reward7`` (:747) is a mean over EVERY env and frame pair."""
    tilt = curent_obs[:, -1, 0]
    dtilt = curent_obs[:, -1, 2]
    yaw_err = curent_obs[:, -1, 5]
    act_buff_ = act_buff / 0.5                                              # :722
    max_tilt = 0.38                                                         # :724
    tilt_err = torch.clamp(tilt / max_tilt, -1.0, 1.0)                      # :725-726
    yaw_err = torch.clamp(yaw_err / np.pi, -1.0, 1.0)                       # :728-729
    dtilt_err = torch.clamp(dtilt / 0.3, -1.0, 1.0)                         # :731-732
    reward1 = 1 - yaw_err ** 2                                              # :737
    reward2 = 1 - tilt_err ** 2                                             # :738
    reward4 = 1 - dtilt_err ** 2                                            # :739
    tilt_w = 1 - torch.tanh(50 * tilt_err ** 2)                             # :742
    dtilt_w = 1 - torch.tanh(50 * dtilt_err ** 2)                           # :743
    reward5 = 1 - ((act_buff_[:, -1] ** 2) * (tilt_w * dtilt_w))            # :745
    reward7 = 1 - torch.mean(torch.diff(act_buff_, dim=1) ** 2)             # :747 (scalar: all envs)
    reward = reward1 * 0.45 + reward2 * 0.1 + reward4 * 0.35 + reward5 * 2.0 + reward7 * 0.2   # :753
    finished = progress_buf >= max_episode_length - 1                       # :755
    felt = torch.abs(tilt) >= max_tilt                                      # :757
    reset = torch.where(finished, 1, 0)                                     # :758
    reset = torch.where(felt, 1, reset)                                     # :759
    reward = torch.clip(reward, 0., None)                                   # :760
    reward = torch.where(felt, -1.0, reward)                                # :761
    return reward, reset


def _uniform(unit: Tensor, bounds) -> Tensor:
    return bounds[0] + unit * (bounds[1] - bounds[0])                       # get_randoms :549-550


# --------------------------------------------------------------------------------------
# Task-level restatement over plain tensors (fake-sim semantics, noise passed in)
# --------------------------------------------------------------------------------------

class GogoroPaperOracle:
    """State + method bodies of the paper variant's ``Gogoro`` (:36-165, 342-378, 382-412, 490-545, 557-566,
    570-688) without isaacgym.  The constructor does NOT run the initial ``reset_idx(all)`` (:165)."""

    def __init__(self, consts: GogoroPaperConsts, root_tensor: Tensor, state_dof: Tensor, thormang_pose: Tensor):
        self.c = consts
        n = root_tensor.shape[0]
        self.n_envs = n
        self.num_dof = thormang_pose.shape[1]
        self.root_tensor = root_tensor
        self.state_dof = state_dof
        self.dof_pos = state_dof.view(n, self.num_dof, 2)[..., 0]           # :148
        self.dof_vel = state_dof.view(n, self.num_dof, 2)[..., 1]           # :149
        self.thormang_pose = thormang_pose
        self.root_reset_tensor = root_tensor.clone().detach()                # :155
        self.root_reset_tensor[:, 7:13] = 0                                  # :156
        self.yaw_command = torch.zeros(n)                                    # :81
        self.curent_speed = torch.zeros(n)                                   # :88
        self.curent_command = torch.zeros(n)                                 # :90
        self.steer_offsets = torch.zeros(n)                                  # :91
        self.steer_delay = torch.zeros(n, dtype=torch.long)                  # :92
        self.command_history = torch.zeros((n, consts.command_delay[1]))     # :93
        self.curent_speed_offset = torch.zeros(n)                            # :96
        self.curent_imu_x_offset = torch.zeros(n)                            # :97
        self.speed_no_noise = torch.zeros(n)                                 # :100
        self.buffer_obs = torch.zeros((n, BUFF, NUM_CH))                     # :114
        self.buffer_obs_noisy = torch.zeros((n, BUFF, NUM_CH))               # :115
        self.max_episode_length = int(consts.max_steps)                      # :47
        self.obs_buf = torch.zeros((n, NUM_OBS))
        self.rew_buf = torch.zeros(n)
        self.reset_buf = torch.ones(n, dtype=torch.long)
        self.timeout_buf = torch.zeros(n, dtype=torch.long)
        self.progress_buf = torch.zeros(n, dtype=torch.long)
        self.dof_pos_command = torch.zeros((n, self.num_dof))
        self.dof_speed_command = torch.zeros((n, self.num_dof))

    # :342-378
    def pre_physics_step(self, actions: Tensor):
        c = self.c
        actions = torch.clamp(actions, -1, 1)                                # :344
        self.curent_command = actions[:, 0] * c.max_steering                 # :345
        self.command_history[:, :-1] = self.command_history[:, 1:].clone()   # :352
        self.command_history[:, -1] = self.curent_command                    # :353
        idx = torch.arange(self.n_envs)
        if c.use_steer_delay:
            applied_command = self.command_history[idx, -self.steer_delay]   # :355 (delay 0 -> column 0, sic)
        else:
            applied_command = self.command_history[idx, -3]                  # :357
        dof_pos_command = torch.zeros(self.n_envs, self.num_dof)             # :359
        dof_pos_command[:, c.steering_joint] = applied_command               # :360
        dof_speed_command = torch.zeros((self.n_envs, self.num_dof))         # :375
        dof_speed_command[:, c.rear_wheel_joint] = self.curent_speed         # :376
        self.dof_pos_command, self.dof_speed_command = dof_pos_command, dof_speed_command

    # :490-545; sensor_unit: U(0,1) draws [N,6] in draw order: (N,2) imu_filter :525, (N,2) imu_noise :527,
    # (N) speed noise :528, (N) imu_filter :531
    def compute_obs_rwd(self, sensor_unit: Tensor):
        c = self.c
        delay = (self.steer_delay.clone() - c.command_delay[0]) / (c.command_delay[1] - c.command_delay[0])   # :497
        obs_ = paper_observations(self.root_tensor.clone(), self.yaw_command.clone(), self.curent_command.clone(), delay)
        self.buffer_obs[:, :-1] = self.buffer_obs[:, 1:].clone()             # :502
        self.buffer_obs[:, -1] = obs_.clone()                                # :503
        self.buffer_obs_noisy[:, :-1] = self.buffer_obs_noisy[:, 1:].clone()  # :505
        self.buffer_obs_noisy[:, -1] = obs_.clone()                          # :506
        self.rew_buf[:], self.reset_buf[:] = paper_reward(self.buffer_obs[:, :, 6], self.buffer_obs, self.progress_buf,
                                                          self.max_episode_length)   # :508-514
        self.speed_no_noise = obs_[:, 4].clone()                             # :521
        nz = self.buffer_obs_noisy
        nz[:, -1, 0:2] += _uniform(sensor_unit[:, 0:2], c.imu_filter_noise)  # :525
        nz[:, -1, 0] += self.curent_imu_x_offset                             # :526
        nz[:, -1, 2:4] += _uniform(sensor_unit[:, 2:4], c.imu_noise)         # :527
        nz[:, -1, 4] += _uniform(sensor_unit[:, 4], c.speed_sensor_noise)    # :528
        nz[:, -1, 4] += self.curent_speed_offset                             # :529
        nz[:, -1, 4] = torch.where(nz[:, -1, 4] < 0, 0, nz[:, -1, 4])        # :530
        nz[:, -1, 5] += _uniform(sensor_unit[:, 5], c.imu_filter_noise)      # :531
        nz[:, -1, 0:2] /= np.pi                                              # :534
        nz[:, -1, 2:4] /= 3.0                                                # :535
        nz[:, -1, 4] /= 5.0                                                  # :536
        nz[:, -1, 5] /= np.pi                                                # :537
        nz[:, -1, 6] /= c.max_steering                                       # :538
        nz[:, -1, 2] += self.buffer_obs[:, -2, 6] - self.buffer_obs[:, -1, 6]            # :541
        nz[:, -1, 0] += (self.buffer_obs[:, -2, 6] - self.buffer_obs[:, -1, 6]) * 0.3    # :542
        nz[:, :, 1] = 0.0                                                    # :544
        self.obs_buf = nz.flatten(start_dim=1)                               # :545

    # :599-688 (+ randomize :557-566, generate_spawn_r :570-587).
    # unit: [K,5] U(0,1) draws, row k for env_ids[k]; columns in the order the reference draws them:
    #   0 speed (:562)  1 steer_delay (:563)  2 steer_offsets (:564)  3 speed-sensor offset (:566)  4 imu_x offset (:655)
    def reset_idx(self, env_ids: Tensor, unit: Tensor):
        c = self.c
        if len(env_ids) == 0:                                                # :604-606
            return
        self.curent_speed[env_ids] = _uniform(unit[:, 0], c.speed_range)     # :562
        self.steer_delay[env_ids] = _uniform(unit[:, 1], c.command_delay).to(torch.long)   # :563
        self.steer_offsets[env_ids] = _uniform(unit[:, 2], c.steering_offset)              # :564
        self.curent_speed_offset[env_ids] = _uniform(unit[:, 3], c.speed_sensor_offset)    # :566
        k = len(env_ids)
        initial_target = torch.zeros(k)                                      # :571
        initial_rot = torch.zeros(k)                                         # :572
        from .gogoro_oracle import euler_to_quaternion                       # :590-596, same body as gogoro_new.py:496-502
        r = euler_to_quaternion(torch.zeros(k), torch.zeros(k), initial_rot)  # :573-575
        test = self.root_reset_tensor.clone()                                # :618
        test[env_ids, 2] = 0.03                                              # :619
        test[env_ids, 3:7] = r                                               # :635
        test[env_ids, 7:13] = 0.0                                            # :636
        if c.start_speed != 0.0:                                             # :638-641 DEBUG_START_SPEED
            test[env_ids, 7] = c.start_speed * torch.cos(initial_rot)
            test[env_ids, 8] = c.start_speed * torch.sin(initial_rot)
        self.root_tensor[env_ids] = test[env_ids]                            # :645 (fake sim row copy)
        self.dof_pos[env_ids] = self.thormang_pose[env_ids]                  # :647
        self.dof_vel[env_ids] = 0.0                                          # :648
        self.curent_imu_x_offset[env_ids] = _uniform(unit[:, 4], c.imu_x_offset)   # :653-655 (per-id loop)
        self.progress_buf[env_ids] = 0                                       # :678
        self.reset_buf[env_ids] = 0                                          # :679
        self.obs_buf[env_ids] = 0                                            # :680
        self.buffer_obs[env_ids] = 0.0                                       # :681
        self.buffer_obs_noisy[env_ids] = 0.0                                 # :682
        self.curent_command[env_ids] = 0.0                                   # :683
        self.yaw_command[env_ids] = initial_target.clone()                   # :685
        self.command_history[env_ids] = 0.0                                  # :686
        self.speed_no_noise[env_ids] = 0.0                                   # :688

    # :382-412.  reset_unit [K,5] compact; sensor_unit [N,6]; speed_unit [K'] compact (envs whose progress ==
    # speed_freq_update, ascending); yaw_unit [N] dense (:410 draws n_envs values)
    def post_physics_step(self, sensor_unit: Tensor, reset_unit: Optional[Tensor] = None,
                          speed_unit: Optional[Tensor] = None, yaw_unit: Optional[Tensor] = None):
        c = self.c
        self.progress_buf += 1                                               # :384
        env_ids = self.reset_buf.nonzero(as_tuple=False).squeeze(-1)         # :385
        if len(env_ids) > 0:
            self.reset_idx(env_ids, reset_unit[:len(env_ids)])               # :387
        self.compute_obs_rwd(sensor_unit)                                    # :390
        speed_change = (self.progress_buf == c.speed_freq_update).nonzero(as_tuple=False).flatten()   # :393
        yaw_change = (self.progress_buf == c.yaw_freq_update).nonzero(as_tuple=False).flatten()       # :394
        # :397-400 index each list with itself (identity on prefix hit sets, IndexError otherwise): hit set used as is
        if len(speed_change) > 0:
            self.curent_speed[speed_change] = _uniform(speed_unit[:len(speed_change)], c.speed_range)  # :407
        if len(yaw_change) > 0:
            self.yaw_command[yaw_change] = _uniform(yaw_unit, [-np.pi, np.pi])[yaw_change]            # :410
        self.yaw_command = torch.where(self.yaw_command > np.pi, self.yaw_command - np.pi * 2, self.yaw_command)    # :411
        self.yaw_command = torch.where(self.yaw_command < -np.pi, self.yaw_command + np.pi * 2, self.yaw_command)   # :412
        return env_ids, speed_change, yaw_change

    # vec_task.py:313-359; clipActions / clipObservations default to inf (Gogoro_paper.yaml has neither key)
    def step(self, actions: Tensor, sensor_unit: Tensor, reset_unit=None, speed_unit=None, yaw_unit=None, physics=None):
        self.pre_physics_step(actions)
        if physics is not None:
            physics(self)
        self.post_physics_step(sensor_unit, reset_unit, speed_unit, yaw_unit)
        self.timeout_buf = (self.progress_buf >= self.max_episode_length - 1) & (self.reset_buf != 0)
        return self.obs_buf, self.rew_buf, self.reset_buf, {"time_outs": self.timeout_buf}
