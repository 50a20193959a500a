"""CPU restatement of ``isaacgymenvs/tasks/gogoro_realistic_turning.py``.
TEST INFRASTRUCTURE.

The file is the sibling of ``gogoro_realistic_turning_sim_paper.py`` (same ``Gogoro`` class skeleton, buffers, cfg
keys); everything the two share is inherited from ``oracle.gogoro_paper_oracle`` and only what differs is restated
here, with this file's line numbers.  ``turning_reward`` and the observation function are pinned bit-for-bit
against the reference's own lifted TorchScript functions (tests/test_oracle_vs_reference.py,
tests/golden/gogoro_turning_jit.npz).

Module flags as saved (:22-32): DEBUG = False, hence USE_STEER_DELAY = True (:31), CENTER_ROBOT = False,
PUSH_ROBOT = False, RANDOM_DAMPING = True, DEBUG_START_SPEED = True, IGNORE_ZERO = False.

OUT OF SCOPE and absent (simulator side): ``set_env_dof_prop`` / ``set_actor_dof_properties`` per reset env
(:655-674: RANDOM_DAMPING and the seat offsets draw from ``get_randoms`` there and feed only PhysX properties),
viewer lines (:407-430), the terrain branch.
"""
from typing import Optional, Tuple

import numpy as np
import torch
from torch import Tensor

from . import gogoro_paper_oracle as gp
from .gogoro_paper_oracle import GogoroPaperConsts, _uniform, paper_observations, shortest_angle_distance  # noqa: F401

RESET_NOISE_COLS = 8


def turning_consts(**kw) -> GogoroPaperConsts:
    """Same cfg keys as the paper file (:44-84); USE_STEER_DELAY = not DEBUG = True (:31)."""
    kw.setdefault("use_steer_delay", True)
    return GogoroPaperConsts(**kw)


def turning_reward(act_buff: Tensor, curent_obs: Tensor, progress_buf: Tensor, max_episode_length: int
                   ) -> Tuple[Tensor, Tensor]:
    """:713-765 ``compute_gogoro_reward``.  ``reward5`` (:742) is a mean over EVERY env and frame."""
    tilt = curent_obs[:, -1, 0]                                             # :714
    dtilt = curent_obs[:, -1, 2]                                            # :715
    yaw_err = curent_obs[:, -1, 5]                                          # :718
    max_tilt = 0.38                                                         # :721
    tilt_err = torch.clamp(tilt / max_tilt, -1.0, 1.0)                      # :722-723
    yaw_err = torch.clamp(yaw_err / np.pi, -1.0, 1.0)                       # :725-726
    dtilt_err = torch.clamp(dtilt, -1.0, 1.0)                               # :728-729
    reward1 = 1 / (1 + yaw_err ** 2 * 200)                                  # :734
    reward2 = 1 / (1 + tilt_err ** 2 * 100)                                 # :735
    reward4 = 1 / (1 + dtilt_err ** 2 * 100)                                # :736
    min_, _ = torch.min(act_buff, dim=1)                                    # :738
    max_, _ = torch.min(act_buff, dim=1)                                    # :739 (sic: min again)
    action_err = min_ - max_                                                # :740, identically 0
    reward3 = 1 / (1 + action_err ** 2 * 500)                               # :741
    reward5 = 1 - torch.abs(torch.mean(act_buff))                           # :742 (scalar: all envs, all frames)
    reward = reward1 * 3.0 + reward2 * 1.0 + reward3 * 10.0 + reward4 * 2.0 + reward5 * 10.0   # :744
    finished = progress_buf >= max_episode_length - 1                       # :755
    felt = torch.abs(tilt) >= max_tilt                                      # :756
    reset = torch.where(finished, 1, 0)                                     # :758
    reset = torch.where(felt, 1, reset)                                     # :759
    reward = torch.clip(reward, 0., None)                                   # :761
    reward = torch.where(felt, -1.0, reward)                                # :762
    return reward, reset


def euler_to_quaternion(roll: Tensor, pitch: Tensor, yaw: Tensor) -> Tensor:
    """:592-598."""
    qx = torch.sin(roll / 2) * torch.cos(pitch / 2) * torch.cos(yaw / 2) - torch.cos(roll / 2) * torch.sin(pitch / 2) * torch.sin(yaw / 2)
    qy = torch.cos(roll / 2) * torch.sin(pitch / 2) * torch.cos(yaw / 2) + torch.sin(roll / 2) * torch.cos(pitch / 2) * torch.sin(yaw / 2)
    qz = torch.cos(roll / 2) * torch.cos(pitch / 2) * torch.sin(yaw / 2) - torch.sin(roll / 2) * torch.sin(pitch / 2) * torch.cos(yaw / 2)
    qw = torch.cos(roll / 2) * torch.cos(pitch / 2) * torch.cos(yaw / 2) + torch.sin(roll / 2) * torch.sin(pitch / 2) * torch.sin(yaw / 2)
    return torch.stack((qx, qy, qz, qw), dim=-1)


class GogoroTurningOracle(gp.GogoroPaperOracle):
    """``Gogoro`` of gogoro_realistic_turning.py: the paper oracle with this file's four differences."""

    # :345-383
    def pre_physics_step(self, actions: Tensor):
        c = self.c
        actions = torch.clamp(actions, -1, 1)                                # :349
        self.curent_command = actions[:, 0] * c.max_steering                 # :350
        self.command_history[:, :-1] = self.command_history[:, 1:].clone()   # :357
        self.command_history[:, -1] = self.curent_command                    # :358
        idx = torch.arange(self.n_envs)
        if c.use_steer_delay:
            applied_command = self.command_history[idx, -self.steer_delay]   # :360 (delay 0 -> column 0, sic)
        else:
            applied_command = self.command_history[idx, -3]                  # :362
        dof_pos_command = torch.zeros(self.n_envs, self.num_dof)             # :364
        dof_pos_command[:, c.steering_joint] = applied_command + self.steer_offsets   # :365
        dof_speed_command = torch.zeros((self.n_envs, self.num_dof))         # :380
        dof_speed_command[:, c.rear_wheel_joint] = self.curent_speed         # :381
        self.dof_pos_command, self.dof_speed_command = dof_pos_command, dof_speed_command

    # :480-547; sensor_unit as in the paper oracle (same draws in the same order, :518-524)
    def compute_obs_rwd(self, sensor_unit: Tensor):
        c = self.c
        delay = (self.steer_delay.clone() - c.command_delay[0]) / (c.command_delay[1] - c.command_delay[0])   # :489
        obs_ = paper_observations(self.root_tensor.clone(), self.yaw_command.clone(), self.curent_command.clone(), delay)
        self.buffer_obs[:, :-1] = self.buffer_obs[:, 1:].clone()             # :495
        self.buffer_obs[:, -1] = obs_.clone()                                # :496
        self.buffer_obs_noisy[:, :-1] = self.buffer_obs_noisy[:, 1:].clone()  # :498
        self.buffer_obs_noisy[:, -1] = obs_.clone()                          # :499
        self.rew_buf[:], self.reset_buf[:] = turning_reward(self.buffer_obs[:, :, 6], self.buffer_obs, self.progress_buf,
                                                            self.max_episode_length)   # :501-507
        self.speed_no_noise = obs_[:, 4].clone()                             # :514
        nz = self.buffer_obs_noisy
        nz[:, -1, 0:2] += _uniform(sensor_unit[:, 0:2], c.imu_filter_noise)  # :518
        nz[:, -1, 0] += self.curent_imu_x_offset                             # :519
        nz[:, -1, 2:4] += _uniform(sensor_unit[:, 2:4], c.imu_noise)         # :520
        nz[:, -1, 4] += _uniform(sensor_unit[:, 4], c.speed_sensor_noise)    # :521
        nz[:, -1, 4] += self.curent_speed_offset                             # :522
        nz[:, -1, 4] = torch.where(nz[:, -1, 4] < 0, 0, nz[:, -1, 4])        # :523
        nz[:, -1, 5] += _uniform(sensor_unit[:, 5], c.imu_filter_noise)      # :524
        nz[:, -1, 0:2] /= np.pi                                              # :527
        nz[:, -1, 2:4] /= 3.0                                                # :528
        nz[:, -1, 4] /= 5.0                                                  # :529
        nz[:, -1, 5] /= np.pi                                                # :530
        nz[:, -1, 6] /= c.max_steering                                       # :531
        nz[:, :, 1] = 0.0                                                    # :543
        self.obs_buf = nz.flatten(start_dim=1)                               # :547

    # :601-686 (+ randomize :559-568, generate_spawn_r :572-588).
    # unit: [K,8] U(0,1) draws, row k for env_ids[k]:
    #   0 speed (:564)  1 steer_delay (:565)  2 steer_offsets (:566)  3 speed-sensor offset (:568)  4 imu_x offset (:653)
    #   5 yaw target (:573)  6 spawn yaw offset (:574)  7 spawn roll (:575)
    def reset_idx(self, env_ids: Tensor, unit: Tensor):
        c = self.c
        k = len(env_ids)
        if k == 0:
            return
        self.curent_speed[env_ids] = _uniform(unit[:, 0], c.speed_range)     # :564
        self.steer_delay[env_ids] = _uniform(unit[:, 1], c.command_delay).to(torch.long)   # :565
        self.steer_offsets[env_ids] = _uniform(unit[:, 2], c.steering_offset)              # :566
        self.curent_speed_offset[env_ids] = _uniform(unit[:, 3], c.speed_sensor_offset)    # :568
        initial_target = (unit[:, 5] * 2 - 1) * np.pi                        # :573
        initial_rot = initial_target + _uniform(unit[:, 6], [-np.pi, np.pi])  # :574
        r = euler_to_quaternion((unit[:, 7] * 2 - 1) * 0.1, torch.zeros(k), initial_rot)   # :575-577
        test = self.root_reset_tensor.clone()                                # :614
        test[env_ids, 2] = 0.03                                              # :615
        test[env_ids, 3:7] = r                                               # :633
        test[env_ids, 7:13] = 0.0                                            # :634
        if c.start_speed != 0.0:                                             # :636-639 DEBUG_START_SPEED
            test[env_ids, 7] = c.start_speed * torch.cos(initial_rot)
            test[env_ids, 8] = c.start_speed * torch.sin(initial_rot)
        self.root_tensor[env_ids] = test[env_ids]                            # :643 (fake sim row copy)
        self.dof_pos[env_ids] = self.thormang_pose[env_ids]                  # :645
        self.dof_vel[env_ids] = 0.0                                          # :646
        self.curent_imu_x_offset[env_ids] = _uniform(unit[:, 4], c.imu_x_offset)   # :651-653 (per-id loop)
        self.progress_buf[env_ids] = 0                                       # :676
        self.reset_buf[env_ids] = 0                                          # :677
        self.obs_buf[env_ids] = 0                                            # :678
        self.buffer_obs[env_ids] = 0.0                                       # :679
        self.buffer_obs_noisy[env_ids] = 0.0                                 # :680
        self.curent_command[env_ids] = 0.0                                   # :681
        self.yaw_command[env_ids] = initial_target.clone()                   # :683
        self.command_history[env_ids] = 0.0                                  # :684
        self.speed_no_noise[env_ids] = 0.0                                   # :686

    # :387-405: the hit lists are used as they are (no self-indexing in this file)
    def post_physics_step(self, sensor_unit: Tensor, reset_unit: Optional[Tensor] = None,
                          speed_unit: Optional[Tensor] = None, yaw_unit: Optional[Tensor] = None):
        return super().post_physics_step(sensor_unit, reset_unit, speed_unit, yaw_unit)
