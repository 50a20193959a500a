"""``Gogoro`` -- drop-in for the registered scooter task ``isaacgymenvs/tasks/gogoro_new.py::Gogoro``
on the env-step path (Thormang3 rider locked to a riding pose, 39-DOF actor, 6 observations, 1 action).

Same constructor signature, cfg keys (``cfg["env"]``, ``cfg["noises"]``, ``cfg["joints_pos"]``,
``cfg["sim"]``), attribute names (spelling included: ``curent_command`` ...) and method surface as
the reference class (gogoro_new.py:30-601); method bodies are single calls into the sm_100a CUDA
extension (include/thormang_b200.h).  No TorchScript, no CPU fallback.

Left out with PhysX / the graphics stack: ``apply_randomizations``
(:162,476), the per-env ``set_actor_dof_properties`` loops and their damping draw (:560-578,
595-601), the viewer debug lines and their ``.item()`` syncs (:392-420), the terrain branch
(``USE_TERAIN`` False, :26).  ``dof_pos_command`` / ``dof_speed_command`` are persistent [N,D]
tensors whose two live columns are rewritten each step (the other columns stay zero, like the
reference's fresh ``torch.zeros``); ``cfg["env"]["zeroFillTargets"]`` restores the full rewrite.
"""
import ctypes as C
from typing import Optional

import numpy as np
import torch

from .. import _native as nat
from ..sim.fake_sim import FakeSim
from .base.vec_task import VecTask

NUM_OBS = 6
NUM_ACT = 1
HISTORY = 5
RESET_NOISE_COLS = 10


class Gogoro(VecTask):

    def __init__(self, cfg, rl_device, sim_device, graphics_device_id, headless,
                 virtual_screen_capture=False, force_render=False):
        self.curent_step = 0
        self.cfg = cfg
        nz = cfg["noises"]
        self.n_envs = cfg["env"]["numEnvs"]
        self.max_episode_length = int(cfg["env"]["max_steps"])                    # gogoro_new.py:39 (0-d tensor there)
        self.randomization_params = cfg.get("task", {}).get("randomization_params", {})
        # gogoro_new.py:46-68
        self.imu_filter_noise = nz["imu_filter_noise"]
        self.imu_noise = nz["imu_noise"]
        self.speed_sensor_noise = nz["speed_sensor_noise"]
        self.steering_action_noise = nz["steering_action_noise"]
        self.spawn_x_angle = nz.get("spawn_x_angle", [0, 0.05])
        self.speed_range = nz["speed_range"]
        self.speed_freq_update = nz["speed_freq_update"]
        self.yaw_freq_update = nz["yaw_freq_update"]
        self.seat_offset_x_range = nz["seat_offset_x_range"]
        self.seat_offset_y_range = nz["seat_offset_y_range"]
        self.seat_offset_z_range = nz["seat_offset_z_range"]
        self.seat_offset_xR_range = nz["seat_offset_xr_range"]
        self.steering_damping_range = nz.get("steering_damping_range", [100, 1000])
        self.steering_offset = nz["steering_offset"]
        self.speed_sensor_offset = nz["speed_sensor_offset"]
        # gogoro_new.py:84-88
        self.min_speed, self.max_speed = 0.0, 10.0
        self.max_steering, self.max_steering_change = 0.5, 0.2
        self.buff_size = 1                                                        # gogoro_new.py:103
        cfg["env"]["numObservations"] = NUM_OBS * self.buff_size                  # gogoro_new.py:107
        cfg["env"]["numActions"] = NUM_ACT                                        # gogoro_new.py:108

        super().__init__(config=cfg, rl_device=rl_device, sim_device=sim_device,
                         graphics_device_id=graphics_device_id, headless=headless,
                         virtual_screen_capture=virtual_screen_capture, force_render=force_render)

        sim, dev, n, d = self.sim, self.device, self.n_envs, self.num_dof
        f = dict(device=dev, dtype=torch.float)
        self.dt = cfg["sim"]["dt"]
        # gogoro_new.py:70-92
        self.imu_offsets = torch.zeros(n, **f)
        self.steer_offsets = torch.zeros(n, **f)
        self.curent_speed_offset = torch.zeros(n, **f)
        self.config_vector = torch.zeros((n, 5), **f)
        self.yaw_command = torch.zeros(n, **f)
        self.curent_speed = self.get_randoms(n, self.speed_range)
        self.curent_command = torch.zeros(n, **f)
        self.action_history = torch.zeros((n, HISTORY), **f)
        self.buffer_obs = torch.zeros((n, self.buff_size, NUM_OBS), **f)
        # gogoro_new.py:125-145
        self.root_tensor = sim.acquire_actor_root_state_tensor()
        self.state_dof = sim.acquire_dof_state_tensor()
        self.root_positions = self.root_tensor[:, 0:3]
        self.root_orientations = self.root_tensor[:, 3:7]
        self.root_angular_vels = self.root_tensor[:, 10:13]
        self.dof_pos = self.state_dof.view(n, d, 2)[..., 0]
        self.dof_vel = self.state_dof.view(n, d, 2)[..., 1]
        self.root_reset_tensor = self.root_tensor.clone().detach()
        self.root_reset_tensor[:, 7:13] = 0
        self.dof_pos_command = torch.zeros((n, d), **f)
        self.dof_speed_command = torch.zeros((n, d), **f)
        self.zero_fill_targets = bool(cfg["env"].get("zeroFillTargets", False))
        self.episode_return = torch.zeros(n, **f) if cfg["env"].get("episodeStatistics", True) else None
        self.episode_stats = torch.zeros(nat.TK_STATS, device=dev, dtype=torch.float64)
        self.noise_seed = int(cfg.get("seed", 0)) & 0xFFFFFFFFFFFFFFFF
        self.env_id_offset = int(cfg["env"].get("envIdOffset", 0))
        self.step_count = 0        # keys the in-kernel Philox streams together with the seed and the global env id
        self._handle = None
        self._create_native()
        if cfg["env"].get("resetOnInit", True):
            self.reset_idx(torch.arange(0, n, device=dev, dtype=torch.long))       # gogoro_new.py:150

    # ------------------------------------------------------------------------------------------
    def create_sim(self):
        """gogoro_new.py:155-340 without PhysX: injected simulator or the fake-sim shim."""
        n = self.cfg["env"]["numEnvs"]
        sim = self.cfg.get("sim_backend")
        if sim is None:
            names = ["steering_joint", "rear_wheel_joint", "front_wheel_joint", "base_x", "base_y", "base_z"]
            names += sorted(self.cfg.get("joints_pos", {}).keys())
            names += [f"aux_{i}" for i in range(39 - len(names))]
            sim = FakeSim(n, 39, self.device, num_sensor_floats=0, dof_names=names[:39])
        self.sim = sim
        self.num_dof = sim.num_dof
        self.dof_names = list(sim.dof_names)
        self.dof_name_to_id = {k: v for v, k in enumerate(self.dof_names)}        # gogoro_new.py:219
        # gogoro_new.py:234,256-262: joints locked to the riding pose
        self.thormang_pose = torch.zeros(n, self.num_dof, device=self.device)
        for j_name, value in self.cfg.get("joints_pos", {}).items():
            if j_name in self.dof_name_to_id:
                self.thormang_pose[:, self.dof_name_to_id[j_name]] = value + 0.0001 / 2

    def _create_native(self):
        lib = nat.lib()
        p = nat.GogoroParams()
        p.device = self.device_id
        p.num_envs, p.num_dof = self.n_envs, self.num_dof
        p.steering_joint = self.dof_name_to_id["steering_joint"]
        p.rear_wheel_joint = self.dof_name_to_id["rear_wheel_joint"]
        p.max_episode_length = self.max_episode_length
        p.speed_freq_update, p.yaw_freq_update = int(self.speed_freq_update), int(self.yaw_freq_update)
        p.max_steering, p.max_steering_change = self.max_steering, self.max_steering_change
        for name, pair in (("imu_filter_noise", self.imu_filter_noise), ("imu_noise", self.imu_noise),
                           ("speed_sensor_noise", self.speed_sensor_noise),
                           ("steering_action_noise", self.steering_action_noise), ("speed_range", self.speed_range),
                           ("seat_offset_x_range", self.seat_offset_x_range),
                           ("seat_offset_y_range", self.seat_offset_y_range),
                           ("seat_offset_z_range", self.seat_offset_z_range),
                           ("seat_offset_xr_range", self.seat_offset_xR_range),
                           ("steering_offset", self.steering_offset),
                           ("speed_sensor_offset", self.speed_sensor_offset)):
            arr = getattr(p, name)
            arr[0], arr[1] = float(pair[0]), float(pair[1])
        handle = C.c_void_p()
        nat.check(lib.tk_gogoro_create(C.byref(p), C.byref(handle)), "tk_gogoro_create")
        self._handle, self._lib = handle, lib
        self._post_physics = lib.tk_gogoro_post_physics
        self._torch_device = torch.device(self.device)
        self._philox_noise = nat.GogoroNoise()
        self._philox_noise.reset.mode = nat.NOISE_PHILOX
        self._philox_noise_ref = C.byref(self._philox_noise)
        self._bind()

    def _bind(self):
        b = nat.GogoroBuffers()
        for name, t in (
                ("root_tensor", self.root_tensor), ("state_dof", self.state_dof), ("obs_buf", self.obs_buf),
                ("buffer_obs", self.buffer_obs), ("rew_buf", self.rew_buf), ("reset_buf", self.reset_buf),
                ("progress_buf", self.progress_buf), ("timeout_buf", self.timeout_buf),
                ("yaw_command", self.yaw_command), ("curent_command", self.curent_command),
                ("curent_speed", self.curent_speed), ("steer_offsets", self.steer_offsets),
                ("imu_offsets", self.imu_offsets), ("curent_speed_offset", self.curent_speed_offset),
                ("action_history", self.action_history), ("config_vector", self.config_vector),
                ("thormang_pose", self.thormang_pose), ("root_reset_tensor", self.root_reset_tensor),
                ("dof_pos_command", self.dof_pos_command), ("dof_speed_command", self.dof_speed_command),
                ("episode_return", self.episode_return)):
            if t is not None and (not t.is_contiguous() or t.device != torch.device(self.device)):
                raise ValueError(f"{name}: the kernels need a contiguous tensor on {self.device}")
            setattr(b, name, nat.ptr(t))
        nat.check(self._lib.tk_gogoro_bind(self._handle, C.byref(b)), "tk_gogoro_bind")

    def rebind(self):
        self._bind()

    def _stream(self):
        return nat.current_stream_ptr(self.device)

    def __del__(self):
        try:
            if getattr(self, "_handle", None):
                self._lib.tk_gogoro_destroy(self._handle)
                self._handle = None
        except Exception:
            pass

    # gogoro_new.py:466-470
    def get_randoms(self, shape, bounds):
        return bounds[0] + torch.rand(shape, device=self.device) * (bounds[1] - bounds[0])

    def get_randoms_norm(self, shape, mean_cov):
        return mean_cov[0] + torch.randn(shape, device=self.device) * mean_cov[1]

    # ------------------------------------------------------------------------------------------
    def pre_physics_step(self, actions: torch.Tensor, steer_unit: Optional[torch.Tensor] = None):
        """gogoro_new.py:347-369.  ``steer_unit``: the N(0,1) draw of :362 (in-kernel Philox when omitted)."""
        a = actions.to(self.device, dtype=torch.float).contiguous()
        nat.check(self._lib.tk_gogoro_pre_physics(self._handle, nat.ptr(a), nat.ptr(steer_unit), self.noise_seed,
                                                  self.step_count, self.env_id_offset,
                                                  1 if self.zero_fill_targets else 0, self._stream()),
                  "tk_gogoro_pre_physics")
        self.sim.set_dof_position_target_tensor(self.dof_pos_command)
        self.sim.set_dof_velocity_target_tensor(self.dof_speed_command)

    def compute_observations(self) -> torch.Tensor:
        """compute_gogoro_observations (gogoro_new.py:692-723) on the current state -> clean obs [N,6]."""
        out = torch.empty((self.n_envs, NUM_OBS), device=self.device, dtype=torch.float)
        nat.check(self._lib.tk_gogoro_compute_observations(self._handle, nat.ptr(out), self._stream()),
                  "tk_gogoro_compute_observations")
        return out

    def compute_reward(self, cur_obs: Optional[torch.Tensor] = None):
        """compute_gogoro_reward (gogoro_new.py:645-684) on ``cur_obs`` (default: ``buffer_obs[:, -1]``)."""
        o = (self.buffer_obs[:, -1] if cur_obs is None else cur_obs).contiguous()
        nat.check(self._lib.tk_gogoro_compute_reward(self._handle, nat.ptr(o), self._stream()), "tk_gogoro_compute_reward")

    def _push_resets(self, ids_int32):
        """gogoro_new.py:547,552: the reset rows reach a simulator that keeps its own state only through these two calls
        (the reference asserts both)."""
        if ids_int32 is None:
            return
        assert self.sim.set_actor_root_state_tensor_indexed(self.root_tensor, ids_int32)
        assert self.sim.set_dof_state_tensor_indexed(self.state_dof, ids_int32)

    def compute_obs_rwd(self, sensor_unit: Optional[torch.Tensor] = None):
        """gogoro_new.py:424-462.  ``sensor_unit`` [N,5]: the five N(0,1) draws of :451-460 in draw order."""
        self.sim.refresh_actor_root_state_tensor()
        self.sim.refresh_dof_state_tensor()
        su = None if sensor_unit is None else sensor_unit.contiguous()
        nat.check(self._lib.tk_gogoro_compute_obs_rwd(self._handle, nat.ptr(su), self.noise_seed, self.step_count,
                                                      self.env_id_offset, self._stream()), "tk_gogoro_compute_obs_rwd")
        self.curent_step += 1

    def reset_idx(self, env_ids: torch.Tensor, unit: Optional[torch.Tensor] = None):
        """gogoro_new.py:505-590.  ``unit`` [K,10]: the draws in reference order (see include/thormang_b200.h);
        drawn here with torch.rand / torch.randn when omitted."""
        k = len(env_ids)
        if k == 0:
            return
        if unit is None:
            unit = torch.randn((k, RESET_NOISE_COLS), device=self.device)
            uni = torch.rand((k, RESET_NOISE_COLS), device=self.device)
            for col in (0, 2, 3, 4):
                unit[:, col] = uni[:, col]
        ids = env_ids.to(device=self.device, dtype=torch.long).contiguous()
        nat.check(self._lib.tk_gogoro_reset_idx(self._handle, nat.ptr(ids), k, nat.ptr(unit.contiguous()), self._stream()),
                  "tk_gogoro_reset_idx")
        if self._sim_detached():
            self._push_resets(ids.to(torch.int32))                                # gogoro_new.py:506

    def post_physics_step(self, sensor_unit: Optional[torch.Tensor] = None, reset_unit: Optional[torch.Tensor] = None,
                          speed_unit: Optional[torch.Tensor] = None, yaw_unit: Optional[torch.Tensor] = None,
                          noise_mode: Optional[int] = None):
        """gogoro_new.py:373-389 + vec_task.py:345 as one fused kernel (Philox noise unless the draws are supplied)."""
        if self._sim_detached():                # a simulator with its own state (sim/detached_sim.py, PhysX)
            self.sim.refresh_actor_root_state_tensor()                            # gogoro_new.py:425-426
            self.sim.refresh_dof_state_tensor()
            self._collect_reset_ids()                                             # :377 -- the fused kernel consumes reset_buf
            self._post_physics_fused(sensor_unit, reset_unit, speed_unit, yaw_unit, noise_mode)
            self._push_resets(self._collected_reset_ids())
            return
        self._post_physics_fused(sensor_unit, reset_unit, speed_unit, yaw_unit, noise_mode)

    def _post_physics_fused(self, sensor_unit, reset_unit, speed_unit, yaw_unit, noise_mode):
        supplied = sensor_unit is not None or reset_unit is not None
        if not supplied and noise_mode is None and speed_unit is None and yaw_unit is None:
            # production path: in-kernel Philox, nothing allocated per step
            nz = self._philox_noise
            nz.reset.seed, nz.reset.step, nz.reset.env_id_offset = self.noise_seed, self.step_count, self.env_id_offset
            rc = self._post_physics(self._handle, self._philox_noise_ref,
                                    nat.current_stream_ptr(self._torch_device))
            if rc:
                nat.check(rc, "tk_gogoro_post_physics")
            self.step_count += 1
            self.curent_step += 1
            return
        nz = nat.GogoroNoise()
        if noise_mode is None:
            noise_mode = nat.NOISE_COMPACT if supplied else nat.NOISE_PHILOX
        nz.reset.mode = noise_mode
        keep = []   # keeps temporaries alive until the launch is enqueued
        for field, t in (("sensor_unit", sensor_unit), ("speed_unit", speed_unit), ("yaw_unit", yaw_unit)):
            if t is not None:
                t = t.contiguous()
                keep.append(t)
            setattr(nz, field, nat.ptr(t))
        if noise_mode in (nat.NOISE_COMPACT, nat.NOISE_DENSE):
            if reset_unit is None:   # no rows: a set reset bit falls back to the mean draw (u = 0.5 / z = 0)
                reset_unit = torch.zeros((1, RESET_NOISE_COLS), device=self.device)
                nz.reset.rows = 0
            else:
                reset_unit = reset_unit.contiguous()
                nz.reset.rows = reset_unit.shape[0]
            keep.append(reset_unit)
            nz.reset.unit_a = nat.ptr(reset_unit)
        nz.reset.seed, nz.reset.step, nz.reset.env_id_offset = self.noise_seed, self.step_count, self.env_id_offset
        nat.check(self._lib.tk_gogoro_post_physics(self._handle, C.byref(nz), self._stream()), "tk_gogoro_post_physics")
        self.step_count += 1
        self.curent_step += 1

    def philox_noise(self, step: Optional[int] = None):
        """Materialise the in-kernel draws of ``step``: dict(steer_unit, sensor_unit, reset_unit, speed_unit, yaw_unit)."""
        n, dev = self.n_envs, self.device
        out = dict(steer_unit=torch.empty(n, device=dev), sensor_unit=torch.empty((n, 5), device=dev),
                   reset_unit=torch.empty((n, RESET_NOISE_COLS), device=dev), speed_unit=torch.empty(n, device=dev),
                   yaw_unit=torch.empty(n, device=dev))
        nat.check(self._lib.tk_gogoro_philox_noise(
            self._handle, self.noise_seed, self.step_count if step is None else step, self.env_id_offset,
            nat.ptr(out["steer_unit"]), nat.ptr(out["sensor_unit"]), nat.ptr(out["reset_unit"]),
            nat.ptr(out["speed_unit"]), nat.ptr(out["yaw_unit"]), self._stream()), "tk_gogoro_philox_noise")
        return out

    def advance_noise_epoch(self):
        """Capture ONE call of this inside a CUDA graph of steps: ``step_count`` is frozen in the captured launches, the
        device-side epoch bumped here makes every replay draw fresh in-kernel noise."""
        nat.check(self._lib.tk_gogoro_advance_epoch(self._handle, self._stream()), "tk_gogoro_advance_epoch")

    def episode_statistics(self, accumulate: bool = False) -> torch.Tensor:
        """fp64[8]: sum reward, #reset, #timeout, sum return/length of finished episodes, sum |tilt|, sum |yaw err|, N."""
        nat.check(self._lib.tk_gogoro_stats(self._handle, nat.ptr(self.episode_stats), 1 if accumulate else 0,
                                            self._stream()), "tk_gogoro_stats")
        return self.episode_stats
