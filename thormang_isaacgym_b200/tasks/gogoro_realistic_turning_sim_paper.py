"""``Gogoro`` of ``isaacgymenvs/tasks/gogoro_realistic_turning_sim_paper.py`` -- the 8-channel x 20-frame history
scooter task used for the paper runs (cfg/task/Gogoro_paper.yaml) -- on the env-step path.
Registered here as ``"GogoroPaper"`` (the reference file is not in ``isaacgym_task_map``).

Same constructor signature, cfg keys, attribute names and method surface as the reference class (:36-700); method
bodies are calls into the sm_100a CUDA extension (include/thormang_b200.h, csrc/tk_gogoro_paper.cu).  No
TorchScript, no CPU fallback.

Storage (see the header): ``obs_buf`` [N,160] IS ``buffer_obs_noisy`` (time-ordered, shifted in place; the
reference's ``obs_buf`` is a flatten() view of it too, :545); the clean frames are a ring and ``buffer_obs`` is a
property that materialises the reference's time-ordered [N,20,8] tensor on demand (the step path never needs it).

The reward's ``reward7`` (:747) is a mean over every env: a step is the fused kernel plus a small finalize kernel,
and with ``cfg["env"]["worldReduce"]`` (a ``torch.distributed`` process group or True for the default group) the two
sums are all-reduced in between, so that range-sharded multi-GPU runs compute the reference's GLOBAL mean.

Left out with PhysX / the graphics stack: the PUSH_ROBOT force block (:441-486), viewer lines (:415-438),
``set_actor_dof_properties`` (:657-675), terrain.  Module flags of the reference file as saved (DEBUG = True):
fixed delay tap ``-3`` (``cfg["env"]["useSteerDelay"]`` restores :355), start speed 1.3 (``startSpeed``).
"""
import ctypes as C
from typing import Optional

import numpy as np
import torch

from .. import _native as nat
from ..sim.fake_sim import FakeSim
from .base.vec_task import VecTask

NUM_CH = 8            # :110
BUFF = 20             # :112
NUM_ACT = 1
RESET_NOISE_COLS = 5
SENSOR_NOISE_COLS = 6


class Gogoro(VecTask):
    # what tasks/gogoro_realistic_turning.py (the sibling reference file, same class) changes
    VARIANT = nat.GOGORO_PAPER_VARIANT_PAPER
    RESET_NOISE_COLS = RESET_NOISE_COLS
    USE_STEER_DELAY_DEFAULT = False            # USE_STEER_DELAY = not DEBUG, :31, with DEBUG = True as saved (:22)

    def __init__(self, cfg, rl_device, sim_device, graphics_device_id, headless,
                 virtual_screen_capture=False, force_render=False):
        self.steering_sensitivity = 0.1                                           # :40
        self.curent_step = 0
        self.cfg = cfg
        nz = cfg["noises"]
        self.n_envs = cfg["env"]["numEnvs"]
        self.max_episode_length = int(cfg["env"]["max_steps"])                    # :47 (0-d tensor there)
        # :55-80
        self.imu_filter_noise = nz["imu_filter_noise"]
        self.imu_noise = nz["imu_noise"]
        self.speed_sensor_noise = nz["speed_sensor_noise"]
        self.speed_sensor_offset = nz["speed_sensor_offset"]
        self.seat_offset_x_range = nz.get("seat_offset_x_range", [-0.1, 0.1])
        self.seat_offset_y_range = nz.get("seat_offset_y_range", [-0.1, 0.1])
        self.seat_offset_z_range = nz.get("seat_offset_z_range", [-0.05, 0.05])
        self.spawn_x_angle = nz.get("spawn_x_angle", [-0.02, 0.02])
        self.imu_x_offset = nz["imu_x_offset"]
        self.steering_damping_range = nz.get("steering_damping_range", [50, 1000])
        self.steering_action_noise = nz.get("steering_action_noise", [-0.05, 0.05])
        self.speed_range = nz["speed_range"]
        self.speed_freq_update = nz["speed_freq_update"]
        self.yaw_freq_update = nz["yaw_freq_update"]
        self.steering_offset = nz["steering_offset"]
        self.command_delay = nz["command_delay"]
        self.min_speed, self.max_speed, self.max_steering = 0.0, 10.0, 0.5        # :84-86
        self.use_steer_delay = bool(cfg["env"].get("useSteerDelay", self.USE_STEER_DELAY_DEFAULT))       # USE_STEER_DELAY = not DEBUG, :31
        self.start_speed = float(cfg["env"].get("startSpeed", 1.3))               # DEBUG_START_SPEED, :26,638
        self.buff_size = BUFF
        cfg["env"]["numObservations"] = NUM_CH * self.buff_size                   # :117
        cfg["env"]["numActions"] = NUM_ACT                                        # :118

        super().__init__(config=cfg, rl_device=rl_device, sim_device=sim_device,
                         graphics_device_id=graphics_device_id, headless=headless,
                         virtual_screen_capture=virtual_screen_capture, force_render=force_render)

        sim, dev, n, d = self.sim, self.device, self.n_envs, self.num_dof
        f = dict(device=dev, dtype=torch.float)
        self.dt = cfg["sim"]["dt"]
        # :81-102
        self.yaw_command = torch.zeros(n, **f)
        self.curent_speed = self.get_randoms(n, self.speed_range)
        self.last_err = torch.zeros(n, **f)
        self.curent_command = torch.zeros(n, **f)
        self.steer_offsets = self.get_randoms(n, self.steering_offset)
        self.steer_delay = self.get_randoms(n, self.command_delay).to(torch.long)
        self.command_history = torch.zeros((n, int(self.command_delay[1])), **f)
        self.curent_speed_offset = self.get_randoms(n, self.speed_sensor_offset)
        self.curent_imu_x_offset = self.get_randoms(n, self.imu_x_offset)
        self.speed_no_noise = torch.zeros(n, **f)
        self.integral_error = torch.zeros(n, **f)
        self.last_err_speed = torch.zeros(n, **f)
        # clean frames: ring [N,20,8] + compact command ring [N,20] (see the module docstring)
        self.buffer_obs_ring = torch.zeros((n, self.buff_size, NUM_CH), **f)
        self.command_ring = torch.zeros((n, self.buff_size), **f)
        # :133-156
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
        self.diff_sums = torch.zeros(2, device=dev, dtype=torch.float64)
        self.noise_seed = int(cfg.get("seed", 0)) & 0xFFFFFFFFFFFFFFFF
        self.env_id_offset = int(cfg["env"].get("envIdOffset", 0))
        self.step_count = 0
        wr = cfg["env"].get("worldReduce", None)
        self._world_group = None if wr in (None, False) else wr     # True = default process group
        self._handle = None
        self._create_native()
        if cfg["env"].get("resetOnInit", True):
            self.reset_idx(torch.arange(0, n, device=dev, dtype=torch.long))       # :165

    # ------------------------------------------------------------------------------------------
    def create_sim(self):
        """:167-340 without PhysX: injected simulator or the fake-sim shim."""
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
        self.dof_name_to_id = {k: v for v, k in enumerate(self.dof_names)}        # :260
        self.thormang_pose = torch.zeros(n, self.num_dof, device=self.device)     # :275
        for j_name, value in self.cfg.get("joints_pos", {}).items():              # :296-302
            if j_name in self.dof_name_to_id:
                self.thormang_pose[:, self.dof_name_to_id[j_name]] = value + 0.0001 / 2
        for j_name in ("base_x", "base_y", "base_z"):                             # :305-317 (CENTER_ROBOT)
            if j_name in self.dof_name_to_id:
                self.thormang_pose[:, self.dof_name_to_id[j_name]] = 0.0 + 0.0001 / 2

    def _create_native(self):
        lib = nat.lib()
        p = nat.GogoroPaperParams()
        p.device = self.device_id
        p.num_envs, p.num_dof = self.n_envs, self.num_dof
        p.steering_joint = self.dof_name_to_id["steering_joint"]
        p.rear_wheel_joint = self.dof_name_to_id["rear_wheel_joint"]
        p.max_episode_length = self.max_episode_length
        p.speed_freq_update, p.yaw_freq_update = int(self.speed_freq_update), int(self.yaw_freq_update)
        p.buff_size, p.delay_width = self.buff_size, int(self.command_delay[1])
        p.use_steer_delay = 1 if self.use_steer_delay else 0
        p.variant = self.VARIANT
        p.max_steering, p.start_speed = self.max_steering, self.start_speed
        for name, pair in (("imu_filter_noise", self.imu_filter_noise), ("imu_noise", self.imu_noise),
                           ("speed_sensor_noise", self.speed_sensor_noise),
                           ("speed_sensor_offset", self.speed_sensor_offset), ("imu_x_offset", self.imu_x_offset),
                           ("steering_offset", self.steering_offset), ("command_delay", self.command_delay),
                           ("speed_range", self.speed_range)):
            arr = getattr(p, name)
            arr[0], arr[1] = float(pair[0]), float(pair[1])
        handle = C.c_void_p()
        nat.check(lib.tk_gogoro_paper_create(C.byref(p), C.byref(handle)), "tk_gogoro_paper_create")
        self._handle, self._lib = handle, lib
        self._torch_device = torch.device(self.device)
        self._post_physics, self._finalize_reward = lib.tk_gogoro_paper_post_physics, lib.tk_gogoro_paper_finalize_reward
        self._philox_noise = nat.GogoroPaperNoise()
        self._philox_noise.mode = nat.NOISE_PHILOX
        self._philox_noise_ref = C.byref(self._philox_noise)
        self._bind()

    def _bind(self):
        b = nat.GogoroPaperBuffers()
        for name, t in (
                ("root_tensor", self.root_tensor), ("state_dof", self.state_dof), ("obs_buf", self.obs_buf),
                ("buffer_obs_ring", self.buffer_obs_ring), ("command_ring", self.command_ring),
                ("rew_buf", self.rew_buf), ("reset_buf", self.reset_buf), ("progress_buf", self.progress_buf),
                ("timeout_buf", self.timeout_buf), ("yaw_command", self.yaw_command),
                ("curent_command", self.curent_command), ("curent_speed", self.curent_speed),
                ("steer_offsets", self.steer_offsets), ("curent_speed_offset", self.curent_speed_offset),
                ("curent_imu_x_offset", self.curent_imu_x_offset), ("speed_no_noise", self.speed_no_noise),
                ("steer_delay", self.steer_delay), ("command_history", self.command_history),
                ("thormang_pose", self.thormang_pose), ("root_reset_tensor", self.root_reset_tensor),
                ("dof_pos_command", self.dof_pos_command), ("dof_speed_command", self.dof_speed_command),
                ("episode_return", self.episode_return)):
            if t is not None and (not t.is_contiguous() or t.device != torch.device(self.device)):
                raise ValueError(f"{name}: the kernels need a contiguous tensor on {self.device}")
            setattr(b, name, nat.ptr(t))
        nat.check(self._lib.tk_gogoro_paper_bind(self._handle, C.byref(b)), "tk_gogoro_paper_bind")

    def rebind(self):
        self._bind()

    def _stream(self):
        return nat.current_stream_ptr(self.device)

    def __del__(self):
        try:
            if getattr(self, "_handle", None):
                self._lib.tk_gogoro_paper_destroy(self._handle)
                self._handle = None
        except Exception:
            pass

    # :549-553
    def get_randoms(self, shape, bounds):
        return bounds[0] + torch.rand(shape, device=self.device) * (bounds[1] - bounds[0])

    def get_randoms_norm(self, shape, mean_cov):
        return mean_cov[0] + torch.randn(shape, device=self.device) * mean_cov[1]

    # the reference's frame buffers -----------------------------------------------------------
    @property
    def buffer_obs_noisy(self) -> torch.Tensor:
        """[N,20,8] view of ``obs_buf`` (the reference's relation turned around, :545)."""
        return self.obs_buf.view(self.n_envs, self.buff_size, NUM_CH)

    @property
    def buffer_obs(self) -> torch.Tensor:
        """Clean frames in time order, materialised from the ring (a copy: off the step path)."""
        out = torch.empty((self.n_envs, self.buff_size, NUM_CH), device=self.device, dtype=torch.float)
        nat.check(self._lib.tk_gogoro_paper_clean_frames(self._handle, nat.ptr(out), self._stream()),
                  "tk_gogoro_paper_clean_frames")
        return out

    # ------------------------------------------------------------------------------------------
    def pre_physics_step(self, actions: torch.Tensor):
        """:342-378."""
        a = actions.to(self.device, dtype=torch.float).contiguous()
        nat.check(self._lib.tk_gogoro_paper_pre_physics(self._handle, nat.ptr(a), 1 if self.zero_fill_targets else 0,
                                                        self._stream()), "tk_gogoro_paper_pre_physics")
        self.sim.set_dof_position_target_tensor(self.dof_pos_command)
        self.sim.set_dof_velocity_target_tensor(self.dof_speed_command)

    def compute_observations(self) -> torch.Tensor:
        """compute_gogoro_observations (:770-807) on the current state -> clean obs [N,8]."""
        out = torch.empty((self.n_envs, NUM_CH), device=self.device, dtype=torch.float)
        nat.check(self._lib.tk_gogoro_paper_compute_observations(self._handle, nat.ptr(out), self._stream()),
                  "tk_gogoro_paper_compute_observations")
        return out

    def _noise(self, sensor_unit, reset_unit, speed_unit, yaw_unit, keep):
        if sensor_unit is None and reset_unit is None and speed_unit is None and yaw_unit is None:
            nz = self._philox_noise
            nz.seed, nz.step, nz.env_id_offset = self.noise_seed, self.step_count, self.env_id_offset
            return self._philox_noise_ref
        nz = nat.GogoroPaperNoise()
        nz.mode = nat.NOISE_DENSE
        for field, t in (("sensor_unit", sensor_unit), ("reset_unit", reset_unit), ("speed_unit", speed_unit),
                         ("yaw_unit", yaw_unit)):
            if t is not None:
                t = t.to(self.device, dtype=torch.float).contiguous()
                if t.shape[0] != self.n_envs:
                    raise ValueError(f"{field}: supplied draws are addressed by env id, need {self.n_envs} rows")
                want = {"sensor_unit": (SENSOR_NOISE_COLS,), "reset_unit": (self.RESET_NOISE_COLS,)}.get(field, ())
                if tuple(t.shape[1:]) != want:      # the kernels index these with a fixed row stride
                    raise ValueError(f"{field}: need shape {(self.n_envs,) + want}, got {tuple(t.shape)}")
                keep.append(t)
            setattr(nz, field, nat.ptr(t))
        keep.append(nz)
        return C.byref(nz)

    def _finalize(self, stream=None):
        """Second kernel of a step: the cross-env reward term (:747), with the all-reduce when range-sharded."""
        if stream is None:
            stream = nat.current_stream_ptr(self._torch_device)
        sums = 0
        if self._world_group is not None:
            import torch.distributed as dist
            nat.check(self._lib.tk_gogoro_paper_diff_sums(self._handle, nat.ptr(self.diff_sums), stream),
                      "tk_gogoro_paper_diff_sums")
            dist.all_reduce(self.diff_sums, op=dist.ReduceOp.SUM,
                            group=None if self._world_group is True else self._world_group)
            sums = self.diff_sums.data_ptr()
        rc = self._finalize_reward(self._handle, sums, stream)
        if rc:
            nat.check(rc, "tk_gogoro_paper_finalize_reward")

    def compute_obs_rwd(self, sensor_unit: Optional[torch.Tensor] = None):
        """:490-545.  ``sensor_unit`` [N,6]: the U(0,1) draws of :525-531 in draw order (in-kernel Philox if omitted)."""
        self.sim.refresh_actor_root_state_tensor()
        self.sim.refresh_dof_state_tensor()
        keep = []
        nz = self._noise(sensor_unit, None, None, None, keep)
        nat.check(self._lib.tk_gogoro_paper_compute_obs_rwd(self._handle, nz, self._stream()),
                  "tk_gogoro_paper_compute_obs_rwd")
        self._finalize()
        self.curent_step += 1

    def reset_idx(self, env_ids: torch.Tensor, unit: Optional[torch.Tensor] = None):
        """:599-688.  ``unit`` [K,RESET_NOISE_COLS] U(0,1) draws in reference order (drawn with torch.rand when omitted)."""
        k = len(env_ids)
        if k == 0:
            return
        if unit is None:
            unit = torch.rand((k, self.RESET_NOISE_COLS), device=self.device)
        ids = env_ids.to(device=self.device, dtype=torch.long).contiguous()
        unit = unit.to(self.device, dtype=torch.float).contiguous()
        if tuple(unit.shape) != (k, self.RESET_NOISE_COLS):      # the kernel reads RESET_NOISE_COLS floats per row
            raise ValueError(f"unit: need shape {(k, self.RESET_NOISE_COLS)}, got {tuple(unit.shape)}")
        nat.check(self._lib.tk_gogoro_paper_reset_idx(self._handle, nat.ptr(ids), k, nat.ptr(unit), self._stream()),
                  "tk_gogoro_paper_reset_idx")
        if self._sim_detached():
            self._push_resets(ids.to(torch.int32))                                # :604

    def _push_resets(self, ids_int32):
        """:645,650 (gogoro_realistic_turning.py:643,648): the reset rows reach a simulator that keeps its own state only
        through these two calls (the reference asserts both)."""
        if ids_int32 is None:
            return
        assert self.sim.set_actor_root_state_tensor_indexed(self.root_tensor, ids_int32)
        assert self.sim.set_dof_state_tensor_indexed(self.state_dof, ids_int32)

    def post_physics_step(self, sensor_unit: Optional[torch.Tensor] = None, reset_unit: Optional[torch.Tensor] = None,
                          speed_unit: Optional[torch.Tensor] = None, yaw_unit: Optional[torch.Tensor] = None):
        """:382-412 + vec_task.py:345: the fused kernel and the finalize kernel.  Supplied draws are addressed by
        env id ([N, ...] tensors); in-kernel Philox noise when none is given."""
        keep = []
        detached = self._sim_detached()
        if detached:                            # a simulator with its own state (sim/detached_sim.py, PhysX)
            self.sim.refresh_actor_root_state_tensor()                            # :491-492
            self.sim.refresh_dof_state_tensor()
            self._collect_reset_ids()                                             # :386 -- the fused kernel consumes reset_buf
        nz = self._noise(sensor_unit, reset_unit, speed_unit, yaw_unit, keep)
        stream = nat.current_stream_ptr(self._torch_device)
        rc = self._post_physics(self._handle, nz, stream)
        if rc:
            nat.check(rc, "tk_gogoro_paper_post_physics")
        self._finalize(stream)
        self.step_count += 1
        self.curent_step += 1
        if detached:
            self._push_resets(self._collected_reset_ids())

    def philox_noise(self, step: Optional[int] = None):
        """Materialise the in-kernel draws of ``step``: dict(sensor_unit, reset_unit, speed_unit, yaw_unit)."""
        n, dev = self.n_envs, self.device
        out = dict(sensor_unit=torch.empty((n, SENSOR_NOISE_COLS), device=dev),
                   reset_unit=torch.empty((n, self.RESET_NOISE_COLS), device=dev),
                   speed_unit=torch.empty(n, device=dev), yaw_unit=torch.empty(n, device=dev))
        nat.check(self._lib.tk_gogoro_paper_philox_noise(
            self._handle, self.noise_seed, self.step_count if step is None else step, self.env_id_offset,
            nat.ptr(out["sensor_unit"]), nat.ptr(out["reset_unit"]), nat.ptr(out["speed_unit"]),
            nat.ptr(out["yaw_unit"]), self._stream()), "tk_gogoro_paper_philox_noise")
        return out

    def advance_noise_epoch(self):
        nat.check(self._lib.tk_gogoro_paper_advance_epoch(self._handle, self._stream()), "tk_gogoro_paper_advance_epoch")

    def episode_statistics(self, accumulate: bool = False) -> torch.Tensor:
        """fp64[8]: sum reward, #reset, #timeout, sum return/length of finished episodes, sum |tilt|, sum |yaw err|, N."""
        nat.check(self._lib.tk_gogoro_paper_stats(self._handle, nat.ptr(self.episode_stats), 1 if accumulate else 0,
                                                  self._stream()), "tk_gogoro_paper_stats")
        return self.episode_stats
