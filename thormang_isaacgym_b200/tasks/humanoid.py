"""``Humanoid`` -- drop-in for ``isaacgymenvs/tasks/humanoid.py::Humanoid`` on the env-step path.

Same constructor signature, cfg keys, attribute names and method surface as the reference
class (humanoid.py:41-296); every method body is ONE call into the sm_100a CUDA extension
through the C ABI (include/thormang_b200.h).  No TorchScript, no CPU fallback.

What differs from the reference, on purpose:

* the simulator is injected (``cfg["sim_backend"]``) or the fake-sim shim is used: isaacgym /
  PhysX cannot exist here, and the viewer / debug-line code (humanoid.py:72-75,298-316) is
  left out with the graphics stack;
* ``post_physics_step`` is a single fused kernel: progress++, reset of flagged envs, observations,
  reward, reset mask and timeout mask -- no ``nonzero()``, no host sync.  Reset noise is drawn
  in-kernel (Philox) by default; ``post_physics_step(unit_pos=..., unit_vel=...)`` feeds the
  reference's ``[K, D]`` draws (row k <-> k-th reset env) for bit-level parity;
* ``reset_idx(env_ids)`` draws its noise with the same two ``torch.rand`` calls as
  ``torch_rand_float`` (humanoid.py:258-259), so a seeded torch generator gives the same reset.
"""
import ctypes as C
from typing import Optional

import numpy as np
import torch

from .. import _native as nat
from ..sim.fake_sim import FakeSim
from ..synthetic import humanoid_asset
from .base.vec_task import VecTask

NUM_OBS = 110
NUM_DOF = 21


class Humanoid(VecTask):
    fuses_action_clip = True
    # what the sibling task built from the same kernel template overrides (tasks/ant.py)
    _NUM_OBS, _NUM_DOF, _NUM_SENSOR = NUM_OBS, NUM_DOF, 12
    _VARIANT = nat.VARIANT_HUMANOID
    _HAS_DOF_FORCE = True
    _EFFORT_KEY = "motor_efforts"                                                 # humanoid.py:160-171
    _asset = staticmethod(humanoid_asset)

    def __init__(self, cfg, rl_device, sim_device, graphics_device_id, headless,
                 virtual_screen_capture=False, force_render=False):
        self.cfg = cfg
        env = cfg["env"]
        self.randomization_params = cfg.get("task", {}).get("randomization_params", {})
        self.randomize = cfg.get("task", {}).get("randomize", False)
        if self.randomize:
            raise NotImplementedError("PhysX domain randomisation (vec_task.py:538-768) is outside the hot path")
        # humanoid.py:48-65
        self.dof_vel_scale = env["dofVelocityScale"]
        self.angular_velocity_scale = env.get("angularVelocityScale", 0.1)
        self.contact_force_scale = env["contactForceScale"]
        self.power_scale = env["powerScale"]
        self.heading_weight = env["headingWeight"]
        self.up_weight = env["upWeight"]
        self.actions_cost_scale = env["actionsCost"]
        self.energy_cost_scale = env["energyCost"]
        self.joints_at_limit_cost_scale = env["jointsAtLimitCost"]
        self.death_cost = env["deathCost"]
        self.termination_height = env["terminationHeight"]
        self.debug_viz = False
        self.max_episode_length = env["episodeLength"]
        env["numObservations"] = self._NUM_OBS                                    # humanoid.py:67
        env["numActions"] = self._NUM_DOF                                         # humanoid.py:68

        super().__init__(config=cfg, rl_device=rl_device, sim_device=sim_device,
                         graphics_device_id=graphics_device_id, headless=headless,
                         virtual_screen_capture=virtual_screen_capture, force_render=force_render)

        sim, dev, n, d = self.sim, self.device, self.num_envs, self.num_dof
        # humanoid.py:78-98: aliases of simulator memory
        self.vec_sensor_tensor = sim.acquire_force_sensor_tensor().view(n, self._NUM_SENSOR)
        self.dof_force_tensor = sim.acquire_dof_force_tensor().view(n, d) if self._HAS_DOF_FORCE else None
        self.root_states = sim.acquire_actor_root_state_tensor()
        self.initial_root_states = self.root_states.clone()
        self.initial_root_states[:, 7:13] = 0
        self.dof_state = sim.acquire_dof_state_tensor()
        self.dof_pos = self.dof_state.view(n, d, 2)[..., 0]
        self.dof_vel = self.dof_state.view(n, d, 2)[..., 1]
        # humanoid.py:99-103
        zero = torch.tensor([0.0], device=dev)
        self.initial_dof_pos = torch.zeros((n, d), device=dev, dtype=torch.float)
        self.initial_dof_pos = torch.where(self.dof_limits_lower > zero, self.dof_limits_lower,
                                           torch.where(self.dof_limits_upper < zero, self.dof_limits_upper,
                                                       self.initial_dof_pos)).contiguous()
        self.initial_dof_vel = torch.zeros((n, d), device=dev, dtype=torch.float)
        # humanoid.py:106-117
        f = dict(device=dev, dtype=torch.float)
        self.up_vec = torch.tensor([0.0, 0.0, 1.0], **f).repeat((n, 1))
        self.heading_vec = torch.tensor([1.0, 0.0, 0.0], **f).repeat((n, 1))
        self.inv_start_rot = torch.cat((-self.start_rotation[:3], self.start_rotation[3:])).repeat((n, 1))
        self.basis_vec0 = self.heading_vec.clone()
        self.basis_vec1 = self.up_vec.clone()
        self.targets = torch.tensor([1000.0, 0.0, 0.0], **f).repeat((n, 1))
        self.target_dirs = torch.tensor([1.0, 0.0, 0.0], **f).repeat((n, 1))
        self.dt = cfg["sim"]["dt"]
        self.potentials = torch.tensor([-1000. / self.dt], **f).repeat(n)
        self.prev_potentials = self.potentials.clone()
        # a slab-backed simulator (host-driven pipelines) provides the storage of self.actions next to its state
        self.actions = sim.actions_view if getattr(sim, "actions_view", None) is not None else torch.zeros((n, d), **f)
        self.forces = torch.zeros((n, d), **f)
        self.episode_return = torch.zeros(n, **f) if env.get("episodeStatistics", True) else None
        # humanoid.py:289 bumps randomize_buf every step, but only domain randomisation (out of scope, randomize is
        # False in Humanoid.yaml:64) ever reads it: the kernel skips that 16 B/env read-modify-write unless asked
        self.track_randomize_buf = bool(env.get("trackRandomizeBuf", False))
        # up_vec / heading_vec (humanoid.py:246) are read by the debug drawing only (:298-316); True keeps the reference's
        # behaviour of refreshing them every step (24 B per env-step), False leaves them at their initial values
        self.write_debug_vectors = bool(env.get("writeDebugVectors", True))
        self.episode_stats = torch.zeros(nat.TK_STATS, device=dev, dtype=torch.float64)
        # in-kernel reset noise
        self.noise_seed = int(cfg.get("seed", 0)) & 0xFFFFFFFFFFFFFFFF
        self.env_id_offset = int(env.get("envIdOffset", 0))   # global id of local env 0 (multi-GPU range sharding)
        self.step_count = 0        # keys the in-kernel Philox streams together with the seed and the global env id
        self._handle = None
        self._create_native()

    # ------------------------------------------------------------------------------------------
    def create_sim(self):
        """humanoid.py:119-217 without PhysX: take the injected simulator or build the fake-sim shim."""
        self.up_axis_idx = 2
        n = self.cfg["env"]["numEnvs"]
        sim = self.cfg.get("sim_backend")
        if sim is None:
            sim = FakeSim(n, self._NUM_DOF, self.device, num_sensor_floats=self._NUM_SENSOR, asset=self._asset())
        self.sim = sim
        self.num_dof = sim.num_dof
        if self.num_dof != self._NUM_DOF:
            raise ValueError(f"{type(self).__name__} kernels are built for {self._NUM_DOF} DOFs, simulator has {self.num_dof}")
        dev = self.device
        lo = sim.asset["dof_limits_lower"].to(dev).float()
        hi = sim.asset["dof_limits_upper"].to(dev).float()
        self.dof_limits_lower = torch.minimum(lo, hi)                             # humanoid.py:205-215 (swap if lower > upper)
        self.dof_limits_upper = torch.maximum(lo, hi)
        self.motor_efforts = sim.asset[self._EFFORT_KEY].to(dev).float()          # humanoid.py:160-171 | ant.py:160-162
        self.max_motor_effort = float(self.motor_efforts.max())
        self.start_rotation = torch.tensor([0.0, 0.0, 0.0, 1.0], device=dev)      # humanoid.py:180-182

    def _create_native(self):
        lib = nat.lib()
        p = nat.HumanoidParams()
        p.device = self.device_id
        p.num_envs, p.num_dof, p.variant = self.num_envs, self.num_dof, self._VARIANT
        p.dt = self.dt
        p.dof_vel_scale, p.angular_velocity_scale = self.dof_vel_scale, self.angular_velocity_scale
        p.contact_force_scale, p.power_scale = self.contact_force_scale, self.power_scale
        p.heading_weight, p.up_weight = self.heading_weight, self.up_weight
        p.actions_cost_scale, p.energy_cost_scale = self.actions_cost_scale, self.energy_cost_scale
        p.joints_at_limit_cost_scale = self.joints_at_limit_cost_scale
        p.death_cost, p.termination_height = self.death_cost, self.termination_height
        p.max_episode_length = float(self.max_episode_length)
        p.clip_actions = float(self.clip_actions) if not np.isinf(self.clip_actions) else 3.0e38
        p.max_motor_effort = self.max_motor_effort
        for name, src in (("targets", self.targets[0]), ("inv_start_rot", self.inv_start_rot[0]),
                          ("basis_vec0", self.basis_vec0[0]), ("basis_vec1", self.basis_vec1[0])):
            arr = getattr(p, name)
            for i, v in enumerate(src.tolist()):
                arr[i] = v
        for name, src in (("dof_limits_lower", self.dof_limits_lower), ("dof_limits_upper", self.dof_limits_upper),
                          ("motor_efforts", self.motor_efforts)):
            arr = getattr(p, name)
            for i, v in enumerate(src.tolist()):
                arr[i] = v
        handle = C.c_void_p()
        nat.check(lib.tk_humanoid_create(C.byref(p), C.byref(handle)), "tk_humanoid_create")
        self._handle = handle
        self._lib = lib
        self._post_physics = lib.tk_humanoid_post_physics
        self._torch_device = torch.device(self.device)
        self._noise = nat.ResetNoise()
        self._noise.seed, self._noise.env_id_offset = self.noise_seed, self.env_id_offset
        self._noise_ref = C.byref(self._noise)
        self._noise_is_philox = False
        self._bind()

    def _bind(self):
        b = nat.HumanoidBuffers()
        for name, t in (
                ("root_states", self.root_states), ("dof_state", self.dof_state), ("dof_force", self.dof_force_tensor),
                ("sensors", self.vec_sensor_tensor), ("actions", self.actions), ("obs_buf", self.obs_buf),
                ("rew_buf", self.rew_buf), ("reset_buf", self.reset_buf), ("progress_buf", self.progress_buf),
                ("timeout_buf", self.timeout_buf),
                ("randomize_buf", self.randomize_buf if self.track_randomize_buf else None),
                ("potentials", self.potentials), ("prev_potentials", self.prev_potentials),
                ("up_vec", self.up_vec if self.write_debug_vectors else None),
                ("heading_vec", self.heading_vec if self.write_debug_vectors else None),
                ("initial_root_states", self.initial_root_states),
                ("initial_dof_pos", self.initial_dof_pos), ("episode_return", self.episode_return)):
            if t is not None:
                if not t.is_contiguous() or t.device != torch.device(self.device):
                    raise ValueError(f"{name}: the kernels need a contiguous tensor on {self.device}")
            setattr(b, name, nat.ptr(t))
        nat.check(self._lib.tk_humanoid_bind(self._handle, C.byref(b)), "tk_humanoid_bind")

    def _obs_copy_hook(self, dst):
        """N4: the fused kernel stores its obs tile to ``dst`` too (a rollout slot); None switches the copy off."""
        nat.check(self._lib.tk_humanoid_set_obs_copy(self._handle, nat.ptr(dst)), "tk_humanoid_set_obs_copy")
        return True

    def rebind(self):
        """Call after replacing any buffer attribute (e.g. ``env.reset_buf = ...``) by a new tensor."""
        self._bind()

    def _stream(self):
        return nat.current_stream_ptr(self.device)

    def __del__(self):
        try:
            if getattr(self, "_handle", None):
                self._lib.tk_humanoid_destroy(self._handle)
                self._handle = None
        except Exception:
            pass

    # ------------------------------------------------------------------------------------------
    def pre_physics_step(self, actions: torch.Tensor):
        """humanoid.py:281-285 (+ the clamp of vec_task.py:327 when called from ``step``)."""
        a = actions.to(self.device, dtype=torch.float)
        if not a.is_contiguous():
            a = a.contiguous()
        nat.check(self._lib.tk_humanoid_pre_physics(self._handle, nat.ptr(a), nat.ptr(self.forces),
                                                    1 if self._clip_in_kernel else 0, self._stream()),
                  "tk_humanoid_pre_physics")
        self.sim.set_dof_actuation_force_tensor(self.forces)

    def _refresh_sim_tensors(self):
        """humanoid.py:240-243."""
        self.sim.refresh_dof_state_tensor()
        self.sim.refresh_actor_root_state_tensor()
        self.sim.refresh_force_sensor_tensor()
        if self._HAS_DOF_FORCE:
            self.sim.refresh_dof_force_tensor()

    def _push_resets(self, ids_int32):
        """humanoid.py:264-271: the reset rows reach a simulator that keeps its own state only through these two calls."""
        if ids_int32 is None:
            return
        self.sim.set_actor_root_state_tensor_indexed(self.initial_root_states, ids_int32)
        self.sim.set_dof_state_tensor_indexed(self.dof_state, ids_int32)

    def compute_observations(self):
        """humanoid.py:239-251."""
        self._refresh_sim_tensors()
        nat.check(self._lib.tk_humanoid_compute_observations(self._handle, self._stream()),
                  "tk_humanoid_compute_observations")

    def compute_reward(self, actions=None):
        """humanoid.py:219-237."""
        nat.check(self._lib.tk_humanoid_compute_reward(self._handle, self._stream()), "tk_humanoid_compute_reward")

    def reset_idx(self, env_ids: torch.Tensor, unit_pos: Optional[torch.Tensor] = None,
                  unit_vel: Optional[torch.Tensor] = None):
        """humanoid.py:253-279.  Noise = the two torch.rand draws of torch_rand_float unless supplied."""
        k = len(env_ids)
        if k == 0:
            return
        if unit_pos is None:
            unit_pos = torch.rand(k, self.num_dof, device=self.device)            # humanoid.py:258
        if unit_vel is None:
            unit_vel = torch.rand(k, self.num_dof, device=self.device)            # humanoid.py:259
        ids = env_ids.to(device=self.device, dtype=torch.long).contiguous()
        nat.check(self._lib.tk_humanoid_reset_idx(self._handle, nat.ptr(ids), k, nat.ptr(unit_pos.contiguous()),
                                                  nat.ptr(unit_vel.contiguous()), self._stream()),
                  "tk_humanoid_reset_idx")
        if self._sim_detached():
            self._push_resets(ids.to(torch.int32))                                # humanoid.py:264

    def post_physics_step(self, unit_pos: Optional[torch.Tensor] = None, unit_vel: Optional[torch.Tensor] = None,
                          noise_mode: Optional[int] = None):
        """humanoid.py:287-296 + vec_task.py:345 as one fused kernel."""
        detached = self._sim_detached()
        if detached:                            # a simulator with its own state: fetch it, and note who is about to be reset
            self._refresh_sim_tensors()
            self._collect_reset_ids()           # humanoid.py:290 -- the fused kernel consumes reset_buf
        noise = self._noise                     # one struct per task object: nothing is allocated per step
        if unit_pos is None and noise_mode is None:
            if not self._noise_is_philox:       # the default path touches only what changes from step to step
                noise.mode, noise.unit_a, noise.unit_b, noise.rows = nat.NOISE_PHILOX, None, None, 0
                self._noise_is_philox = True
        else:
            self._noise_is_philox = False
            if unit_pos is not None:
                unit_pos, unit_vel = unit_pos.contiguous(), unit_vel.contiguous()
                noise.mode = nat.NOISE_COMPACT if noise_mode is None else noise_mode
                noise.unit_a, noise.unit_b, noise.rows = nat.ptr(unit_pos), nat.ptr(unit_vel), unit_pos.shape[0]
            else:
                noise.mode = noise_mode
                noise.unit_a, noise.unit_b, noise.rows = None, None, 0
        noise.seed, noise.step, noise.env_id_offset = self.noise_seed, self.step_count, self.env_id_offset
        rc = self._post_physics(self._handle, self._noise_ref, nat.current_stream_ptr(self._torch_device))
        if rc:
            nat.check(rc, "tk_humanoid_post_physics")
        self.step_count += 1
        if detached:
            self._push_resets(self._collected_reset_ids())

    # ------------------------------------------------------------------------------------------
    def reward_terms(self) -> torch.Tensor:
        """[N,7]: heading, up, actions_cost, dof_at_limit_cost, electricity_cost, alive, progress (humanoid.py:344-363)."""
        out = torch.empty((self.num_envs, 7), device=self.device, dtype=torch.float)
        nat.check(self._lib.tk_humanoid_reward_terms(self._handle, nat.ptr(out), self._stream()), "tk_humanoid_reward_terms")
        return out

    def wm_reward(self, obs: torch.Tensor) -> torch.Tensor:
        """Second ``compute_reward`` of the file (humanoid.py:425-472) over obs[..., 110]."""
        flat = obs.to(self.device, dtype=torch.float).reshape(-1, NUM_OBS).contiguous()
        out = torch.empty(flat.shape[0], device=self.device, dtype=torch.float)
        nat.check(self._lib.tk_humanoid_wm_reward(self._handle, nat.ptr(flat), flat.shape[0], nat.ptr(out), self._stream()),
                  "tk_humanoid_wm_reward")
        return out.view(obs.shape[:-1])

    def philox_noise(self, step: Optional[int] = None):
        """The U(0,1) draws the fused kernel uses at ``step`` (default: the step about to run)."""
        ua = torch.empty((self.num_envs, self.num_dof), device=self.device)
        ub = torch.empty_like(ua)
        nat.check(self._lib.tk_humanoid_philox_noise(self._handle, self.noise_seed,
                                                     self.step_count if step is None else step, self.env_id_offset,
                                                     nat.ptr(ua), nat.ptr(ub), self._stream()), "tk_humanoid_philox_noise")
        return ua, ub

    def advance_noise_epoch(self):
        """Capture ONE call of this inside a CUDA graph of steps: ``step_count`` is frozen in the captured launches, the
        device-side epoch bumped here makes every replay draw fresh in-kernel noise."""
        nat.check(self._lib.tk_humanoid_advance_epoch(self._handle, self._stream()), "tk_humanoid_advance_epoch")

    def episode_statistics(self, accumulate: bool = False) -> torch.Tensor:
        """fp64[8] on device: sum reward, #reset, #timeout, sum return/length of finished episodes, sum up_proj,
        sum heading_proj, N -- the all-reduce payload of multi-GPU runs (parallel.py)."""
        nat.check(self._lib.tk_humanoid_stats(self._handle, nat.ptr(self.episode_stats), 1 if accumulate else 0,
                                              self._stream()), "tk_humanoid_stats")
        return self.episode_stats
