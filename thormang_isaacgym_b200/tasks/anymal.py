"""``Anymal`` -- drop-in for ``isaacgymenvs/tasks/anymal.py::Anymal`` on the env-step path:
the canonical ``targets = action_scale * actions + default_dof_pos`` PD-target write (:229-232) and the contact-force
termination (:359-362).  This fork's copy of the task: 61 observations through ``compute_anymal_observations_V2``.

Same constructor signature, cfg keys (``cfg["env"]["learn"|"control"|"baseInitState"|"defaultJointAngles"|
"randomCommandVelocityRanges"]``), attribute names and method surface as the reference class (:40-321); method bodies
are single calls into the sm_100a CUDA extension (include/thormang_b200.h, csrc/tk_anymal.cu).  No TorchScript, no
CPU fallback.  Viewer camera (:99-104), PhysX asset / actor creation (:171-226) and domain randomisation are left out
with PhysX and the graphics stack; body indices come from the simulator object (``sim.base_index``,
``sim.knee_indices``) or ``cfg["env"]["baseIndex"|"kneeIndices"]``.
"""
import ctypes as C
from typing import Optional

import torch

from .. import _native as nat
from ..config import ANYMAL_DOF_NAMES
from ..sim.fake_sim import FakeSim
from .base.vec_task import VecTask

NUM_OBS = 61          # :83
NUM_DOF = 12          # :84
NUM_BODIES = 13


class Anymal(VecTask):
    fuses_action_clip = True

    def __init__(self, cfg, rl_device, sim_device, graphics_device_id, headless,
                 virtual_screen_capture=False, force_render=False):
        self.cfg = cfg
        learn, ctrl = cfg["env"]["learn"], cfg["env"]["control"]
        # :46-66
        self.lin_vel_scale = learn["linearVelocityScale"]
        self.ang_vel_scale = learn["angularVelocityScale"]
        self.dof_pos_scale = learn["dofPositionScale"]
        self.dof_vel_scale = learn["dofVelocityScale"]
        self.action_scale = ctrl["actionScale"]
        self.rew_scales = {"lin_vel_xy": learn["linearVelocityXYRewardScale"],
                           "ang_vel_z": learn["angularVelocityZRewardScale"], "torque": learn["torqueRewardScale"]}
        self.randomization_params = cfg.get("task", {}).get("randomization_params", {})
        self.randomize = cfg.get("task", {}).get("randomize", False)
        rng = cfg["env"]["randomCommandVelocityRanges"]
        self.command_x_range, self.command_y_range, self.command_yaw_range = rng["linear_x"], rng["linear_y"], rng["yaw"]
        b = cfg["env"]["baseInitState"]
        self.base_init_state = b["pos"] + b["rot"] + b["vLinear"] + b["vAngular"]                  # :72-78
        self.named_default_joint_angles = cfg["env"]["defaultJointAngles"]
        cfg["env"]["numObservations"] = NUM_OBS
        cfg["env"]["numActions"] = NUM_DOF

        super().__init__(config=cfg, rl_device=rl_device, sim_device=sim_device,
                         graphics_device_id=graphics_device_id, headless=headless,
                         virtual_screen_capture=virtual_screen_capture, force_render=force_render)

        sim, dev, n, d = self.sim, self.device, self.num_envs, self.num_dof
        self.dt = cfg["sim"]["dt"]
        self.max_episode_length_s = learn["episodeLength_s"]
        self.max_episode_length = int(self.max_episode_length_s / self.dt + 0.5)                    # :92
        self.Kp, self.Kd = ctrl["stiffness"], ctrl["damping"]
        self._rew_scales_cfg = dict(self.rew_scales)
        for key in self.rew_scales.keys():
            self.rew_scales[key] *= self.dt                                                         # :96-97
        # :107-125
        self.root_states = sim.acquire_actor_root_state_tensor()
        self.dof_state = sim.acquire_dof_state_tensor()
        self.dof_pos = self.dof_state.view(n, d, 2)[..., 0]
        self.dof_vel = self.dof_state.view(n, d, 2)[..., 1]
        self.contact_forces = sim.acquire_net_contact_force_tensor().view(n, -1, 3)
        self.torques = sim.acquire_dof_force_tensor().view(n, d)
        # :127-143
        self.commands = torch.zeros(n, 3, dtype=torch.float, device=dev)
        self.commands_x, self.commands_y, self.commands_yaw = self.commands[..., 0], self.commands[..., 1], self.commands[..., 2]
        self.default_dof_pos = torch.zeros((n, d), dtype=torch.float, device=dev)
        for i in range(cfg["env"]["numActions"]):
            self.default_dof_pos[:, i] = self.named_default_joint_angles[self.dof_names[i]]
        self.extras = {}
        self.initial_root_states = self.root_states.clone()
        self.initial_root_states[:] = torch.tensor(self.base_init_state, device=dev, dtype=torch.float)
        self.gravity_vec = torch.tensor([0.0, 0.0, -1.0], device=dev).repeat((n, 1))              # up_axis z, :142
        self.actions = torch.zeros(n, self.num_actions, dtype=torch.float, device=dev)
        self.targets = torch.zeros(n, d, dtype=torch.float, device=dev)
        self.episode_return = torch.zeros(n, device=dev) if cfg["env"].get("episodeStatistics", True) else None
        self.episode_stats = torch.zeros(nat.TK_STATS, device=dev, dtype=torch.float64)
        self.noise_seed = int(cfg.get("seed", 0)) & 0xFFFFFFFFFFFFFFFF
        self.env_id_offset = int(cfg["env"].get("envIdOffset", 0))
        self.step_count = 0
        self._handle = None
        self._create_native()
        if cfg["env"].get("resetOnInit", True):
            self.reset_idx(torch.arange(n, device=dev))                                             # :145

    def create_sim(self):
        """:147-226 without PhysX: injected simulator or the fake-sim shim."""
        n = self.cfg["env"]["numEnvs"]
        sim = self.cfg.get("sim_backend")
        if sim is None:
            sim = FakeSim(n, NUM_DOF, self.device, num_sensor_floats=0, dof_names=list(ANYMAL_DOF_NAMES), num_bodies=NUM_BODIES)
        self.sim = sim
        self.up_axis_idx = 2
        self.num_dof = sim.num_dof
        self.num_bodies = sim.num_bodies
        self.dof_names = list(sim.dof_names)
        env = self.cfg["env"]
        self.base_index = int(getattr(sim, "base_index", env.get("baseIndex", 0)))                  # :215,226
        knees = getattr(sim, "knee_indices", env.get("kneeIndices", (2, 5, 8, 11)))                 # :212-213,223-224
        self.knee_indices = torch.tensor(list(knees), dtype=torch.long, device=self.device)

    def _create_native(self):
        lib = nat.lib()
        p = nat.AnymalParams()
        p.device, p.num_envs, p.num_dof, p.num_bodies = self.device_id, self.num_envs, self.num_dof, self.num_bodies
        p.max_episode_length, p.base_index = self.max_episode_length, self.base_index
        knees = self.knee_indices.tolist()
        p.num_knee = len(knees)
        for i, k in enumerate(knees):
            p.knee_indices[i] = int(k)
        p.dof_pos_scale, p.dof_vel_scale, p.action_scale = self.dof_pos_scale, self.dof_vel_scale, self.action_scale
        p.clip_actions = float(self.clip_actions)
        p.dt = float(self.dt)
        p.rew_lin_vel_xy, p.rew_ang_vel_z = float(self._rew_scales_cfg["lin_vel_xy"]), float(self._rew_scales_cfg["ang_vel_z"])
        p.rew_torque = float(self._rew_scales_cfg["torque"])
        for name, pair in (("command_x_range", self.command_x_range), ("command_y_range", self.command_y_range),
                           ("command_yaw_range", self.command_yaw_range)):
            arr = getattr(p, name)
            arr[0], arr[1] = float(pair[0]), float(pair[1])
        for j, v in enumerate(self.default_dof_pos[0].tolist()):
            p.default_dof_pos[j] = v
        for c, v in enumerate(self.base_init_state):
            p.base_init_state[c] = float(v)
        p.gravity_vec[0], p.gravity_vec[1], p.gravity_vec[2] = 0.0, 0.0, -1.0
        handle = C.c_void_p()
        nat.check(lib.tk_anymal_create(C.byref(p), C.byref(handle)), "tk_anymal_create")
        self._handle, self._lib = handle, lib
        self._post_physics = lib.tk_anymal_post_physics
        self._torch_device = torch.device(self.device)
        self._philox_noise = nat.AnymalNoise()
        self._philox_noise.mode = nat.NOISE_PHILOX
        self._philox_noise_ref = C.byref(self._philox_noise)
        self._bind()

    def _bind(self):
        b = nat.AnymalBuffers()
        for name, t in (("root_states", self.root_states), ("dof_state", self.dof_state), ("torques", self.torques),
                        ("contact_forces", self.contact_forces), ("commands", self.commands), ("actions", self.actions),
                        ("obs_buf", self.obs_buf), ("rew_buf", self.rew_buf), ("reset_buf", self.reset_buf),
                        ("progress_buf", self.progress_buf), ("timeout_buf", self.timeout_buf),
                        ("episode_return", self.episode_return)):
            if t is not None and (not t.is_contiguous() or t.device != torch.device(self.device)):
                raise ValueError(f"{name}: the kernels need a contiguous tensor on {self.device}")
            setattr(b, name, nat.ptr(t))
        nat.check(self._lib.tk_anymal_bind(self._handle, C.byref(b)), "tk_anymal_bind")

    def rebind(self):
        self._bind()

    def _stream(self):
        return nat.current_stream_ptr(self.device)

    def __del__(self):
        try:
            if getattr(self, "_handle", None):
                self._lib.tk_anymal_destroy(self._handle)
                self._handle = None
        except Exception:
            pass

    # ------------------------------------------------------------------------------------------
    def pre_physics_step(self, actions: torch.Tensor):
        """:229-232 (+ the clamp of vec_task.py:327 when called from ``step``)."""
        a = actions.to(self.device, dtype=torch.float).contiguous()
        nat.check(self._lib.tk_anymal_pre_physics(self._handle, nat.ptr(a), nat.ptr(self.targets),
                                                  1 if self._clip_in_kernel else 0, self._stream()), "tk_anymal_pre_physics")
        self.sim.set_dof_position_target_tensor(self.targets)

    def _refresh_sim_tensors(self):
        """:263-266."""
        for f in ("refresh_dof_state_tensor", "refresh_actor_root_state_tensor", "refresh_net_contact_force_tensor",
                  "refresh_dof_force_tensor"):
            getattr(self.sim, f)()

    def _push_resets(self, ids_int32):
        """anymal.py:302-308: the reset rows reach a simulator that keeps its own state only through these two calls."""
        if ids_int32 is None:
            return
        self.sim.set_actor_root_state_tensor_indexed(self.initial_root_states, ids_int32)
        self.sim.set_dof_state_tensor_indexed(self.dof_state, ids_int32)

    def compute_observations(self):
        """:262-283."""
        self._refresh_sim_tensors()
        nat.check(self._lib.tk_anymal_compute_observations(self._handle, self._stream()), "tk_anymal_compute_observations")

    def compute_reward(self, actions=None):
        """:244-259."""
        nat.check(self._lib.tk_anymal_compute_reward(self._handle, self._stream()), "tk_anymal_compute_reward")

    def reset_idx(self, env_ids: torch.Tensor, unit_pos: Optional[torch.Tensor] = None,
                  unit_vel: Optional[torch.Tensor] = None, unit_cmd: Optional[torch.Tensor] = None):
        """:285-311.  The unit tensors are the U(0,1) draws of :288-289 and :305-307 (drawn here when omitted)."""
        k = len(env_ids)
        if k == 0:
            return
        dev = self.device
        unit_pos = torch.rand((k, NUM_DOF), device=dev) if unit_pos is None else unit_pos.to(dev).contiguous()
        unit_vel = torch.rand((k, NUM_DOF), device=dev) if unit_vel is None else unit_vel.to(dev).contiguous()
        unit_cmd = torch.rand((k, 3), device=dev) if unit_cmd is None else unit_cmd.to(dev).contiguous()
        ids = env_ids.to(device=dev, dtype=torch.long).contiguous()
        nat.check(self._lib.tk_anymal_reset_idx(self._handle, nat.ptr(ids), k, nat.ptr(unit_pos), nat.ptr(unit_vel),
                                                nat.ptr(unit_cmd), self._stream()), "tk_anymal_reset_idx")
        if self._sim_detached():
            self._push_resets(ids.to(torch.int32))                                # anymal.py:301

    def post_physics_step(self, unit_pos: Optional[torch.Tensor] = None, unit_vel: Optional[torch.Tensor] = None,
                          unit_cmd: Optional[torch.Tensor] = None):
        """:234-242 + vec_task.py:345 as one fused kernel.  Supplied draws are addressed by env id ([N, ...]);
        in-kernel Philox noise when none is given."""
        detached = self._sim_detached()
        if detached:                            # a simulator with its own state (sim/detached_sim.py, PhysX)
            self._refresh_sim_tensors()
            self._collect_reset_ids()           # anymal.py:238 -- the fused kernel consumes reset_buf
        stream = nat.current_stream_ptr(self._torch_device)
        if unit_pos is None and unit_vel is None and unit_cmd is None:
            nz = self._philox_noise
            nz.seed, nz.step, nz.env_id_offset = self.noise_seed, self.step_count, self.env_id_offset
            ref = self._philox_noise_ref
        else:
            nz = nat.AnymalNoise()
            nz.mode = nat.NOISE_DENSE
            keep = []
            for field, t in (("unit_pos", unit_pos), ("unit_vel", unit_vel), ("unit_cmd", unit_cmd)):
                if t is not None:
                    t = t.to(self.device, dtype=torch.float).contiguous()
                    if t.shape[0] != self.num_envs:
                        raise ValueError(f"{field}: supplied draws are addressed by env id, need {self.num_envs} rows")
                    keep.append(t)
                setattr(nz, field, nat.ptr(t))
            ref = C.byref(nz)
        rc = self._post_physics(self._handle, ref, stream)
        if rc:
            nat.check(rc, "tk_anymal_post_physics")
        self.step_count += 1
        if detached:
            self._push_resets(self._collected_reset_ids())

    def philox_noise(self, step: Optional[int] = None):
        n, dev = self.num_envs, self.device
        out = (torch.empty((n, NUM_DOF), device=dev), torch.empty((n, NUM_DOF), device=dev), torch.empty((n, 3), device=dev))
        nat.check(self._lib.tk_anymal_philox_noise(self._handle, self.noise_seed, self.step_count if step is None else step,
                                                   self.env_id_offset, nat.ptr(out[0]), nat.ptr(out[1]), nat.ptr(out[2]),
                                                   self._stream()), "tk_anymal_philox_noise")
        return out

    def advance_noise_epoch(self):
        nat.check(self._lib.tk_anymal_advance_epoch(self._handle, self._stream()), "tk_anymal_advance_epoch")

    def episode_statistics(self, accumulate: bool = False) -> torch.Tensor:
        """fp64[8]: sum reward, #reset, #timeout, sum return / length of finished episodes, sum lin / ang velocity error, N."""
        nat.check(self._lib.tk_anymal_stats(self._handle, nat.ptr(self.episode_stats), 1 if accumulate else 0,
                                            self._stream()), "tk_anymal_stats")
        return self.episode_stats
