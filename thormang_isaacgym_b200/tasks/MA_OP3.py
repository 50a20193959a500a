"""``MA_OP3`` -- drop-in for ``isaacgymenvs/tasks/MA_OP3.py::MA_OP3`` on the env-step path:
two 22-DOF ROBOTIS OP3 humanoids carrying a table, tensors laid out ``[N, agents, ...]``.

Same constructor signature, cfg keys and attribute names as the reference class (:14-148); ``compute_observations``
(:411-438), ``compute_reward`` (:360-409) and ``reset_idx`` (:440-489) are single calls into the sm_100a CUDA extension
(include/thormang_b200.h, csrc/tk_ma_op3.cu).  No TorchScript, no CPU fallback.

The reference's step hooks are stubs (``pre_physics_step`` is ``pass``, ``post_physics_step`` zeroes the buffers,
:309-317) and three things they would need are undefined there; this class states its assumptions (the CPU oracle
``oracle/ma_op3_oracle.py`` makes the same ones):

* the step is the ordering of the task the file derives from (anymal.py:229-242);
* ``prev_torques`` (read at :373, never assigned) and ``last_actions`` (only zeroed, :486) are refreshed at the end of
  each step;
* actions are ``[N, agents, 22]`` and an agent's observation row has 88 entries (the reference forces
  ``numObservations = numActions = 1`` at :62-63, with which its own ``torch.cat`` (:697) cannot run).

``_reward_gripper_contact_hold`` (:346-358) reduces over ALL envs; ``grip_counts()`` runs the count kernel and, in a
sharded run (``cfg["env"]["gripAllReduce"]``), all-reduces the four counts over ranks before the reward kernel.
"""
import ctypes as C
from typing import Optional

import torch

from .. import _native as nat
from ..config import OP3_DOF_NAMES
from ..sim.fake_sim import FakeSim
from .base.vec_task import VecTask

NUM_AGENTS = 2
NUM_DOF = 22
NUM_OBS = 88
ACTORS_PER_ENV = 3
DEFAULT_BODIES = 25                      # rigid bodies of one OP3 in the fake-sim layout (the URDF is not in the tree)
DEFAULT_FEET = (10, 11, 16, 17)          # "*ank*" links (:237-238), actor-local
DEFAULT_GRIPPERS = (20, 24)              # "*gr*" links (:241-242), actor-local: left, right


class MA_OP3(VecTask):
    fuses_action_clip = True

    def __init__(self, cfg, rl_device, sim_device, graphics_device_id, headless,
                 virtual_screen_capture=False, force_render=False):
        self.cfg = cfg
        learn, ctrl = cfg["env"]["learn"], cfg["env"]["control"]
        # :17-35
        self.lin_vel_scale = learn["linearVelocityScale"]
        self.ang_vel_scale = learn["angularVelocityScale"]
        self.dof_pos_scale = learn["dofPositionScale"]
        self.dof_vel_scale = learn["dofVelocityScale"]
        self.action_scale = ctrl["actionScale"]
        self.rew_scales = {
            "lin_vel_xy": learn["linearVelocityXYRewardScale"], "ang_vel_z": learn["angularVelocityZRewardScale"],
            "torque": learn["torqueRewardScale"], "heading_scale": learn["headingScale"], "up_scale": learn["upScale"],
            "air_time": learn["feetAirTimeRewardScale"], "syns_hip": learn["syncronizeHipRewardScale"],
            "no_fly": learn["noflyRewardScale"], "stand_scale": learn["standRewardScale"],
            "action_rate": learn["actionRateRewardScale"]}
        self.randomization_params = cfg.get("task", {}).get("randomization_params", {})
        self.randomize = cfg.get("task", {}).get("randomize", False)
        rng = cfg["env"]["randomCommandVelocityRanges"]
        self.command_x_range, self.command_y_range, self.command_yaw_range = rng["linear_x"], rng["linear_y"], rng["yaw"]
        self.command_heading_range = rng.get("heading")
        b = cfg["env"]["baseInitState"]
        self.base_init_state = b["pos"] + b["rot"] + b["vLinear"] + b["vAngular"]                  # :52-58
        self.named_default_joint_angles = cfg["env"]["defaultJointAngles"]
        cfg["env"]["numObservations"] = NUM_OBS          # the reference writes 1 here (:62-63), see the module docstring
        cfg["env"]["numActions"] = NUM_DOF
        self.dt = cfg["sim"]["dt"]
        self.max_episode_length_s = learn["episodeLength_s"]
        self.max_episode_length = int(self.max_episode_length_s / self.dt + 0.5)                    # :69
        self.Kp, self.Kd = ctrl["stiffness"], ctrl["damping"]

        super().__init__(config=cfg, rl_device=rl_device, sim_device=sim_device,
                         graphics_device_id=graphics_device_id, headless=headless,
                         virtual_screen_capture=virtual_screen_capture, force_render=force_render)

        sim, dev, n, a, d = self.sim, self.device, self.num_envs, self.num_agents, self.num_dof
        self._rew_scales_cfg = dict(self.rew_scales)
        for key in self.rew_scales.keys():
            self.rew_scales[key] *= self.dt                                                         # :80-81
        f32 = dict(dtype=torch.float, device=dev)
        # :88-115
        self.root_states = sim.acquire_actor_root_state_tensor().view(n, ACTORS_PER_ENV, 13)
        self.initial_root_states = self.root_states.clone()
        self.initial_root_states[:, :, 7:13] = 0
        self.goal_pos = torch.tensor([0.0, -10.0, 0.0], **f32).repeat((n, 1))
        self.table_state = self.root_states[:, 2]
        self.table_positions = self.table_state[:, 0:3]
        self.table_linvels = self.table_state[:, 7:10]
        self.table_potentials = torch.tensor([-1000. / self.dt], **f32).repeat(n)
        self.prev_table_potentials = self.table_potentials.clone()
        self.dof_state = sim.acquire_dof_state_tensor()
        self.dof_pos = self.dof_state.view(n, a, d, 2)[..., 0]
        self.dof_vel = self.dof_state.view(n, a, d, 2)[..., 1]
        self.default_dof_pos = torch.zeros((n, a, d), **f32)
        self.contact_forces = sim.acquire_net_contact_force_tensor().view(n, -1, 3)
        self.torques = sim.acquire_dof_force_tensor().view(n, a, d)
        self.prev_torques = torch.zeros((n, a, d), **f32)                                           # assumption, see docstring
        # :117-148
        self.commands = torch.zeros(n, 3, **f32)
        self.commands_x, self.commands_y, self.commands_yaw = self.commands[..., 0], self.commands[..., 1], self.commands[..., 2]
        for i in range(d):
            self.default_dof_pos[:, :, i] = self.named_default_joint_angles[self.dof_names[i]]
        self.extras = {}
        self.gravity_vec = torch.tensor([0.0, 0.0, -1.0], **f32).repeat((n, a, 1))
        self.actions = torch.zeros(n, a, d, **f32)
        nf = self.feet_indices.shape[1]
        self.feet_air_time = torch.zeros(n, a, nf, **f32)
        self.last_contacts = torch.zeros(n, a, nf, dtype=torch.bool, device=dev)
        self.last_actions = self.actions.clone()
        self.up_vec = torch.tensor([0.0, 0.0, 1.0], **f32).repeat((n, a + 1, 1))
        self.heading_vec = torch.tensor([1.0, 0.0, 0.0], **f32).repeat((n, a + 1, 1))
        self.basis_vec0, self.basis_vec1 = self.heading_vec.clone(), self.up_vec.clone()
        q = self.start_rotation
        self.inv_start_rot = torch.tensor([-q[0], -q[1], -q[2], q[3]], **f32).repeat((n, a, 1))
        self.targets = torch.tensor([0.0, 10.0, 0.0], **f32).repeat((n, 1))
        self.potentials = torch.tensor([-1000. / self.dt], **f32).repeat(n, a)
        self.prev_potentials = self.potentials.clone()
        self.pd_targets = torch.zeros(n, a, d, **f32)
        self.episode_return = torch.zeros(n, device=dev) if cfg["env"].get("episodeStatistics", True) else None
        self.episode_stats = torch.zeros(nat.TK_STATS, device=dev, dtype=torch.float64)
        self.noise_seed = int(cfg.get("seed", 0)) & 0xFFFFFFFFFFFFFFFF
        self.env_id_offset = int(cfg["env"].get("envIdOffset", 0))
        self.grip_all_reduce = bool(cfg["env"].get("gripAllReduce", False))
        self.grip_count_buf = torch.zeros((a, 2), dtype=torch.long, device=dev)
        self.step_count = 0
        self._handle = None
        self._create_native()
        if cfg["env"].get("resetOnInit", True):
            self.reset_idx(torch.arange(n, device=dev))                                             # :148

    def allocate_buffers(self):
        """multi_vec_task.py:102-125: obs ``[N, agents, obs]``, rewards ``[N, agents]``."""
        super().allocate_buffers()
        n, dev = self.num_envs, self.device
        self.obs_buf = torch.zeros((n, self.num_agents, self.num_obs), device=dev, dtype=torch.float)
        self.rew_buf = torch.zeros((n, self.num_agents), device=dev, dtype=torch.float)

    def zero_actions(self) -> torch.Tensor:
        """multi_vec_task.py:209-217."""
        return torch.zeros([self.num_envs, self.num_agents, self.num_actions], dtype=torch.float32, device=self.rl_device)

    def create_sim(self):
        """:150-306 without PhysX: injected simulator or the fake-sim shim."""
        n = self.cfg["env"]["numEnvs"]
        sim = self.cfg.get("sim_backend")
        if sim is None:
            sim = FakeSim(n, NUM_AGENTS * NUM_DOF, self.device, num_sensor_floats=0, dof_names=list(OP3_DOF_NAMES),
                          num_bodies=NUM_AGENTS * DEFAULT_BODIES + 1, actors_per_env=ACTORS_PER_ENV)
            base = torch.tensor([[-0.31, 0.0, 0.27], [0.30, 0.0, 0.27], [0.0, 0.0, 0.30]], device=self.device)   # :221-231
            sim.root_states.view(n, ACTORS_PER_ENV, 13)[:, :, 0:3] = base
        self.sim = sim
        self.up_axis_idx = 2
        self.num_dof = NUM_DOF
        self.num_bodies_total = sim.num_bodies
        self.dof_names = list(sim.dof_names)[:NUM_DOF]
        env = self.cfg["env"]
        per = (sim.num_bodies - 1) // NUM_AGENTS
        feet = getattr(sim, "feet_indices", env.get("feetIndices", [[a * per + b for b in DEFAULT_FEET] for a in range(NUM_AGENTS)]))
        grip = getattr(sim, "gripper_indices", env.get("gripperIndices", [[a * per + b for b in DEFAULT_GRIPPERS] for a in range(NUM_AGENTS)]))
        self.feet_indices = torch.as_tensor(feet, dtype=torch.long, device=self.device)              # :294-296
        self.gripper_indices = torch.as_tensor(grip, dtype=torch.long, device=self.device)           # :300-302
        self.start_rotation = list(getattr(sim, "start_rotation", (0.0, 0.0, 0.0, 1.0)))             # :233-236

    def _create_native(self):
        lib = nat.lib()
        p = nat.Op3Params()
        p.device, p.num_envs, p.num_agents, p.num_dof = self.device_id, self.num_envs, self.num_agents, self.num_dof
        p.num_bodies_total, p.max_episode_length = self.num_bodies_total, self.max_episode_length
        feet, grip = self.feet_indices.tolist(), self.gripper_indices.tolist()
        p.num_feet, p.num_grippers = len(feet[0]), len(grip[0])
        for a in range(self.num_agents):
            for i, v in enumerate(feet[a]):
                p.feet_indices[a][i] = int(v)
            for i, v in enumerate(grip[a][:nat.OP3_MAX_GRIPPERS]):
                p.gripper_indices[a][i] = int(v)
        p.lin_vel_scale, p.ang_vel_scale = self.lin_vel_scale, self.ang_vel_scale
        p.dof_pos_scale, p.dof_vel_scale, p.action_scale = self.dof_pos_scale, self.dof_vel_scale, self.action_scale
        p.clip_actions = float(self.clip_actions)
        p.dt = float(self.dt)
        s = self._rew_scales_cfg
        p.rew_torque, p.rew_up, p.rew_air_time = float(s["torque"]), float(s["up_scale"]), float(s["air_time"])
        p.rew_syns_hip, p.rew_no_fly, p.rew_stand = float(s["syns_hip"]), float(s["no_fly"]), float(s["stand_scale"])
        p.rew_action_rate = float(s["action_rate"])
        for name, pair in (("command_x_range", self.command_x_range), ("command_y_range", self.command_y_range),
                           ("command_yaw_range", self.command_yaw_range)):
            arr = getattr(p, name)
            arr[0], arr[1] = float(pair[0]), float(pair[1])
        for j, v in enumerate(self.default_dof_pos[0, 0].tolist()):
            p.default_dof_pos[j] = v
        for c, v in enumerate(self.inv_start_rot[0, 0].tolist()):
            p.inv_start_rot[c] = v
        for c in range(3):
            p.gravity_vec[c] = (0.0, 0.0, -1.0)[c]
            p.up_vec[c] = (0.0, 0.0, 1.0)[c]
            p.targets[c] = (0.0, 10.0, 0.0)[c]
            p.goal_pos[c] = (0.0, -10.0, 0.0)[c]
        handle = C.c_void_p()
        nat.check(lib.tk_ma_op3_create(C.byref(p), C.byref(handle)), "tk_ma_op3_create")
        self._handle, self._lib = handle, lib
        self._torch_device = torch.device(self.device)
        self._philox_noise = nat.Op3Noise()
        self._philox_noise.mode = nat.NOISE_PHILOX
        self._bind()

    def _bind(self):
        b = nat.Op3Buffers()
        self._last_contacts_u8 = self.last_contacts.view(torch.uint8)
        for name, t in (("root_states", self.root_states), ("initial_root_states", self.initial_root_states),
                        ("dof_state", self.dof_state), ("torques", self.torques), ("prev_torques", self.prev_torques),
                        ("contact_forces", self.contact_forces), ("commands", self.commands), ("actions", self.actions),
                        ("last_actions", self.last_actions), ("feet_air_time", self.feet_air_time),
                        ("last_contacts", self._last_contacts_u8), ("potentials", self.potentials),
                        ("prev_potentials", self.prev_potentials), ("table_potentials", self.table_potentials),
                        ("prev_table_potentials", self.prev_table_potentials), ("obs_buf", self.obs_buf),
                        ("rew_buf", self.rew_buf), ("reset_buf", self.reset_buf), ("progress_buf", self.progress_buf),
                        ("timeout_buf", self.timeout_buf), ("episode_return", self.episode_return),
                        ("grip_counts", self.grip_count_buf)):
            if t is not None and (not t.is_contiguous() or t.device != torch.device(self.device)):
                raise ValueError(f"{name}: the kernels need a contiguous tensor on {self.device}")
            setattr(b, name, nat.ptr(t))
        nat.check(self._lib.tk_ma_op3_bind(self._handle, C.byref(b)), "tk_ma_op3_bind")

    def rebind(self):
        self._bind()

    def _stream(self):
        return nat.current_stream_ptr(self.device)

    def __del__(self):
        try:
            if getattr(self, "_handle", None):
                self._lib.tk_ma_op3_destroy(self._handle)
                self._handle = None
        except Exception:
            pass

    # ------------------------------------------------------------------------------------------
    def pre_physics_step(self, actions: torch.Tensor):
        """Un-stubbed (:309-310 is ``pass``): anymal.py:229-232 + the clamp of vec_task.py:327 when called from ``step``."""
        a = actions.to(self.device, dtype=torch.float).contiguous()
        nat.check(self._lib.tk_ma_op3_pre_physics(self._handle, nat.ptr(a), nat.ptr(self.pd_targets),
                                                  1 if self._clip_in_kernel else 0, self._stream()), "tk_ma_op3_pre_physics")
        self.sim.set_dof_position_target_tensor(self.pd_targets)

    def grip_counts(self) -> torch.Tensor:
        """int64 [agents, 2] (device): envs pressing the left / right gripper (:351-352), summed over ranks when sharded."""
        nat.check(self._lib.tk_ma_op3_grip_counts(self._handle, self._stream()), "tk_ma_op3_grip_counts")
        self._all_reduce_counts()
        return self.grip_count_buf

    def _all_reduce_counts(self):
        if self.grip_all_reduce and torch.distributed.is_available() and torch.distributed.is_initialized() \
                and torch.distributed.get_world_size() > 1:
            torch.distributed.all_reduce(self.grip_count_buf)        # the one exchange step of this path (NCCL, 32 bytes)

    def compute_observations(self):
        """:411-438."""
        for f in ("refresh_dof_state_tensor", "refresh_actor_root_state_tensor", "refresh_net_contact_force_tensor",
                  "refresh_dof_force_tensor"):
            getattr(self.sim, f)()
        nat.check(self._lib.tk_ma_op3_compute_observations(self._handle, self._stream()), "tk_ma_op3_compute_observations")

    def compute_reward(self, actions=None):
        """:360-409."""
        self.grip_counts()
        nat.check(self._lib.tk_ma_op3_compute_reward(self._handle, self._stream()), "tk_ma_op3_compute_reward")

    def reset_idx(self, env_ids: torch.Tensor, unit_cmd: Optional[torch.Tensor] = None):
        """:440-489.  ``unit_cmd`` [K,3]: the U(0,1) draws of :472-474 (drawn here when omitted)."""
        k = len(env_ids)
        if k == 0:
            return
        dev = self.device
        unit_cmd = torch.rand((k, 3), device=dev) if unit_cmd is None else unit_cmd.to(dev, dtype=torch.float).contiguous()
        ids = env_ids.to(device=dev, dtype=torch.long).contiguous()
        nat.check(self._lib.tk_ma_op3_reset_idx(self._handle, nat.ptr(ids), k, nat.ptr(unit_cmd), self._stream()),
                  "tk_ma_op3_reset_idx")

    def post_physics_step(self, unit_cmd: Optional[torch.Tensor] = None):
        self.launch_post_physics(unit_cmd)
        self._all_reduce_counts()
        self.finalize_reward()

    def finalize_reward(self):
        """Second half of the step: the batch-global hold term, after ``grip_count_buf`` holds the GLOBAL counts."""
        nat.check(self._lib.tk_ma_op3_finalize_reward(self._handle, self._stream()), "tk_ma_op3_finalize_reward")

    def launch_post_physics(self, unit_cmd: Optional[torch.Tensor] = None):
        """progress += 1, reset, compute_observations, compute_reward, prev_torques / last_actions refresh and the timeout
        line (multi_vec_task.py:193): ONE fused kernel that also counts the gripper contacts, then (after the all-reduce
        of the four counts in a sharded run) a finalize grid that adds the batch-global hold term -- it exits at once
        unless exactly one env presses each gripper.  ``unit_cmd`` [N,3] is addressed by env id; in-kernel Philox noise
        when omitted."""
        if unit_cmd is None:
            nz = self._philox_noise
            nz.seed, nz.step, nz.env_id_offset = self.noise_seed, self.step_count, self.env_id_offset
        else:
            nz = nat.Op3Noise()
            nz.mode = nat.NOISE_DENSE
            unit_cmd = unit_cmd.to(self.device, dtype=torch.float).contiguous()
            if unit_cmd.shape[0] != self.num_envs:
                raise ValueError(f"unit_cmd: supplied draws are addressed by env id, need {self.num_envs} rows")
            nz.unit_cmd = nat.ptr(unit_cmd)
        nat.check(self._lib.tk_ma_op3_post_physics(self._handle, C.byref(nz), self._stream()), "tk_ma_op3_post_physics")
        self.step_count += 1

    def philox_noise(self, step: Optional[int] = None) -> torch.Tensor:
        out = torch.empty((self.num_envs, 3), device=self.device)
        nat.check(self._lib.tk_ma_op3_philox_noise(self._handle, self.noise_seed, self.step_count if step is None else step,
                                                   self.env_id_offset, nat.ptr(out), self._stream()), "tk_ma_op3_philox_noise")
        return out

    def advance_noise_epoch(self):
        nat.check(self._lib.tk_ma_op3_advance_epoch(self._handle, self._stream()), "tk_ma_op3_advance_epoch")

    def episode_statistics(self, accumulate: bool = False) -> torch.Tensor:
        """fp64[8]: sum reward (both agents), #reset, #timeout, sum return / length of finished episodes, sum distance of
        agent 0 / agent 1 to the table, N."""
        nat.check(self._lib.tk_ma_op3_stats(self._handle, nat.ptr(self.episode_stats), 1 if accumulate else 0,
                                            self._stream()), "tk_ma_op3_stats")
        return self.episode_stats
