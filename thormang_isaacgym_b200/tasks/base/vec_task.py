"""VecTask surface of the reference, kept as the drop-in boundary.

Mirrors ``isaacgymenvs/tasks/base/vec_task.py`` of the reference: ``Env`` (:60-162) and
``VecTask`` (:165-406) -- constructor signature, buffers (``allocate_buffers`` :254-277),
``step`` (:313-359), ``reset`` (:377-389), ``reset_done`` (:391-406), ``zero_actions``
(:361-369), ``get_state`` (:297-299) and the space / count properties (:139-162).

Left out on purpose: viewer and ``render`` (:225-252,408-440;
no graphics stack on this box), PhysX sim-parameter parsing (:442-490) and domain
randomisation through PhysX property setters (:538-768).  The simulator is injected
(``cfg["sim_backend"]``) or replaced by the fake-sim shim.
"""
import abc
import ctypes
from dataclasses import dataclass
from typing import Any, Dict, Optional, Tuple

import numpy as np
import torch


@dataclass
class Box:
    """Minimal stand-in for ``gym.spaces.Box`` (gym is not installed here; vec_task.py:102-105)."""
    low: np.ndarray
    high: np.ndarray

    @property
    def shape(self):
        return self.low.shape


class Env(abc.ABC):
    def __init__(self, config: Dict[str, Any], rl_device: str, sim_device: str, graphics_device_id: int, headless: bool):
        split_device = sim_device.split(":")                                    # vec_task.py:71-73
        self.device_type = split_device[0]
        self.device_id = int(split_device[1]) if len(split_device) > 1 else 0
        self.device = "cpu"
        if config["sim"].get("use_gpu_pipeline", True):                          # vec_task.py:76-81
            if self.device_type.lower() in ("cuda", "gpu"):
                self.device = "cuda:" + str(self.device_id)
            else:
                config["sim"]["use_gpu_pipeline"] = False
        if not str(self.device).startswith("cuda"):
            raise RuntimeError("thormang_isaacgym_b200 runs the env-step path on a CUDA device only "
                               f"(sim_device={sim_device!r}); there is no CPU fallback")
        self.rl_device = rl_device
        self.headless = headless
        self.graphics_device_id = -1                                             # no graphics stack here
        self.num_environments = config["env"]["numEnvs"]
        self.num_agents = config["env"].get("numAgents", 1)
        self.num_observations = config["env"]["numObservations"]
        self.num_states = config["env"].get("numStates", 0)
        self.num_actions = config["env"]["numActions"]
        self.control_freq_inv = config["env"].get("controlFrequencyInv", 1)
        self.obs_space = Box(np.ones(self.num_obs) * -np.inf, np.ones(self.num_obs) * np.inf)
        self.state_space = Box(np.ones(self.num_states) * -np.inf, np.ones(self.num_states) * np.inf)
        self.act_space = Box(np.ones(self.num_actions) * -1., np.ones(self.num_actions) * 1.)
        self.clip_obs = config["env"].get("clipObservations", np.inf)
        self.clip_actions = config["env"].get("clipActions", np.inf)

    @abc.abstractmethod
    def allocate_buffers(self): ...

    @abc.abstractmethod
    def step(self, actions: torch.Tensor) -> Tuple[Dict[str, torch.Tensor], torch.Tensor, torch.Tensor, Dict[str, Any]]: ...

    @abc.abstractmethod
    def reset(self) -> Dict[str, torch.Tensor]: ...

    @abc.abstractmethod
    def reset_idx(self, env_ids: torch.Tensor): ...

    @property
    def observation_space(self): return self.obs_space

    @property
    def action_space(self): return self.act_space

    @property
    def num_envs(self) -> int: return self.num_environments

    @property
    def num_acts(self) -> int: return self.num_actions

    @property
    def num_obs(self) -> int: return self.num_observations


class VecTask(Env):
    # tasks whose pre-physics kernel applies the +-clipActions clamp of vec_task.py:327 itself
    fuses_action_clip = False

    def __init__(self, config, rl_device, sim_device, graphics_device_id, headless,
                 virtual_screen_capture: bool = False, force_render: bool = False):
        super().__init__(config, rl_device, sim_device, graphics_device_id, headless)
        self.virtual_screen_capture = False      # no graphics stack: accepted and ignored
        self.force_render = False
        self.viewer = None
        self.dr_randomizations = {}
        self.sim_initialized = False
        self.create_sim()
        self.sim_initialized = True
        self.allocate_buffers()
        self.obs_dict = {}
        self._clip_in_kernel = False
        self._rollout, self._rollout_t = None, 0
        self.first_randomization, self.last_step, self.last_rand_step = True, -1, -1

    def allocate_buffers(self):
        """vec_task.py:254-277: same names, shapes, dtypes and initial values."""
        dev = self.device
        n = self.num_envs
        self.output_slab = None
        if self.cfg["env"].get("outputSlab", False):
            # host-driven pipelines (host_io.HostBridge): obs | rew | reset carved out of ONE allocation so that a step's
            # results leave the device as a single copy.  Same shapes, dtypes and initial values as below.
            def up(x): return (x + 255) // 256 * 256
            o_obs, o_rew = 0, up(n * self.num_obs * 4)
            o_reset = o_rew + up(n * 4)
            self.output_slab = torch.zeros(o_reset + n * 8, device=dev, dtype=torch.uint8)
            self.obs_buf = self.output_slab[o_obs:o_obs + n * self.num_obs * 4].view(torch.float).view(n, self.num_obs)
            self.rew_buf = self.output_slab[o_rew:o_rew + n * 4].view(torch.float)
            self.reset_buf = self.output_slab[o_reset:o_reset + n * 8].view(torch.long)
            self.reset_buf.fill_(1)
        else:
            self.obs_buf = torch.zeros((n, self.num_obs), device=dev, dtype=torch.float)
            self.rew_buf = torch.zeros(n, device=dev, dtype=torch.float)
            self.reset_buf = torch.ones(n, device=dev, dtype=torch.long)
        self.states_buf = torch.zeros((self.num_envs, self.num_states), device=dev, dtype=torch.float)
        # the reference allocates int64 and rebinds to a bool tensor in step() (vec_task.py:345); the kernels
        # write the bool directly
        self.timeout_buf = torch.zeros(self.num_envs, device=dev, dtype=torch.bool)
        self.progress_buf = torch.zeros(self.num_envs, device=dev, dtype=torch.long)
        self.randomize_buf = torch.zeros(self.num_envs, device=dev, dtype=torch.long)
        self.extras = {}

    @abc.abstractmethod
    def create_sim(self): ...

    def get_state(self):
        return torch.clamp(self.states_buf, -self.clip_obs, self.clip_obs).to(self.rl_device)

    @abc.abstractmethod
    def pre_physics_step(self, actions: torch.Tensor): ...

    @abc.abstractmethod
    def post_physics_step(self): ...

    # ---- reset hand-off to a simulator that keeps its own state -----------------------------------------
    # The reference's reset_idx passes env_ids_int32 to gym.set_actor_root_state_tensor_indexed / set_dof_state_tensor_indexed
    # (humanoid.py:264-271, gogoro_new.py:547,552).  The fused kernels reset rows of the task's tensors and overwrite
    # reset_buf, so the id list is taken (tk_reset_ids) BEFORE the fused step and the rows are pushed behind it.  A backend
    # that shares memory with the task (FakeSim.aliases_task_memory) needs neither, and pays nothing.
    def _sim_detached(self) -> bool:
        d = getattr(self, "_detached", None)
        if d is None:
            d = self._detached = not getattr(self.sim, "aliases_task_memory", False)
        return d

    def _collect_reset_ids(self, stream=None):
        """Enqueue env_ids = reset_buf.nonzero() -> int32 on the task's stream (no sync)."""
        from ... import _native as nat
        if getattr(self, "_reset_ids", None) is None:
            n, dev = self.num_envs, self.device
            lib = nat.lib()
            self._reset_ids = torch.empty(n, device=dev, dtype=torch.int32)
            self._reset_count = torch.zeros(1, device=dev, dtype=torch.int32)
            self._reset_scratch = torch.empty(max(1, lib.tk_reset_ids_scratch_bytes(n) // 4), device=dev, dtype=torch.int32)
            self._reset_ids_fn = lib.tk_reset_ids
        if stream is None:
            stream = nat.current_stream_ptr(torch.device(self.device))
        nat.check(self._reset_ids_fn(nat.ptr(self.reset_buf), self.num_envs, nat.ptr(self._reset_ids), nat.ptr(self._reset_count),
                                     nat.ptr(self._reset_scratch), stream), "tk_reset_ids")

    def _collected_reset_ids(self) -> Optional[torch.Tensor]:
        """The id list of the last _collect_reset_ids (int32, ascending) or None when it is empty.  Reads the count back:
        the same host synchronisation the reference's nonzero() + len(env_ids) makes (humanoid.py:290-291)."""
        k = int(self._reset_count.item())
        return self._reset_ids[:k] if k > 0 else None

    # ---- step prologue / epilogue as one fused kernel each -----------------------------------------
    def _noise_clamp(self, tensor, which, clip, noisy_out, clamped_out, unit=None):
        """``clamp(noise_lambda(tensor), +-clip)`` of vec_task.py:323-327 / :347-353 in ONE pass (csrc/tk_epilogue.cu).
        ``which``: 'observations' | 'actions'.  ``unit``: the draws of this call (tests); in-kernel Philox otherwise."""
        from ... import _native as nat
        params = self.dr_randomizations.get(which)
        lam, corr = None, None
        if params:
            lam = params["_lambda"]
            corr = params.get("corr")
            if corr is None:                       # drawn on first use and kept until the next randomisation (:613-616)
                corr = torch.randn_like(tensor)
                params["corr"] = corr
        self._epilogue_calls = getattr(self, "_epilogue_calls", 0) + 1
        nat.check(nat.lib().tk_step_noise_clamp(
            nat.ptr(tensor), tensor.numel(), None if lam is None else ctypes.byref(lam), nat.ptr(corr), nat.ptr(unit),
            int(self.cfg.get("seed", 0)) & 0xFFFFFFFFFFFFFFFF, (self._epilogue_calls << 1) | (which == "actions"),
            int(self.cfg["env"].get("envIdOffset", 0)) * tensor.shape[-1], float(clip), nat.ptr(noisy_out),
            nat.ptr(clamped_out), nat.current_stream_ptr(self.device)), "tk_step_noise_clamp")

    def set_rollout_buffer(self, buf: Optional[torch.Tensor]):
        """``buf`` [horizon, N, num_obs] on the sim device: every ``step`` writes ``obs_dict["obs"]`` straight into slot
        ``t`` (t = 0, 1, ... wrapping at the horizon) -- rl_games' rollout tensor filled by the kernels instead of a clamp
        copy (vec_task.py:353) plus a copy into the experience buffer.  None switches it off."""
        if buf is not None and (buf.dim() != 3 or buf.shape[1:] != (self.num_envs, self.num_obs) or not buf.is_contiguous()
                                or buf.dtype != torch.float or str(buf.device) != str(self.device)):
            raise ValueError(f"rollout buffer must be a contiguous float [horizon, {self.num_envs}, {self.num_obs}] tensor on {self.device}")
        self._rollout, self._rollout_t = buf, 0

    def _obs_copy_hook(self, dst):
        """Tasks whose fused kernel can store its obs tile to a second destination override this and return True."""
        return False

    def _clipped_obs(self, filled_by_kernel: bool = False):
        """vec_task.py:347-353, called by ``step`` only: observation noise lambda + clamp (+ the rollout slot) in one pass.
        ``filled_by_kernel``: the fused post-physics kernel of THIS step has already stored its obs tile into the slot."""
        rollout = getattr(self, "_rollout", None)
        noisy = bool(self.dr_randomizations.get("observations"))
        if rollout is not None:
            slot = rollout[self._rollout_t]
            if not filled_by_kernel:
                self._noise_clamp(self.obs_buf, "observations", self.clip_obs, self.obs_buf if noisy else None, slot)
            self._rollout_t = (self._rollout_t + 1) % rollout.shape[0]
            return slot.to(self.rl_device)
        if not noisy and np.isinf(self.clip_obs):
            # clamp(+-inf) is the identity: skip the extra full-obs copy the reference makes (vec_task.py:353)
            return self.obs_buf.to(self.rl_device)
        out = torch.empty_like(self.obs_buf)
        self._noise_clamp(self.obs_buf, "observations", self.clip_obs, self.obs_buf if noisy else None, out)
        return out.to(self.rl_device)

    def _clamped_obs_only(self):
        """vec_task.py:386 / :400 (``reset`` / ``reset_done``): clamp only -- no noise lambda, no rollout slot."""
        if np.isinf(self.clip_obs):
            return self.obs_buf.to(self.rl_device)
        from ... import _native as nat
        out = torch.empty_like(self.obs_buf)
        nat.check(nat.lib().tk_step_noise_clamp(
            nat.ptr(self.obs_buf), self.obs_buf.numel(), None, None, None, 0, 0, 0, float(self.clip_obs), None,
            nat.ptr(out), nat.current_stream_ptr(self.device)), "tk_step_noise_clamp")
        return out.to(self.rl_device)

    def step(self, actions: torch.Tensor):
        """vec_task.py:313-359 without render(); the DR lambdas and clamps run as fused kernels."""
        if self.dr_randomizations.get("actions"):                                    # vec_task.py:323-327
            a = actions.to(self.device, dtype=torch.float).contiguous()
            action_tensor = torch.empty_like(a)
            self._noise_clamp(a, "actions", self.clip_actions, None, action_tensor)
            self.pre_physics_step(action_tensor)
        elif self.fuses_action_clip:
            self._clip_in_kernel = True
            self.pre_physics_step(actions)
            self._clip_in_kernel = False
        else:
            action_tensor = actions
            if not np.isinf(self.clip_actions):
                a = actions.to(self.device, dtype=torch.float).contiguous()      # the kernel takes a device fp32 pointer
                action_tensor = torch.empty_like(a)
                self._noise_clamp(a, "actions", self.clip_actions, None, action_tensor)
            self.pre_physics_step(action_tensor)
        for _ in range(self.control_freq_inv):
            self.sim.simulate()
        # the fused kernel stores its obs tile into the rollout slot as well when nothing has to be applied to it
        rollout = getattr(self, "_rollout", None)
        filled = False                      # local to this step: reset()/reset_done() never see a stale flag
        if rollout is not None and np.isinf(self.clip_obs) and not self.dr_randomizations.get("observations"):
            filled = self._obs_copy_hook(rollout[self._rollout_t])
        # post_physics_step also fills timeout_buf (vec_task.py:345) inside the fused kernel
        self.post_physics_step()
        if filled:
            self._obs_copy_hook(None)
        self.extras["time_outs"] = self.timeout_buf.to(self.rl_device)
        self.obs_dict["obs"] = self._clipped_obs(filled)
        if self.num_states > 0:
            self.obs_dict["states"] = self.get_state()
        return self.obs_dict, self.rew_buf.to(self.rl_device), self.reset_buf.to(self.rl_device), self.extras

    def apply_randomizations(self, dr_params):
        """vec_task.py:538-646, the NON-PHYSICAL part: the observation / action noise lambdas (gaussian | uniform,
        additive | scaling, linear | constant schedule, correlated component).  Same ``dr_randomizations`` entries as the
        reference ('mu', 'var', 'mu_corr', 'var_corr' | 'lo', 'hi', 'lo_corr', 'hi_corr', 'noise_lambda'); the lambda is
        one fused kernel.  The PhysX property setters (:648-768) are out of scope."""
        from ... import _native as nat
        rand_freq = dr_params.get("frequency", 1)                                    # :547
        self.last_step = int(getattr(self.sim, "frame_count", self.last_step + 1))   # :553 gym.get_frame_count (FakeSim counts simulate() calls)
        if self.first_randomization:
            do_nonenv_randomize = True                                               # :555
        else:
            do_nonenv_randomize = (self.last_step - self.last_rand_step) >= rand_freq    # :558
        if do_nonenv_randomize:
            self.last_rand_step = self.last_step                                     # :565
        for name in ("observations", "actions"):                                     # :576
            if name not in dr_params or not do_nonenv_randomize:
                continue
            prm = dr_params[name]
            dist, op_type = prm["distribution"], prm["operation"]
            sched_type = prm["schedule"] if "schedule" in prm else None
            sched_step = prm["schedule_steps"] if "schedule" in prm else None
            if sched_type == "linear":
                sched_scaling = 1.0 / sched_step * min(self.last_step, sched_step)   # :585-586
            elif sched_type == "constant":
                sched_scaling = 0 if self.last_step < sched_step else 1              # :588
            else:
                sched_scaling = 1
            a, b = prm["range"]
            a_corr, b_corr = prm.get("range_correlated", [0., 0.])
            if dist == "gaussian":                                                   # a = mu, b = var
                if op_type == "additive":
                    a, b, a_corr, b_corr = a * sched_scaling, b * sched_scaling, a_corr * sched_scaling, b_corr * sched_scaling
                elif op_type == "scaling":
                    b = b * sched_scaling
                    a = a * sched_scaling + 1.0 * (1.0 - sched_scaling)
                    b_corr = b_corr * sched_scaling
                    a_corr = a_corr * sched_scaling + 1.0 * (1.0 - sched_scaling)
                entry = {"mu": a, "var": b, "mu_corr": a_corr, "var_corr": b_corr}
            elif dist == "uniform":                                                  # a = lo, b = hi
                if op_type == "additive":
                    a, b, a_corr, b_corr = a * sched_scaling, b * sched_scaling, a_corr * sched_scaling, b_corr * sched_scaling
                elif op_type == "scaling":
                    a = a * sched_scaling + 1.0 * (1.0 - sched_scaling)
                    b = b * sched_scaling + 1.0 * (1.0 - sched_scaling)
                    a_corr = a_corr * sched_scaling + 1.0 * (1.0 - sched_scaling)
                    b_corr = b_corr * sched_scaling + 1.0 * (1.0 - sched_scaling)
                entry = {"lo": a, "hi": b, "lo_corr": a_corr, "hi_corr": b_corr}
            else:
                raise ValueError(f"unknown distribution {dist!r}")
            lam = nat.NoiseLambda()
            lam.distribution = 1 if dist == "gaussian" else 2
            lam.operation = 0 if op_type == "additive" else 1
            lam.range[0], lam.range[1] = float(a), float(b)
            lam.range_correlated[0], lam.range_correlated[1] = float(a_corr), float(b_corr)
            entry["_lambda"] = lam

            def noise_lambda(tensor, param_name=name):                               # :612-620 / :635-641
                out = torch.empty_like(tensor)
                self._noise_clamp(tensor.contiguous(), param_name, float("inf"), out, None)
                return out
            entry["noise_lambda"] = noise_lambda
            self.dr_randomizations[name] = entry
        self.first_randomization = False

    def zero_actions(self) -> torch.Tensor:
        return torch.zeros([self.num_envs, self.num_actions], dtype=torch.float32, device=self.rl_device)

    def reset_idx(self, env_idx):
        pass

    def reset(self):
        self.obs_dict["obs"] = self._clamped_obs_only()                              # vec_task.py:386
        if self.num_states > 0:
            self.obs_dict["states"] = self.get_state()
        return self.obs_dict

    def reset_done(self):
        done_env_ids = self.reset_buf.nonzero(as_tuple=False).flatten()
        if len(done_env_ids) > 0:
            self.reset_idx(done_env_ids)
        self.obs_dict["obs"] = self._clamped_obs_only()                              # vec_task.py:400
        if self.num_states > 0:
            self.obs_dict["states"] = self.get_state()
        return self.obs_dict, done_env_ids

    def render(self, mode="rgb_array"):
        """Viewer / screen capture: unavailable on this box (no graphics stack)."""
        return None
