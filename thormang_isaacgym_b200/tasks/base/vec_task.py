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
from dataclasses import dataclass
from typing import Any, Dict, Tuple

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

    def _clipped_obs(self):
        # clamp(+-inf) is the identity: skip the extra full-obs copy the reference makes (vec_task.py:353)
        if np.isinf(self.clip_obs):
            return self.obs_buf.to(self.rl_device)
        return torch.clamp(self.obs_buf, -self.clip_obs, self.clip_obs).to(self.rl_device)

    def step(self, actions: torch.Tensor):
        """vec_task.py:313-359 without the DR lambdas and render()."""
        if self.fuses_action_clip:
            self._clip_in_kernel = True
            self.pre_physics_step(actions)
            self._clip_in_kernel = False
        else:
            action_tensor = actions if np.isinf(self.clip_actions) else torch.clamp(actions, -self.clip_actions, self.clip_actions)
            self.pre_physics_step(action_tensor)
        for _ in range(self.control_freq_inv):
            self.sim.simulate()
        # post_physics_step also fills timeout_buf (vec_task.py:345) inside the fused kernel
        self.post_physics_step()
        self.extras["time_outs"] = self.timeout_buf.to(self.rl_device)
        self.obs_dict["obs"] = self._clipped_obs()
        if self.num_states > 0:
            self.obs_dict["states"] = self.get_state()
        return self.obs_dict, self.rew_buf.to(self.rl_device), self.reset_buf.to(self.rl_device), self.extras

    def zero_actions(self) -> torch.Tensor:
        return torch.zeros([self.num_envs, self.num_actions], dtype=torch.float32, device=self.rl_device)

    def reset_idx(self, env_idx):
        pass

    def reset(self):
        self.obs_dict["obs"] = self._clipped_obs()
        if self.num_states > 0:
            self.obs_dict["states"] = self.get_state()
        return self.obs_dict

    def reset_done(self):
        done_env_ids = self.reset_buf.nonzero(as_tuple=False).flatten()
        if len(done_env_ids) > 0:
            self.reset_idx(done_env_ids)
        self.obs_dict["obs"] = self._clipped_obs()
        if self.num_states > 0:
            self.obs_dict["states"] = self.get_state()
        return self.obs_dict, done_env_ids

    def render(self, mode="rgb_array"):
        """Viewer / screen capture: unavailable on this box (no graphics stack)."""
        return None
