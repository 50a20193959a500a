"""Fake-sim shim: plain torch tensors standing in for the PhysX-owned buffers.

The reference tasks never own their state: ``gymtorch.wrap_tensor`` hands them views of
simulator memory (isaacgymenvs/tasks/humanoid.py:78-98, gogoro_new.py:125-138) and the
``gymapi`` calls are device-pointer hand-offs.  PhysX / isaacgym cannot be installed here,
so this class provides the same tensors in the same layout and the handful of calls the
hot path makes:

* ``refresh_*``                                 no-op (the tensors are always current);
* ``set_actor_root_state_tensor_indexed(src, ids)``   row copy ``root[ids] = src[ids]``;
* ``set_dof_state_tensor_indexed(src, ids)``    row copy unless ``src`` IS the dof-state tensor;
* ``set_dof_actuation_force_tensor`` / ``set_dof_position_target_tensor`` /
  ``set_dof_velocity_target_tensor``             keep a reference (``sim.actuation_force`` ...);
* ``simulate()``                                 calls an optional user ``physics(sim)`` callback
  (identity by default), i.e. the "fake physics" of BASELINE config 5.

Any object with these attributes can be injected instead (``cfg["sim_backend"]``), e.g. a
real simulator exposing its device buffers through ``torch.as_tensor``/DLPack.
"""
from typing import Callable, Dict, List, Optional

import torch


class FakeSim:
    # The task's tensors ARE this simulator's state: resets written by the kernels need no push and nothing has to be
    # refreshed.  A backend without this attribute (or with False) is treated as keeping its own state: the tasks then call
    # refresh_* in front of the fused step and set_*_indexed with the reset id list behind it (sim/detached_sim.py).
    aliases_task_memory = True

    def __init__(self, num_envs: int, num_dof: int, device, num_sensor_floats: int = 12,
                 dof_names: Optional[List[str]] = None, asset: Optional[Dict[str, torch.Tensor]] = None,
                 physics: Optional[Callable[["FakeSim"], None]] = None, slab: bool = False, num_bodies: int = 0,
                 actors_per_env: int = 1):
        self.num_envs, self.num_dof, self.device = num_envs, num_dof, torch.device(device)
        self.actors_per_env = actors_per_env
        f = dict(dtype=torch.float32, device=self.device)
        # net contact forces [N*bodies, 3] (anymal.py:113,124), only for tasks that read them
        self.num_bodies = num_bodies
        self.net_contact_force = torch.zeros((num_envs * num_bodies, 3), **f) if num_bodies else None
        self.state_slab = None
        self.actions_view = None
        if slab:
            # one allocation for everything a host-resident simulator uploads per step (host_io.HostBridge):
            # root | dof_state | dof_force | sensors | actions, each region 256-byte aligned -> a single H2D copy
            sizes = [num_envs * 13, num_envs * num_dof * 2, num_envs * num_dof, num_envs * num_sensor_floats,
                     num_envs * num_dof]
            offs, total = [], 0
            for sz in sizes:
                offs.append(total)
                total += (sz + 63) // 64 * 64
            self.state_slab = torch.zeros(total, **f)
            cut = lambda i, shape: self.state_slab[offs[i]:offs[i] + sizes[i]].view(shape)
            self.root_states = cut(0, (num_envs, 13))
            self.dof_state = cut(1, (num_envs * num_dof, 2))
            self.dof_force = cut(2, (num_envs, num_dof))
            self.force_sensor = cut(3, (num_envs, num_sensor_floats))
            self.actions_view = cut(4, (num_envs, num_dof))
            self.slab_layout = dict(zip(("root_states", "dof_state", "dof_force", "sensors", "actions"), zip(offs, sizes)))
        else:
            # actors_per_env > 1 (MA_OP3.py:78-98: two robots + a table): [N * actors, 13], num_dof = DOFs per env
            self.root_states = torch.zeros((num_envs * actors_per_env, 13), **f)
            self.dof_state = torch.zeros((num_envs * num_dof, 2), **f)
            self.dof_force = torch.zeros((num_envs, num_dof), **f)
            self.force_sensor = torch.zeros((num_envs, num_sensor_floats), **f)
        self.root_states[:, 6] = 1.0                     # identity quaternion (xyzw)
        self.dof_names = dof_names or [f"dof_{i}" for i in range(num_dof)]
        self.asset = {k: v.to(self.device) for k, v in (asset or {}).items()}
        self.physics = physics
        self.actuation_force = None
        self.dof_position_target = None
        self.dof_velocity_target = None
        self.dt = None
        self.frame_count = 0

    # -- tensor API -------------------------------------------------------------------------
    def acquire_actor_root_state_tensor(self): return self.root_states
    def acquire_dof_state_tensor(self): return self.dof_state
    def acquire_dof_force_tensor(self): return self.dof_force
    def acquire_force_sensor_tensor(self): return self.force_sensor
    def acquire_net_contact_force_tensor(self): return self.net_contact_force
    def refresh_net_contact_force_tensor(self): return True
    def refresh_actor_root_state_tensor(self): return True
    def refresh_dof_state_tensor(self): return True
    def refresh_dof_force_tensor(self): return True
    def refresh_force_sensor_tensor(self): return True

    def set_actor_root_state_tensor_indexed(self, src: torch.Tensor, ids: torch.Tensor) -> bool:
        idx = ids.long()
        self.root_states[idx] = src[idx]
        return True

    def set_dof_state_tensor_indexed(self, src: torch.Tensor, ids: torch.Tensor) -> bool:
        if src.data_ptr() != self.dof_state.data_ptr():
            idx = ids.long()
            d = self.num_dof
            self.dof_state.view(self.num_envs, d, 2)[idx] = src.view(self.num_envs, d, 2)[idx]
        return True

    def set_dof_actuation_force_tensor(self, t): self.actuation_force = t; return True
    def set_dof_position_target_tensor(self, t): self.dof_position_target = t; return True
    def set_dof_velocity_target_tensor(self, t): self.dof_velocity_target = t; return True

    def simulate(self):
        self.frame_count += 1                            # gym.get_frame_count (vec_task.py:553): one per sim frame
        if self.physics is not None:
            self.physics(self)

    def get_frame_count(self): return self.frame_count

    def fetch_results(self, wait=True): return True

    # -- helpers for tests / benches --------------------------------------------------------
    def load_state(self, state: Dict[str, torch.Tensor]):
        """Copy a synthetic snapshot (thormang_isaacgym_b200.synthetic) into the sim buffers."""
        alias = {"root_states": "root_states", "root_tensor": "root_states", "dof_state": "dof_state",
                 "state_dof": "dof_state", "dof_force": "dof_force", "sensors": "force_sensor",
                 "torques": "dof_force", "contact_forces": "net_contact_force"}
        for key, attr in alias.items():
            if key in state:
                dst = getattr(self, attr)
                dst.copy_(state[key].to(self.device).reshape(dst.shape))
