"""A simulator double that does NOT share memory with the task -- the hand-off the real PhysX tensor API has.

With ``gymtorch.wrap_tensor`` the task holds staging tensors: PhysX writes them on ``refresh_*`` and reads rows of
them only when ``set_actor_root_state_tensor_indexed`` / ``set_dof_state_tensor_indexed`` are called with an int32 id
list (isaacgymenvs/tasks/humanoid.py:240-243,264-271; gogoro_new.py:410-411,547,552).  ``FakeSim`` cannot show a
missing refresh or a missing push, because there the task's tensors ARE the simulator state.  Here the simulator
state is private:

* ``acquire_*``     hand out the staging tensors (same shapes as FakeSim);
* ``refresh_*``     overwrite the staging tensor with the private state (whatever the task wrote there is lost);
* ``set_*_indexed`` copy rows ``ids`` of ``src`` into the private state -- the only way task writes reach it;
* ``simulate()``    runs the ``physics`` callback on the PRIVATE state.

``calls`` counts every API call; ``pushed_ids`` keeps the id lists of the last push, for tests.
"""
from collections import Counter
from types import SimpleNamespace

import torch

from .fake_sim import FakeSim


class DetachedSim(FakeSim):
    aliases_task_memory = False

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.calls = Counter()
        self.pushed_ids = {}
        self._px = SimpleNamespace(num_envs=self.num_envs, num_dof=self.num_dof)
        self._sync_private()

    _STATE = ("root_states", "dof_state", "dof_force", "force_sensor", "net_contact_force")

    def _sync_private(self):
        for name in self._STATE:
            t = getattr(self, name)
            setattr(self._px, name, None if t is None else t.clone())

    def load_state(self, state):
        super().load_state(state)
        self._sync_private()

    @property
    def private(self):
        """The simulator's own state (what PhysX would hold)."""
        return self._px

    def _refresh(self, name):
        self.calls["refresh_" + name] += 1
        getattr(self, name).copy_(getattr(self._px, name))
        return True

    def refresh_actor_root_state_tensor(self): return self._refresh("root_states")
    def refresh_dof_state_tensor(self): return self._refresh("dof_state")
    def refresh_dof_force_tensor(self): return self._refresh("dof_force")
    def refresh_force_sensor_tensor(self): return self._refresh("force_sensor")
    def refresh_net_contact_force_tensor(self): return self._refresh("net_contact_force")

    def _ids(self, ids, what):
        if ids.dtype != torch.int32:
            raise TypeError(f"{what}: the gym API takes int32 actor indices (humanoid.py:264), got {ids.dtype}")
        self.pushed_ids[what] = ids.clone()
        return ids.long()

    def set_actor_root_state_tensor_indexed(self, src, ids):
        self.calls["set_actor_root_state_tensor_indexed"] += 1
        idx = self._ids(ids, "root")
        if tuple(src.shape) != tuple(self._px.root_states.shape):
            raise ValueError(f"root state source has shape {tuple(src.shape)}, the simulator's is {tuple(self._px.root_states.shape)}")
        self._px.root_states[idx] = src[idx]
        return True

    def set_dof_state_tensor_indexed(self, src, ids):
        self.calls["set_dof_state_tensor_indexed"] += 1
        idx = self._ids(ids, "dof")
        d = self.num_dof
        if src.numel() != self._px.dof_state.numel():
            raise ValueError(f"dof state source has {src.numel()} elements, the simulator's {self._px.dof_state.numel()}")
        self._px.dof_state.view(self.num_envs, d, 2)[idx] = src.reshape(self.num_envs, d, 2)[idx]
        return True

    def simulate(self):
        self.calls["simulate"] += 1
        self.frame_count += 1
        if self.physics is not None:
            self._px.actuation_force = self.actuation_force
            self._px.dof_position_target = self.dof_position_target
            self._px.dof_velocity_target = self.dof_velocity_target
            self.physics(self._px)
