"""The reset hand-off to a simulator that keeps its own state.

The reference's reset_idx gives the int32 id list to gym.set_actor_root_state_tensor_indexed and
gym.set_dof_state_tensor_indexed (humanoid.py:264-271, gogoro_new.py:547,552) and re-reads the state with refresh_*
(humanoid.py:240-243, gogoro_new.py:425-426).  FakeSim shares memory with the task and cannot notice when either is
missing; sim/detached_sim.py does not.  Here every task runs twice from one snapshot -- on FakeSim and on DetachedSim,
same physics callback, same in-kernel noise -- and the two runs must stay bit-identical, the detached simulator's PRIVATE
state included; the id lists it receives must be reset_buf.nonzero() as int32.
"""
import pytest
import torch

import parity_util as U
from thormang_isaacgym_b200 import _native as nat
from thormang_isaacgym_b200.sim.detached_sim import DetachedSim

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("n", [1, 31, 1024, 1025, 4103, 131072 + 5, 1 << 20])
@pytest.mark.parametrize("density", [0.0, 0.03, 1.0])
def test_reset_ids_is_nonzero_as_int32(n, density):
    lib = nat.lib()
    g = torch.Generator().manual_seed(n % 9973)
    reset = (torch.rand(n, generator=g) < density).long()
    if density == 0.03:
        reset[-1] = 1
        reset *= torch.randint(1, 5, (n,), generator=g)          # any non-zero value counts, like nonzero()
    reset = reset.to(DEV)
    ids = torch.full((n,), -7, device=DEV, dtype=torch.int32)
    count = torch.full((1,), -7, device=DEV, dtype=torch.int32)
    scratch = torch.empty(max(1, lib.tk_reset_ids_scratch_bytes(n) // 4), device=DEV, dtype=torch.int32)
    assert lib.tk_reset_ids_scratch_bytes(n) == 4 * ((n + 1023) // 1024)
    stream = nat.current_stream_ptr(torch.device(DEV))
    for _ in range(2):                                            # the scratch is reusable as it is
        nat.check(lib.tk_reset_ids(nat.ptr(reset), n, nat.ptr(ids), nat.ptr(count), nat.ptr(scratch), stream), "tk_reset_ids")
        want = reset.nonzero(as_tuple=False).flatten().to(torch.int32)
        k = int(count.item())
        assert k == want.numel()
        assert torch.equal(ids[:k], want)
        assert bool((ids[k:] == -7).all()), "wrote past the list"


def _physics(root, dof_state, forces, dt=0.0166):
    root[:, 0:3] += dt * root[:, 7:10]
    root[:, 2] -= 0.02
    dof_state[:, 0] += dt * dof_state[:, 1]
    if forces is not None:
        dof_state[:, 1] += 0.001 * forces.reshape(-1)


def _humanoid_twins(n, seed):
    env_a, _, st = U.make_humanoid_pair(n, seed=seed)
    env_d, _, _ = U.make_humanoid_pair(n, seed=seed, state={k: v.clone() for k, v in st.items()}, sim_cls=DetachedSim)
    env_a.sim.physics = lambda s: _physics(s.root_states, s.dof_state, env_a.forces)
    env_d.sim.physics = lambda s: _physics(s.root_states, s.dof_state, env_d.forces)
    return env_a, env_d


def test_humanoid_on_a_detached_simulator_equals_the_aliasing_run():
    n, horizon = 1500, 10
    env_a, env_d = _humanoid_twins(n, seed=31)
    assert not env_a._sim_detached() and env_d._sim_detached()
    g = torch.Generator().manual_seed(3)
    total_resets = 0
    for t in range(horizon):
        a = (torch.rand(n, 21, generator=g) * 2.4 - 1.2).to(DEV)
        want_ids = env_a.reset_buf.nonzero(as_tuple=False).flatten().to(torch.int32)
        for env in (env_a, env_d):
            env._clip_in_kernel = True
            env.pre_physics_step(a)
            env._clip_in_kernel = False
            env.sim.simulate()
            env.post_physics_step()
        torch.cuda.synchronize()
        for name in ("obs_buf", "rew_buf", "reset_buf", "progress_buf", "timeout_buf", "potentials", "prev_potentials",
                     "root_states", "dof_state"):
            U.exact(getattr(env_d, name), getattr(env_a, name), f"t={t} {name}")
        U.exact(env_d.sim.private.root_states, env_a.sim.root_states, f"t={t} simulator-side root state")
        U.exact(env_d.sim.private.dof_state, env_a.sim.dof_state, f"t={t} simulator-side dof state")
        if want_ids.numel():
            total_resets += want_ids.numel()
            for what in ("root", "dof"):
                got = env_d.sim.pushed_ids[what]
                assert got.dtype == torch.int32 and torch.equal(got, want_ids), f"t={t}: ids handed to the {what} setter"
    c = env_d.sim.calls
    assert total_resets > n // 20, "the rollout must contain resets"
    assert c["refresh_root_states"] == horizon and c["refresh_dof_state"] == horizon and c["refresh_force_sensor"] == horizon
    assert c["set_actor_root_state_tensor_indexed"] == c["set_dof_state_tensor_indexed"] > 0


def test_the_detached_double_notices_a_missing_push():
    """Sensitivity of the double: with the push switched off the simulator keeps stepping the un-reset bodies."""
    n = 600
    env_a, env_d = _humanoid_twins(n, seed=32)
    env_d._push_resets = lambda ids: None
    a = torch.zeros(n, 21, device=DEV)
    assert int(env_a.reset_buf.sum()) > 0
    for env in (env_a, env_d):
        env.pre_physics_step(a)
        env.sim.simulate()
        env.post_physics_step()
    torch.cuda.synchronize()
    assert not torch.equal(env_d.sim.private.dof_state, env_a.sim.dof_state)


def test_humanoid_reset_idx_method_pushes_its_rows():
    n = 256
    _, env_d = _humanoid_twins(n, seed=33)
    ids = torch.tensor([3, 17, 200], device=DEV)
    env_d.reset_idx(ids)
    torch.cuda.synchronize()
    assert torch.equal(env_d.sim.pushed_ids["root"], ids.to(torch.int32))
    U.exact(env_d.sim.private.root_states[ids], env_d.initial_root_states[ids], "pushed root rows")
    U.exact(env_d.sim.private.dof_state.view(n, 21, 2)[ids], env_d.dof_state.view(n, 21, 2)[ids], "pushed dof rows")


def test_gogoro_on_a_detached_simulator_equals_the_aliasing_run():
    n, horizon = 1100, 8
    env_a, _, st = U.make_gogoro_pair(n, seed=41)
    env_d, _, _ = U.make_gogoro_pair(n, seed=41, state={k: v.clone() for k, v in st.items()}, sim_cls=DetachedSim)
    for env in (env_a, env_d):
        env.sim.physics = lambda s: _physics(s.root_states, s.dof_state, None)
    g = torch.Generator().manual_seed(4)
    total_resets = 0
    for t in range(horizon):
        a = (torch.rand(n, 1, generator=g) * 2 - 1).to(DEV)
        want_ids = env_a.reset_buf.nonzero(as_tuple=False).flatten().to(torch.int32)
        for env in (env_a, env_d):
            env.pre_physics_step(a)
            env.sim.simulate()
            env.post_physics_step()
        torch.cuda.synchronize()
        for name in ("obs_buf", "buffer_obs", "rew_buf", "reset_buf", "progress_buf", "root_tensor", "state_dof", "yaw_command",
                     "curent_speed", "config_vector"):
            U.exact(getattr(env_d, name), getattr(env_a, name), f"t={t} {name}")
        U.exact(env_d.sim.private.root_states, env_a.sim.root_states, f"t={t} simulator-side root state")
        U.exact(env_d.sim.private.dof_state, env_a.sim.dof_state, f"t={t} simulator-side dof state")
        if want_ids.numel():
            total_resets += want_ids.numel()
            assert torch.equal(env_d.sim.pushed_ids["root"], want_ids) and torch.equal(env_d.sim.pushed_ids["dof"], want_ids)
    assert total_resets > 0
    assert env_d.sim.calls["refresh_root_states"] == horizon


def _twin_rollout(env_a, env_d, act_dim, horizon, names, root_name, dof_name, seed):
    g = torch.Generator().manual_seed(seed)
    n, total = env_a.num_envs, 0
    for env in (env_a, env_d):
        env.sim.physics = lambda s: _physics(s.root_states, s.dof_state, None)
    for t in range(horizon):
        a = (torch.rand(n, act_dim, generator=g) * 2 - 1).to(DEV)
        want_ids = env_a.reset_buf.nonzero(as_tuple=False).flatten().to(torch.int32)
        for env in (env_a, env_d):
            env.pre_physics_step(a)
            env.sim.simulate()
            env.post_physics_step()
        torch.cuda.synchronize()
        for name in names + (root_name, dof_name):
            U.exact(getattr(env_d, name), getattr(env_a, name), f"t={t} {name}")
        U.exact(env_d.sim.private.root_states, env_a.sim.root_states, f"t={t} simulator-side root state")
        U.exact(env_d.sim.private.dof_state, env_a.sim.dof_state, f"t={t} simulator-side dof state")
        if want_ids.numel():
            total += want_ids.numel()
            assert torch.equal(env_d.sim.pushed_ids["root"], want_ids) and torch.equal(env_d.sim.pushed_ids["dof"], want_ids)
    assert total > 0, "the rollout must contain resets"
    assert env_d.sim.calls["refresh_root_states"] == horizon and env_d.sim.calls["refresh_dof_state"] == horizon


def test_anymal_on_a_detached_simulator_equals_the_aliasing_run():
    n = 900
    env_a, _, st = U.make_anymal_pair(n, seed=51)
    env_d, _, _ = U.make_anymal_pair(n, seed=51, state={k: v.clone() for k, v in st.items()}, sim_cls=DetachedSim)
    _twin_rollout(env_a, env_d, 12, 6, ("obs_buf", "rew_buf", "reset_buf", "progress_buf", "commands"), "root_states", "dof_state", 6)
    assert env_d.sim.calls["refresh_net_contact_force"] == 6


@pytest.mark.parametrize("turning", [False, True], ids=["paper", "turning"])
def test_gogoro_8x20_on_a_detached_simulator_equals_the_aliasing_run(turning):
    n = 700
    env_a, _, st = U.make_gogoro_paper_pair(n, seed=61, turning=turning)
    env_d, _, _ = U.make_gogoro_paper_pair(n, seed=61, state={k: v.clone() for k, v in st.items()}, turning=turning,
                                           sim_cls=DetachedSim)
    _twin_rollout(env_a, env_d, 1, 6, ("obs_buf", "rew_buf", "reset_buf", "progress_buf", "command_ring", "yaw_command"),
                  "root_tensor", "state_dof", 7)
