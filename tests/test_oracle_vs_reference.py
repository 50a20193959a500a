"""Oracle vs the reference's own TorchScript functions, lifted live from /root/reference.

Runs only where the reference tree exists (the build container); on the GPU box the
same check is carried by the golden fixtures (test_oracle_golden.py).  Same torch
operators on the same CPU => bit-exact equality is required.
"""
import numpy as np
import pytest
import torch

from oracle import gogoro_oracle as go
from oracle import humanoid_oracle as ho
from oracle import lift_reference as L
from thormang_isaacgym_b200 import synthetic as syn

pytestmark = pytest.mark.skipif(not L.reference_available(), reason="/root/reference not present")


@pytest.mark.parametrize("n", [1, 31, 32, 33, 4096])
def test_humanoid_bit_exact(n):
    ref = L.humanoid()
    asset = syn.humanoid_asset()
    st = syn.humanoid_state(n, seed=1234 + n, asset=asset)
    d = 21
    dof_pos = st["dof_state"].view(n, d, 2)[..., 0]
    dof_vel = st["dof_state"].view(n, d, 2)[..., 1]
    targets = torch.tensor([1000.0, 0.0, 0.0]).repeat(n, 1)
    inv = torch.tensor([0.0, 0.0, 0.0, 1.0]).repeat(n, 1)
    b0 = torch.tensor([1.0, 0.0, 0.0]).repeat(n, 1)
    b1 = torch.tensor([0.0, 0.0, 1.0]).repeat(n, 1)
    lo, hi, eff = asset["dof_limits_lower"], asset["dof_limits_upper"], asset["motor_efforts"]
    want = ref.compute_humanoid_observations(torch.zeros(n, 110), st["root_states"], targets, st["potentials"], inv,
                                             dof_pos, dof_vel, st["dof_force"], lo, hi, 0.1, st["sensors"],
                                             st["actions"], 0.0166, 0.01, 0.25, b0, b1)
    got = ho.humanoid_observations(st["root_states"], targets, st["potentials"], inv, dof_pos, dof_vel,
                                   st["dof_force"], lo, hi, 0.1, st["sensors"], st["actions"], 0.0166, 0.01, 0.25,
                                   b0, b1)
    for w, g_ in zip(want, got):
        assert torch.equal(w, g_)
    rw = ref.compute_humanoid_reward(want[0], st["reset_buf"], st["progress_buf"], st["actions"], 0.1, 0.5, want[1],
                                     want[2], 0.01, 0.05, 0.25, float(eff.max()), eff, 0.8, -1.0, 1000.0)
    rg = ho.humanoid_reward(got[0], st["reset_buf"], st["progress_buf"], st["actions"], 0.1, 0.5, got[1], got[2],
                            0.01, 0.05, 0.25, float(eff.max()), eff, 0.8, -1.0, 1000.0)
    assert torch.equal(rw[0], rg[0]) and torch.equal(rw[1], rg[1])


@pytest.mark.parametrize("n", [1, 33, 4096])
def test_gogoro_bit_exact(n):
    ref = L.gogoro()
    st = syn.gogoro_state(n, seed=77 + n)
    want = ref.compute_gogoro_observations(st["root_tensor"].clone(), st["yaw_command"].clone(),
                                           st["curent_command"].clone())
    got = go.gogoro_observations(st["root_tensor"].clone(), st["yaw_command"].clone(), st["curent_command"].clone())
    assert torch.equal(want, got)
    rw = ref.compute_gogoro_reward(curent_obs=want.view(n, 1, 6), progress_buf=st["progress_buf"],
                                   action_history=st["action_history"], max_episode_length=torch.tensor(1000))
    rg = go.gogoro_reward(got.view(n, 1, 6), st["progress_buf"], st["action_history"], 1000)
    assert torch.equal(rw[0], rg[0]) and torch.equal(rw[1], rg[1])


@pytest.mark.parametrize("n", [1, 33, 4096])
def test_ant_bit_exact(n):
    from oracle import ant_oracle as ao
    ref = L.ant()
    asset = syn.ant_asset()
    st = syn.ant_state(n, seed=500 + n, asset=asset)
    d = 8
    dof_pos = st["dof_state"].view(n, d, 2)[..., 0]
    dof_vel = st["dof_state"].view(n, d, 2)[..., 1]
    targets = torch.tensor([1000.0, 0.0, 0.0]).repeat(n, 1)
    inv = torch.tensor([0.0, 0.0, 0.0, 1.0]).repeat(n, 1)
    b0 = torch.tensor([1.0, 0.0, 0.0]).repeat(n, 1)
    b1 = torch.tensor([0.0, 0.0, 1.0]).repeat(n, 1)
    lo, hi = asset["dof_limits_lower"], asset["dof_limits_upper"]
    want = ref.compute_ant_observations(torch.zeros(n, 60), st["root_states"], targets, st["potentials"], inv, dof_pos,
                                        dof_vel, lo, hi, 0.2, st["sensors"], st["actions"], 0.0166, 0.1, b0, b1, 2)
    got = ao.ant_observations(st["root_states"], targets, st["potentials"], inv, dof_pos, dof_vel, lo, hi, 0.2,
                              st["sensors"], st["actions"], 0.0166, 0.1, b0, b1, 2)
    for w, g_ in zip(want, got):
        assert torch.equal(w, g_)
    rw = ref.compute_ant_reward(want[0], st["reset_buf"], st["progress_buf"], st["actions"], 0.1, 0.5, want[1], want[2],
                                0.005, 0.05, 0.1, 0.31, -2.0, 1000.0)
    rg = ao.ant_reward(got[0], st["reset_buf"], st["progress_buf"], st["actions"], 0.1, 0.5, got[1], got[2], 0.005, 0.05,
                       0.1, 0.31, -2.0, 1000.0)
    assert torch.equal(rw[0], rg[0]) and torch.equal(rw[1], rg[1])


@pytest.mark.parametrize("n", [1, 33, 4096])
def test_gogoro_paper_bit_exact(n):
    """gogoro_realistic_turning_sim_paper.py:713-807."""
    from oracle import gogoro_paper_oracle as gp
    ref = L.gogoro_paper()
    st = syn.gogoro_paper_state(n, seed=177 + n)
    delay = (st["steer_delay"].clone() - 0) / (5 - 0)
    want = ref.compute_gogoro_observations(st["root_tensor"].clone(), st["yaw_command"].clone(),
                                           st["curent_command"].clone(), delay.clone())
    got = gp.paper_observations(st["root_tensor"].clone(), st["yaw_command"].clone(), st["curent_command"].clone(), delay)
    assert torch.equal(want, got)
    frames = torch.cat([st["buffer_obs"][:, 1:], want.unsqueeze(1)], dim=1)
    rw = ref.compute_gogoro_reward(act_buff=frames[:, :, 6], curent_obs=frames, progress_buf=st["progress_buf"],
                                   max_episode_length=torch.tensor(3600), yaw_update_freq=torch.tensor(300))
    rg = gp.paper_reward(frames[:, :, 6], frames, st["progress_buf"], 3600)
    assert torch.equal(rw[0], rg[0]) and torch.equal(rw[1], rg[1])
    a, b = syn._u(syn._gen(3), (256,), -7.0, 7.0), syn._u(syn._gen(4), (256,), -7.0, 7.0)
    assert torch.equal(ref.shortest_angle_distance(a, b), gp.shortest_angle_distance(a, b))


@pytest.mark.parametrize("n", [1, 33, 4096])
def test_gogoro_turning_bit_exact(n):
    """SURVEY 8(f) N1 sibling: gogoro_realistic_turning.py:713-810 (rational reward, all-env mean of the command history)."""
    from oracle import gogoro_turning_oracle as gt
    ref = L.gogoro_turning()
    st = syn.gogoro_paper_state(n, seed=277 + n)
    delay = (st["steer_delay"].clone() - 0) / (5 - 0)
    want = ref.compute_gogoro_observations(st["root_tensor"].clone(), st["yaw_command"].clone(),
                                           st["curent_command"].clone(), delay.clone())
    got = gt.paper_observations(st["root_tensor"].clone(), st["yaw_command"].clone(), st["curent_command"].clone(), delay)
    assert torch.equal(want, got)
    frames = torch.cat([st["buffer_obs"][:, 1:], want.unsqueeze(1)], dim=1)
    frames[:, :, 6] += 0.05
    rw = ref.compute_gogoro_reward(act_buff=frames[:, :, 6], curent_obs=frames, progress_buf=st["progress_buf"],
                                   max_episode_length=torch.tensor(3600), yaw_update_freq=torch.tensor(300))
    rg = gt.turning_reward(frames[:, :, 6], frames, st["progress_buf"], 3600)
    assert torch.equal(rw[0], rg[0]) and torch.equal(rw[1], rg[1])
    a, b = syn._u(syn._gen(5), (256,), -7.0, 7.0), syn._u(syn._gen(6), (256,), -7.0, 7.0)
    assert torch.equal(ref.shortest_angle_distance(a, b), gt.shortest_angle_distance(a, b))


@pytest.mark.parametrize("n", [1, 33, 4096])
def test_anymal_bit_exact(n):
    """anymal.py:323-431."""
    from oracle import anymal_oracle as ao
    ref = L.anymal()
    st = syn.anymal_state(n, seed=700 + n, contact_frac=0.2)
    c = ao.AnymalConsts()
    d = 12
    dof_pos = st["dof_state"].view(n, d, 2)[..., 0]
    dof_vel = st["dof_state"].view(n, d, 2)[..., 1]
    default = torch.tensor([0.03, 0.4, -0.8, 0.03, -0.4, 0.8, -0.03, 0.4, -0.8, -0.03, -0.4, 0.8]).repeat(n, 1)
    grav = torch.tensor([0.0, 0.0, -1.0]).repeat(n, 1)
    want = ref.compute_anymal_observations_V2(st["root_states"], st["commands"], st["torques"], st["contact_forces"], dof_pos,
                                              default, dof_vel, grav, st["actions"], 2.0, 0.25, 1.0, 0.05)
    got = ao.anymal_observations(st["root_states"], st["commands"], st["torques"], st["contact_forces"], dof_pos, default,
                                 dof_vel, grav, 1.0, 0.05)
    assert torch.equal(want, got)
    knee = torch.tensor([2, 5, 8, 11])
    rw = ref.compute_anymal_reward(st["root_states"], st["commands"], st["torques"], st["contact_forces"], knee,
                                   st["progress_buf"], c.rew_scales(), 0, c.max_episode_length)
    rg = ao.anymal_reward(st["root_states"], st["commands"], st["torques"], st["contact_forces"], knee, st["progress_buf"],
                          c.rew_scales(), 0, c.max_episode_length)
    assert torch.equal(rw[0], rg[0]) and torch.equal(rw[1], rg[1])
    if n >= 33:
        b, t = max(1, n // 64), 64 if n >= 64 else n
        m = b * t
        lin, ang = got[:m, 0:3].reshape(b, t, 3), got[:m, 3:6].reshape(b, t, 3)
        tq, cmd = st["torques"][:m].reshape(b, t, 12), st["commands"][:m].reshape(b, t, 3)
        assert torch.equal(ref.compute_reward(lin, ang, tq, cmd, c.rew_scales()), ao.anymal_wm_reward(lin, ang, tq, cmd, c.rew_scales()))


@pytest.mark.parametrize("n", [1, 33, 4096])
def test_ma_op3_bit_exact(n):
    """MA_OP3.py:496-708 through the method bodies of compute_observations / compute_reward."""
    from oracle import ma_op3_oracle as mo
    ref = L.ma_op3()
    st = syn.ma_op3_state(n, seed=900 + n)
    got = mo.run_jit_level(mo.oracle_from_state(st))
    want = mo.run_jit_level(mo.oracle_from_state(st), ref.compute_op3_observations, ref.compute_op3_reward,
                            ref.compute_objective_reward)
    for k in want:
        assert torch.equal(want[k], got[k]), k
