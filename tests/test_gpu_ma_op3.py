"""GPU parity of MA_OP3 (two OP3 humanoids + table, ``[N, agents, ...]`` tensors) against the CPU
oracle (oracle/ma_op3_oracle.py, pinned bit-for-bit to MA_OP3.py's JIT functions).

Bar: reset / progress / timeout / last_contacts masks and the gripper counts bit-exact; reset-state writes exact; fp32
observations, potentials, air times and rewards within 1e-5 relative (tests/parity_util.py).
"""
import os

import numpy as np
import pytest
import torch

import parity_util as U
from thormang_isaacgym_b200 import synthetic as syn

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _reward_close(got, want, what):
    # rewards are sums of O(1) terms (alive 2, foot-force excess up to tens) minus small penalties and a progress term
    # (potential differences of ~600-magnitude values, ulp 6e-5, times 5): the absolute floor is that cancellation
    U.close(got, want, rtol=U.RTOL, atol=2e-5, what=what)


def _check_obs_side(env, orc, tag):
    U.close(env.obs_buf, orc.obs_buf, what=f"{tag} obs")
    U.close(env.potentials, orc.potentials, what=f"{tag} potentials")
    U.close(env.prev_potentials, orc.prev_potentials, what=f"{tag} prev_potentials")
    U.close(env.prev_table_potentials, orc.prev_table_potentials, what=f"{tag} prev_table_potentials")


def _check_reward_side(env, orc, tag):
    _reward_close(env.rew_buf, orc.rew_buf, f"{tag} rew")
    U.exact(env.reset_buf, orc.reset_buf, f"{tag} reset")
    U.exact(env.last_contacts, orc.last_contacts, f"{tag} last_contacts")
    U.close(env.feet_air_time, orc.feet_air_time, what=f"{tag} feet_air_time")


@pytest.mark.parametrize("n", [1, 31, 32, 33, 1000, 4096])
def test_observations_and_reward_split(n):
    env, orc, st = U.make_ma_op3_pair(n, seed=100 + n)
    env.compute_observations()
    env.compute_reward()
    orc.compute_observations()
    orc.compute_reward()
    torch.cuda.synchronize()
    _check_obs_side(env, orc, f"n={n}")
    _check_reward_side(env, orc, f"n={n}")


def test_golden_vectors(golden_dir):
    from oracle import ma_op3_oracle as mo
    z = np.load(os.path.join(golden_dir, "ma_op3_jit.npz"))
    g = {k: torch.from_numpy(z[k]) for k in z.files}
    n = g["progress_buf"].shape[0]
    st = {k: v for k, v in g.items() if not k.startswith("out_") and k != "grip_counts"}
    env, orc, _ = U.make_ma_op3_pair(n, seed=1, state=st)
    env.compute_observations()
    env.compute_reward()
    torch.cuda.synchronize()
    U.exact(env.grip_count_buf, g["grip_counts"], "golden grip counts (agent 0 holds, agent 1 does not)")
    U.close(env.obs_buf, g["out_obs"], what="golden obs")
    U.close(env.potentials, g["out_potentials"], what="golden potentials")
    _reward_close(env.rew_buf, g["out_rew"], "golden reward")
    U.exact(env.reset_buf, g["out_reset"], "golden reset (0.25 m / 0.40 m / 0.90 thresholds, episode_lengths == 0 line)")
    U.exact(env.last_contacts, g["out_last_contacts"], "golden last_contacts")
    U.close(env.feet_air_time, g["out_feet_air_time"], what="golden feet_air_time")


@pytest.mark.parametrize("clip", [False, True])
def test_pre_physics(clip):
    n = 1000
    env, orc, st = U.make_ma_op3_pair(n, seed=7)
    env._clip_in_kernel = clip
    env.pre_physics_step(st["actions"].to(DEV))
    env._clip_in_kernel = False
    orc.pre_physics_step(st["actions"], clip=clip)
    torch.cuda.synchronize()
    U.exact(env.actions, orc.actions, "actions")
    U.exact(env.pd_targets, orc.pd_targets, "PD targets")
    assert env.sim.dof_position_target.data_ptr() == env.pd_targets.data_ptr()


def test_reset_idx():
    n = 300
    env, orc, st = U.make_ma_op3_pair(n, seed=9)
    ids = torch.tensor([0, 5, 31, 32, 33, 299])
    uc = syn.ma_op3_reset_noise(n, seed=4)[:len(ids)]
    env.reset_idx(ids.to(DEV), uc.to(DEV))
    orc.reset_idx(ids, uc)
    torch.cuda.synchronize()
    for name in ("root_states", "dof_state", "commands", "progress_buf", "reset_buf", "potentials", "prev_potentials",
                 "table_potentials", "prev_table_potentials", "last_actions", "feet_air_time"):
        U.exact(getattr(env, name), getattr(orc, name), name)


@pytest.mark.parametrize("n", [1, 33, 1000, 4096, 65536])
def test_fused_post_physics(n):
    env, orc, st = U.make_ma_op3_pair(n, seed=500 + n)
    uc = syn.ma_op3_reset_noise(n, seed=n)
    ids = orc.reset_buf.nonzero(as_tuple=False).squeeze(-1)
    env.post_physics_step(uc.to(DEV))
    got_ids = orc.post_physics_step(uc[ids])
    torch.cuda.synchronize()
    assert torch.equal(got_ids, ids)
    if n >= 1000:
        assert len(ids) > 0 and int(orc.reset_buf.sum()) > 0
    U.exact(env.reset_buf, orc.reset_buf, "reset_buf")
    U.exact(env.progress_buf, orc.progress_buf, "progress_buf")
    U.exact(env.timeout_buf, orc.timeout_buf.bool(), "timeout_buf")
    U.exact(env.root_states, orc.root_states, "root_states (reset writes)")
    U.exact(env.dof_state, orc.dof_state, "dof_state (reset writes)")
    U.exact(env.commands, orc.commands, "commands")
    U.exact(env.table_potentials, orc.table_potentials, "table_potentials")
    U.exact(env.prev_torques, orc.prev_torques, "prev_torques (end-of-step refresh)")
    U.exact(env.last_actions, orc.last_actions, "last_actions (end-of-step refresh)")
    _check_obs_side(env, orc, "fused")
    _check_reward_side(env, orc, "fused")


def _one_holder_state(n, seed):
    """Exactly one env presses each gripper of agent 0 (hold term = 1); agent 1 has two left presses (0)."""
    st = syn.ma_op3_state(n, seed=seed)
    _, grip = syn.ma_op3_body_indices()
    cf = st["contact_forces"]
    cf[:, grip.flatten(), 0] = 0.0
    cf[n // 3, grip[0, 0], 0] = 0.5
    cf[n // 2, grip[0, 1], 0] = 0.7
    cf[1, grip[1, 0], 0] = 0.5
    cf[n - 1, grip[1, 0], 0] = 0.5
    cf[2, grip[1, 1], 0] = 0.9
    return st


@pytest.mark.parametrize("n", [40, 4096])
def test_fused_hold_term(n):
    """_reward_gripper_contact_hold (:346-358) is a batch-global 0/1: counted inside the fused kernel, added by finalize."""
    st = _one_holder_state(n, seed=77)
    env, orc, _ = U.make_ma_op3_pair(n, seed=77, state=st)
    uc = syn.ma_op3_reset_noise(n, seed=5)
    ids = orc.reset_buf.nonzero(as_tuple=False).squeeze(-1)
    env.post_physics_step(uc.to(DEV))
    orc.post_physics_step(uc[ids])
    torch.cuda.synchronize()
    assert env.grip_count_buf.tolist() == [[1, 1], [2, 1]]
    _check_reward_side(env, orc, "hold")
    s = env.episode_statistics().cpu()
    assert abs(float(s[0]) - float(env.rew_buf.double().sum())) < 1e-3 * n
    assert abs(float(env.episode_return.double().sum()) - float(env.rew_buf.double().sum())) < 1e-3 * n


def test_philox_mode_and_rollout():
    n = 2048
    env, orc, st = U.make_ma_op3_pair(n, seed=21)
    g = torch.Generator().manual_seed(3)
    for step in range(6):
        a = torch.rand((n, 2, 22), generator=g) * 3 - 1.5
        env._clip_in_kernel = True
        env.pre_physics_step(a.to(DEV))
        env._clip_in_kernel = False
        orc.pre_physics_step(a, clip=True)
        uc = env.philox_noise().cpu()
        ids = orc.reset_buf.nonzero(as_tuple=False).squeeze(-1)
        env.post_physics_step()                        # in-kernel Philox
        orc.post_physics_step(uc[ids])
        torch.cuda.synchronize()
        U.exact(env.reset_buf, orc.reset_buf, f"step {step} reset_buf")
        U.exact(env.progress_buf, orc.progress_buf, f"step {step} progress_buf")
        U.exact(env.dof_state, orc.dof_state, f"step {step} dof_state")
        U.exact(env.commands, orc.commands, f"step {step} commands")
        U.exact(env.pd_targets, orc.pd_targets, f"step {step} targets")
        _check_obs_side(env, orc, f"step {step}")
        _check_reward_side(env, orc, f"step {step}")


def test_step_api_and_statistics():
    import thormang_isaacgym_b200 as pkg
    env = pkg.make(seed=3, task="MA_OP3", num_envs=128, sim_device=DEV, rl_device=DEV)
    obs_d, rew, reset, extras = env.step(env.zero_actions())
    assert obs_d["obs"].shape == (128, 2, 88) and rew.shape == (128, 2) and reset.dtype == torch.int64
    assert float(obs_d["obs"].abs().max()) <= 5.0                         # clipObservations, MA_OP3.yaml
    assert extras["time_outs"].dtype == torch.bool
    s = env.episode_statistics().cpu()
    assert int(s[7]) == 128 and abs(float(s[0]) - float(env.rew_buf.double().sum())) < 1e-4


def test_unaligned_and_sharded():
    """Two shards (cut not a multiple of 32) with the gripper counts summed across them == the unsharded run."""
    n = 4096 + 17
    st = _one_holder_state(n, seed=33)
    uc = syn.ma_op3_reset_noise(n, seed=34)
    full, _, _ = U.make_ma_op3_pair(n, seed=33, state=st)
    full.post_physics_step(uc.to(DEV))
    cut = 1000
    parts = []
    for sl in (slice(0, cut), slice(cut, n)):
        m = sl.stop - sl.start
        sub = {}
        for k, v in st.items():
            per = v.shape[0] // n
            sub[k] = v[sl.start * per:sl.stop * per].clone()
        sub["initial_root_states"] = full.initial_root_states[sl].cpu().clone()
        env, _, _ = U.make_ma_op3_pair(m, seed=33, state=sub)
        env.launch_post_physics(uc[sl].to(DEV))
        parts.append(env)
    total = parts[0].grip_count_buf + parts[1].grip_count_buf      # what the NCCL all-reduce does (gripAllReduce)
    for env in parts:
        env.grip_count_buf.copy_(total)
        env.finalize_reward()
    torch.cuda.synchronize()
    assert total.tolist() == [[1, 1], [2, 1]] and full.grip_count_buf.tolist() == [[1, 1], [2, 1]]
    for name in ("obs_buf", "rew_buf", "reset_buf", "progress_buf", "commands", "feet_air_time", "potentials"):
        U.exact(torch.cat([getattr(p, name) for p in parts]), getattr(full, name), "sharded " + name)


@pytest.mark.parametrize("feet", [[[10, 11, 16], [35, 36, 41]], [[10, 11, 16, 17, 3, 4], [35, 36, 41, 42, 28, 29]]])
def test_runtime_foot_count(feet):
    """The kernel is specialised for the OP3's four "*ank*" links; any other count takes the run-time path."""
    n = 1000
    env, orc, st = U.make_ma_op3_pair(n, seed=61, feet=feet)
    assert env.feet_air_time.shape == (n, 2, len(feet[0]))
    uc = syn.ma_op3_reset_noise(n, seed=8)
    ids = orc.reset_buf.nonzero(as_tuple=False).squeeze(-1)
    env.post_physics_step(uc.to(DEV))
    orc.post_physics_step(uc[ids])
    torch.cuda.synchronize()
    _check_obs_side(env, orc, "nf")
    _check_reward_side(env, orc, "nf")
    U.exact(env.progress_buf, orc.progress_buf, "progress_buf")


def _shifted(t, elems):
    """The same data at an address `elems` elements past a 512-byte aligned allocation."""
    buf = torch.empty(t.numel() + 64, dtype=t.dtype, device=t.device)
    v = buf[elems:elems + t.numel()].view(t.shape)
    v.copy_(t)
    return v


@pytest.mark.parametrize("shift", [
    {"obs_buf": 4, "contact_forces": 2, "feet_air_time": 1},        # 16-byte obs rows, 8-byte contact rows, scalar feet path
    {"contact_forces": 1, "last_contacts": 1},                       # 4-byte contact rows
    {"root_states": 1, "torques": 2, "contact_forces": 4},          # no bulk tile copies at all; contact block still 16-byte aligned
], ids=["obs16_contact8_feet4", "contact4_lastc1", "no_bulk"])
def test_alignment_fallbacks(shift):
    """Every vector / bulk-copy path has a fallback chosen from the bound addresses: same results from shifted buffers."""
    n = 1000 + 7
    env, orc, st = U.make_ma_op3_pair(n, seed=71)
    for name, k in shift.items():
        setattr(env, name, _shifted(getattr(env, name), k))
    env.rebind()
    uc = syn.ma_op3_reset_noise(n, seed=9)
    for step in range(2):
        ids = orc.reset_buf.nonzero(as_tuple=False).squeeze(-1)
        env.post_physics_step(uc.to(DEV))
        orc.post_physics_step(uc[ids])
        torch.cuda.synchronize()
        _check_obs_side(env, orc, f"shifted step {step}")
        _check_reward_side(env, orc, f"shifted step {step}")
        U.exact(env.progress_buf, orc.progress_buf, "progress_buf")
        U.exact(env.feet_air_time, orc.feet_air_time, "feet_air_time")
