"""GPU parity of Anymal (PD-target write + contact-force termination) against the CPU oracle
(oracle/anymal_oracle.py, pinned bit-for-bit to anymal.py's JIT functions).

Bar: reset / progress / timeout masks bit-exact; reset-state writes and the PD targets exact; fp32 observations and
rewards within 1e-5 relative (tests/parity_util.py).
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
    # rewards are O(1e-2) sums of two exp() terms and a negative torque term clipped at 0: near the clip "relative" has
    # no meaning, so the absolute floor is 1e-5 of the largest term (0.02)
    U.close(got, want, rtol=U.RTOL, atol=2e-7, what=what)


@pytest.mark.parametrize("n", [1, 31, 32, 33, 1000, 4096])
def test_observations_and_reward_split(n):
    env, orc, st = U.make_anymal_pair(n, seed=100 + n)
    env.compute_observations()
    env.compute_reward()
    orc.compute_observations()
    orc.compute_reward()
    torch.cuda.synchronize()
    U.close(env.obs_buf, orc.obs_buf, what=f"anymal obs n={n}")
    _reward_close(env.rew_buf, orc.rew_buf, f"anymal rew n={n}")
    U.exact(env.reset_buf, orc.reset_buf, f"anymal reset n={n}")


def test_golden_vectors(golden_dir):
    z = np.load(os.path.join(golden_dir, "anymal_jit.npz"))
    g = {k: torch.from_numpy(z[k]) for k in z.files}
    n = g["root_states"].shape[0]
    st = {k: g[k] for k in ("root_states", "dof_state", "torques", "contact_forces", "commands", "actions", "progress_buf")}
    st["reset_buf"] = torch.zeros(n, dtype=torch.long)
    env, orc, _ = U.make_anymal_pair(n, seed=1, state=st)
    env.compute_observations()
    env.compute_reward()
    torch.cuda.synchronize()
    U.close(env.obs_buf, g["obs"], what="golden obs")
    _reward_close(env.rew_buf, g["rew"], "golden reward")
    U.exact(env.reset_buf, g["reset"], "golden reset (contact norms at / 1 ulp around 1 N)")


@pytest.mark.parametrize("clip", [False, True])
def test_pre_physics(clip):
    n = 1000
    env, orc, st = U.make_anymal_pair(n, seed=7)
    env._clip_in_kernel = clip
    env.pre_physics_step(st["actions"].to(DEV))
    env._clip_in_kernel = False
    orc.pre_physics_step(st["actions"], clip=clip)
    torch.cuda.synchronize()
    U.exact(env.actions, orc.actions, "actions")
    U.exact(env.targets, orc.targets, "PD targets")
    assert env.sim.dof_position_target.data_ptr() == env.targets.data_ptr()


def test_reset_idx():
    n = 300
    env, orc, st = U.make_anymal_pair(n, seed=9)
    ids = torch.tensor([0, 5, 31, 32, 33, 299])
    up, uv, uc = (t[:len(ids)] for t in syn.anymal_reset_noise(n, seed=4))
    env.reset_idx(ids.to(DEV), up.to(DEV), uv.to(DEV), uc.to(DEV))
    orc.reset_idx(ids, up, uv, uc)
    torch.cuda.synchronize()
    for name in ("root_states", "dof_state", "commands", "progress_buf", "reset_buf"):
        U.exact(getattr(env, name), getattr(orc, name), name)
    assert env.reset_buf[ids.to(DEV)].tolist() == [1] * len(ids)          # anymal.py:310 (sic)


@pytest.mark.parametrize("n", [1, 33, 1000, 4096, 65536])
def test_fused_post_physics(n):
    env, orc, st = U.make_anymal_pair(n, seed=500 + n)
    up, uv, uc = syn.anymal_reset_noise(n, seed=n)
    ids = orc.reset_buf.nonzero(as_tuple=False).squeeze(-1)
    env.post_physics_step(up.to(DEV), uv.to(DEV), uc.to(DEV))
    got_ids = orc.post_physics_step(up[ids], uv[ids], uc[ids])
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
    U.close(env.obs_buf, orc.obs_buf, what="obs_buf")
    _reward_close(env.rew_buf, orc.rew_buf, "rew_buf")


def test_philox_mode_and_rollout():
    n = 2048
    env, orc, st = U.make_anymal_pair(n, seed=21)
    g = torch.Generator().manual_seed(3)
    for step in range(6):
        a = torch.rand((n, 12), generator=g) * 3 - 1.5
        env._clip_in_kernel = True
        env.pre_physics_step(a.to(DEV))
        env._clip_in_kernel = False
        orc.pre_physics_step(a, clip=True)
        up, uv, uc = (t.cpu() for t in env.philox_noise())
        ids = orc.reset_buf.nonzero(as_tuple=False).squeeze(-1)
        env.post_physics_step()                        # in-kernel Philox
        orc.post_physics_step(up[ids], uv[ids], uc[ids])
        torch.cuda.synchronize()
        U.exact(env.reset_buf, orc.reset_buf, f"step {step} reset_buf")
        U.exact(env.progress_buf, orc.progress_buf, f"step {step} progress_buf")
        U.exact(env.dof_state, orc.dof_state, f"step {step} dof_state")
        U.exact(env.commands, orc.commands, f"step {step} commands")
        U.close(env.obs_buf, orc.obs_buf, what=f"step {step} obs")
        U.exact(env.targets, orc.targets, f"step {step} targets")


def test_step_api_and_statistics():
    import thormang_isaacgym_b200 as pkg
    env = pkg.make(seed=3, task="Anymal", num_envs=128, sim_device=DEV, rl_device=DEV)
    obs_d, rew, reset, extras = env.step(env.zero_actions())
    assert obs_d["obs"].shape == (128, 61) and rew.shape == (128,) and reset.dtype == torch.int64
    assert float(obs_d["obs"].abs().max()) <= 5.0                         # clipObservations, Anymal.yaml:10
    assert extras["time_outs"].dtype == torch.bool
    s = env.episode_statistics().cpu()
    assert int(s[7]) == 128 and abs(float(s[0]) - float(env.rew_buf.double().sum())) < 1e-6


def test_unaligned_and_sharded():
    n = 4096 + 17
    st = syn.anymal_state(n, seed=33)
    up, uv, uc = syn.anymal_reset_noise(n, seed=34)
    full, _, _ = U.make_anymal_pair(n, seed=33, state=st)
    full.post_physics_step(up.to(DEV), uv.to(DEV), uc.to(DEV))
    cut = 1000                                   # shard boundary that is not a multiple of 32
    parts = []
    for sl in (slice(0, cut), slice(cut, n)):
        m = sl.stop - sl.start
        sub = {k: (v[sl.start * 12:sl.stop * 12] if k == "dof_state" else v[sl]).clone() for k, v in st.items()}
        env, _, _ = U.make_anymal_pair(m, seed=33, state=sub)
        env.post_physics_step(up[sl].to(DEV), uv[sl].to(DEV), uc[sl].to(DEV))
        parts.append(env)
    torch.cuda.synchronize()
    for name in ("obs_buf", "rew_buf", "reset_buf", "progress_buf", "commands"):
        U.exact(torch.cat([getattr(p, name) for p in parts]), getattr(full, name), "sharded " + name)
