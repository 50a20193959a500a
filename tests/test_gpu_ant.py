"""GPU parity: Ant CUDA path (the Humanoid kernel template instantiated with AntSpec, through the
C ABI) vs the CPU oracle and the vectors frozen from the reference's own JIT functions."""
import os

import numpy as np
import pytest
import torch

import parity_util as U
from thormang_isaacgym_b200 import synthetic as syn

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _check_obs_outputs(env, orc):
    U.ant_obs_close(env.obs_buf, orc.obs_buf)
    U.exact(env.obs_buf[:, 0], orc.obs_buf[:, 0], "obs[:,0] (height, raw copy)")
    # dof_pos_scaled feeds the `> 0.99` count of the reward (ant.py:347): true IEEE division -> identical bits
    U.exact(env.obs_buf[:, 12:20], orc.obs_buf[:, 12:20], "obs[:,12:20] (dof_pos_scaled)")
    U.exact(env.obs_buf[:, 20:52], orc.obs_buf[:, 20:52], "obs[:,20:52] (scaled dof_vel, sensors)")
    U.exact(env.obs_buf[:, 52:60], orc.obs_buf[:, 52:60], "obs[:,52:60] (actions, raw copy)")
    U.close(env.potentials, orc.potentials, atol=0.0, what="potentials")
    U.close(env.prev_potentials, orc.prev_potentials, atol=0.0, what="prev_potentials")
    U.close(env.up_vec, orc.up_vec, what="up_vec")
    U.close(env.heading_vec, orc.heading_vec, what="heading_vec")


# the reward carries potentials - prev_potentials, a difference of two ~6e4-sized fp32 numbers: their ulp (3.9e-3 each)
# is the floor of any comparison that does not start from identical potentials
REW_ATOL = 1e-4


@pytest.mark.parametrize("n", [1, 31, 32, 33, 127, 129, 1000, 4096])
def test_observations_and_reward_split(n):
    env, orc, _ = U.make_ant_pair(n, seed=300 + n)
    env.compute_observations()
    orc.compute_observations()
    _check_obs_outputs(env, orc)
    # same potentials on both sides (they agree to <= 1 ulp above; the reward amplifies that ulp)
    env.potentials.copy_(orc.potentials.to(DEV)); env.prev_potentials.copy_(orc.prev_potentials.to(DEV))
    env.compute_reward(env.actions)
    orc.compute_reward()
    U.exact(env.reset_buf, orc.reset_buf, "reset_buf")
    U.close(env.rew_buf, orc.rew_buf, atol=REW_ATOL, what="rew_buf")
    # on identical obs the only freedom left is the summation order of the 8-term costs
    env.obs_buf.copy_(orc.obs_buf.to(DEV))
    env.reset_buf.zero_(); orc.reset_buf = torch.zeros_like(orc.reset_buf)
    env.compute_reward(env.actions)
    orc.compute_reward()
    U.close(env.rew_buf, orc.rew_buf, what="rew_buf on identical obs")
    U.exact(env.reset_buf, orc.reset_buf, "reset_buf on identical obs")


def test_golden_vectors(golden_dir):
    """CUDA path against the vectors frozen from the reference's compute_ant_observations / compute_ant_reward."""
    z = np.load(os.path.join(golden_dir, "ant_jit.npz"))
    g = {k: torch.from_numpy(z[k]) for k in z.files if z[k].ndim > 0}
    n = g["root_states"].shape[0]
    asset = dict(dof_limits_lower=g["dof_limits_lower"], dof_limits_upper=g["dof_limits_upper"], joint_gears=g["joint_gears"])
    state = dict(root_states=g["root_states"], dof_state=g["dof_state"], sensors=g["sensors"], actions=g["actions"],
                 progress_buf=g["progress_buf"], reset_buf=g["reset_buf"], potentials=g["potentials"],
                 prev_potentials=g["potentials"].clone())
    env, _, _ = U.make_ant_pair(n, state=state, asset=asset)
    env.compute_observations()
    U.ant_obs_close(env.obs_buf, g["obs"], "obs vs golden")
    U.exact(env.obs_buf[:, 12:20], g["obs"][:, 12:20], "dof_pos_scaled vs golden")
    U.close(env.up_vec, g["up_vec"], what="up_vec vs golden")
    U.close(env.heading_vec, g["heading_vec"], what="heading_vec vs golden")
    U.close(env.potentials, g["potentials_out"], atol=0.0, what="potentials vs golden")
    U.exact(env.prev_potentials, g["prev_potentials_out"], "prev_potentials vs golden")
    env.potentials.copy_(g["potentials_out"].to(DEV))
    env.compute_reward(env.actions)
    U.exact(env.reset_buf, g["reset_out"], "reset vs golden")
    U.close(env.rew_buf, g["rew"], atol=REW_ATOL, what="rew vs golden")


@pytest.mark.parametrize("n", [1, 33, 257, 4096])
@pytest.mark.parametrize("mode", ["compact", "dense", "philox"])
def test_fused_post_physics(n, mode):
    from thormang_isaacgym_b200 import _native as nat
    env, orc, st = U.make_ant_pair(n, seed=17 + n)
    if n >= 33:
        for i in (0, 31, 32, n - 1):
            st["reset_buf"][i] = 1
        env.reset_buf.copy_(st["reset_buf"].to(DEV)); orc.reset_buf = st["reset_buf"].clone()
    ids = orc.reset_buf.nonzero().flatten()
    k = len(ids)
    ua, ub = syn.humanoid_reset_noise(max(k, 1), seed=n, num_dof=8)
    if mode == "compact":
        env.post_physics_step(ua.to(DEV), ub.to(DEV))
    elif mode == "dense":
        da, db = torch.zeros(n, 8), torch.zeros(n, 8)
        da[ids], db[ids] = ua[:k], ub[:k]
        env.post_physics_step(da.to(DEV), db.to(DEV), noise_mode=nat.NOISE_DENSE)
    else:
        pa, pb = env.philox_noise()
        ua, ub = pa.cpu()[ids], pb.cpu()[ids]
        env.post_physics_step()
    orc.post_physics_step(ua, ub)
    orc.timeout_buf = (orc.progress_buf >= 999) & (orc.reset_buf != 0)
    _check_obs_outputs(env, orc)
    U.exact(env.reset_buf, orc.reset_buf, "reset_buf")
    U.exact(env.progress_buf, orc.progress_buf, "progress_buf")
    U.exact(env.randomize_buf, orc.randomize_buf, "randomize_buf")
    U.exact(env.timeout_buf, orc.timeout_buf, "timeout_buf")
    # potentials agree to <= 1 ulp (3.9e-3 at 6e4); the progress term of the reward is their difference
    U.close(env.rew_buf, orc.rew_buf, atol=1e-2, what="rew_buf")
    same = (env.potentials.cpu() == orc.potentials) & (env.prev_potentials.cpu() == orc.prev_potentials)
    assert float(same.float().mean()) > 0.5
    U.close(env.rew_buf.cpu()[same], orc.rew_buf[same], atol=REW_ATOL, what="rew_buf where the potentials are bit-equal")
    U.exact(env.dof_state, orc.dof_state, "dof_state after reset")
    U.exact(env.root_states, orc.root_states, "root_states after reset")
    stats = env.episode_statistics().cpu()
    want = torch.tensor([orc.rew_buf.double().sum(), (orc.reset_buf != 0).sum(), orc.timeout_buf.sum(), 0.0,
                         orc.progress_buf[orc.reset_buf != 0].sum(), orc.obs_buf[:, 10].double().sum(),
                         orc.obs_buf[:, 11].double().sum(), n], dtype=torch.float64)
    keep = [0, 1, 2, 4, 5, 6, 7]
    U.close(stats[keep], want[keep], rtol=1e-5, atol=1e-2 * max(1, n // 64), what="episode statistics")


def test_pre_physics_reset_idx_and_step():
    n = 1000
    env, orc, st = U.make_ant_pair(n, seed=5)
    a = (torch.rand(n, 8, generator=torch.Generator().manual_seed(1)) * 3 - 1.5)
    env.pre_physics_step(a.to(DEV))                 # ant.py:273-277: no clamp
    orc.pre_physics_step(a, clip=False)
    U.exact(env.actions, orc.actions, "actions"); U.exact(env.forces, orc.forces, "forces")
    ids = torch.tensor([0, 5, 31, 32, 33, 500, 999])
    ua, ub = syn.humanoid_reset_noise(len(ids), seed=3, num_dof=8)
    env.reset_idx(ids.to(DEV), ua.to(DEV), ub.to(DEV))
    orc.reset_idx(ids, ua, ub)
    for name in ("dof_state", "root_states", "potentials", "prev_potentials", "progress_buf", "reset_buf"):
        U.exact(getattr(env, name), getattr(orc, name), name)
    # VecTask.step: clamp + pre + (identity physics) + fused post + timeout
    pa, pb = env.philox_noise()
    obs_d, rew, reset, extras = env.step(a.to(DEV))
    k_ids = orc.reset_buf.nonzero().flatten()
    orc.step(a, pa.cpu()[k_ids], pb.cpu()[k_ids])
    U.ant_obs_close(obs_d["obs"], orc.obs_buf, "step obs")
    U.exact(reset, orc.reset_buf, "step reset")
    U.exact(extras["time_outs"].cpu().bool(), orc.timeout_buf.bool(), "step time_outs")
    assert obs_d["obs"].shape == (n, 60) and env.num_acts == 8


def test_factory_and_guards():
    import thormang_isaacgym_b200 as pkg
    env = pkg.make(3, "Ant", 64)
    assert env.num_obs == 60 and env.num_acts == 8 and env.obs_buf.shape == (64, 60)
    env.reset()
    env.step(torch.zeros(64, 8, device=DEV))
    torch.cuda.synchronize()
    assert torch.isfinite(env.obs_buf).all()
    with pytest.raises(NotImplementedError):
        env.reward_terms()
