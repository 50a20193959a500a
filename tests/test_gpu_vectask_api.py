"""GPU tests of the drop-in VecTask surface (vec_task.py:313-406): step / reset / reset_done / zero_actions / spaces,
driven exactly like rl_games' RLGPUEnv drives the reference, checked against the oracle's step()."""
import numpy as np
import pytest
import torch

import parity_util as U
from thormang_isaacgym_b200 import synthetic as syn

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_humanoid_step_api_matches_oracle():
    n = 2048
    env, orc, st = U.make_humanoid_pair(n, seed=41)
    assert env.num_envs == n and env.num_obs == 110 and env.num_acts == 21
    assert env.observation_space.shape == (110,) and env.action_space.shape == (21,)
    assert env.zero_actions().shape == (n, 21) and env.reset()["obs"].shape == (n, 110)
    g = torch.Generator().manual_seed(2)
    for t in range(4):
        a = torch.rand(n, 21, generator=g) * 3 - 1.5           # beyond clipActions = 1: the clamp is exercised
        ua, ub = env.philox_noise()                              # the draws the fused kernel is about to use
        ids = orc.reset_buf.nonzero().flatten()
        obs_d, rew, reset, extras = env.step(a.to(DEV))
        o_obs, o_rew, o_reset, o_ex = orc.step(a, ua.cpu()[ids] if len(ids) else ua.cpu()[:1],
                                               ub.cpu()[ids] if len(ids) else ub.cpu()[:1])
        assert obs_d["obs"].data_ptr() == env.obs_buf.data_ptr(), "clip_obs = inf: no extra obs copy"
        U.humanoid_obs_close(obs_d["obs"], o_obs, f"obs t={t}")
        U.exact(reset, o_reset, f"reset t={t}")
        U.exact(extras["time_outs"], o_ex["time_outs"], f"time_outs t={t}")
        U.close(rew, o_rew, what=f"rew t={t}")
        U.exact(env.actions, orc.actions, f"clamped actions t={t}")
        assert env.sim.actuation_force is env.forces
    # reset_done(): resets the flagged envs through reset_idx (torch-drawn noise) and returns their ids
    flagged = env.reset_buf.nonzero().flatten()
    obs_d, done_ids = env.reset_done()
    U.exact(done_ids, flagged, "reset_done ids")
    assert int(env.reset_buf.sum()) == 0 and int(env.progress_buf[flagged].sum()) == 0


def test_humanoid_finite_clip_obs_and_rebind():
    n = 256
    env, orc, _ = U.make_humanoid_pair(n, seed=43, clipObservations=0.5)
    obs_d, *_ = env.step(torch.zeros(n, 21, device=DEV))
    assert float(obs_d["obs"].abs().max()) <= 0.5 and float(env.obs_buf.abs().max()) > 0.5
    # replacing a buffer object requires rebind(); the kernels then write the new tensor
    env.rew_buf = torch.full((n,), 123.0, device=DEV)
    env.rebind()
    env.post_physics_step()
    assert float(env.rew_buf.max()) <= 2.1 + 1e-6


def test_gogoro_step_api():
    n = 1024
    env, orc, st = U.make_gogoro_pair(n, seed=44)
    assert env.num_obs == 6 and env.num_acts == 1
    a = st["actions"]
    for t in range(3):
        nz = {k: v.cpu() for k, v in env.philox_noise().items()}
        ids = orc.reset_buf.nonzero().flatten()
        sp = ((orc.reset_buf == 0) & (orc.progress_buf + 1 == 300)).nonzero().flatten()
        obs_d, rew, reset, extras = env.step(a.to(DEV))
        orc.step(a, nz["steer_unit"], nz["sensor_unit"], nz["reset_unit"][ids] if len(ids) else nz["reset_unit"][:1],
                 nz["speed_unit"][sp] if len(sp) else nz["speed_unit"][:1], nz["yaw_unit"])
        U.exact(reset, orc.reset_buf, f"reset t={t}")
        U.exact(extras["time_outs"], orc.timeout_buf, f"time_outs t={t}")
        U.gogoro_obs_close(obs_d["obs"], orc.obs_buf, f"obs t={t}", noisy=True)
        U.close(rew, orc.rew_buf, rtol=2e-5, atol=2e-5, what=f"rew t={t}")
        # keep both sides bit-identical where libm differences could drift (spawn quaternion of reset envs)
        orc.root_tensor.copy_(env.root_tensor.cpu()); orc.yaw_command.copy_(env.yaw_command.cpu())
        orc.curent_command.copy_(env.curent_command.cpu())
    assert env.sim.dof_position_target is env.dof_pos_command and env.sim.dof_velocity_target is env.dof_speed_command


def test_factory_and_registry():
    import thormang_isaacgym_b200 as pkg
    from thormang_isaacgym_b200.tasks import isaacgym_task_map
    assert {"Humanoid", "Gogoro"} <= set(isaacgym_task_map)
    env = pkg.make(seed=3, task="Humanoid", num_envs=64, sim_device=DEV, rl_device=DEV)
    obs_d, rew, reset, extras = env.step(env.zero_actions())
    assert obs_d["obs"].shape == (64, 110) and rew.shape == (64,) and reset.dtype == torch.int64
    g = pkg.make(seed=3, task="Gogoro", num_envs=64, sim_device=DEV, rl_device=DEV)
    obs_d, rew, reset, extras = g.step(g.zero_actions())
    assert obs_d["obs"].shape == (64, 6) and extras["time_outs"].dtype == torch.bool
