"""Tiny driver for compute-sanitizer (memcheck / racecheck): every kernel once at ragged sizes.

    compute-sanitizer --tool memcheck python tools/sanitize_smoke.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

import parity_util as U  # noqa: E402
from thormang_isaacgym_b200 import synthetic as syn  # noqa: E402

for n in (1, 33, 4096 + 7):
    env, orc, st = U.make_humanoid_pair(n, seed=n)
    a = torch.rand(n, 21) * 2 - 1
    env.pre_physics_step(a.cuda())
    ids = orc.reset_buf.nonzero().flatten()
    ua, ub = syn.humanoid_reset_noise(max(len(ids), 1), seed=1)
    env.post_physics_step(ua.cuda(), ub.cuda())      # COMPACT noise (count + scan kernels)
    env.post_physics_step()                          # Philox
    env.compute_observations(); env.compute_reward()
    env.reset_idx(torch.arange(min(n, 5), device="cuda:0"))
    env.reward_terms(); env.wm_reward(env.obs_buf); env.episode_statistics(); env.advance_noise_epoch()
    g, gorc, gst = U.make_gogoro_pair(n, seed=n)
    nz = {k: v.cuda() for k, v in syn.gogoro_noise(n, seed=2).items()}
    g.pre_physics_step(gst["actions"].cuda(), nz["steer_unit"]); g.pre_physics_step(gst["actions"].cuda())
    g.post_physics_step(nz["sensor_unit"], nz["reset_unit"], nz["speed_unit"], nz["yaw_unit"])
    g.post_physics_step()
    g.compute_obs_rwd(nz["sensor_unit"]); g.compute_observations(); g.compute_reward()
    g.reset_idx(torch.arange(min(n, 5), device="cuda:0")); g.episode_statistics(); g.philox_noise()
    # the four later kernel families (Philox paths: no supplied tensors needed), two steps each so that resets of step 1 run
    ant, _, ast = U.make_ant_pair(n, seed=n)
    ant.pre_physics_step((torch.rand(n, 8) * 2 - 1).cuda()); ant.post_physics_step(); ant.post_physics_step()
    ant.compute_observations(); ant.compute_reward(); ant.episode_statistics()
    for turning in (False, True):
        pp, _, pst = U.make_gogoro_paper_pair(n, seed=n, turning=turning)
        pp.pre_physics_step(pst["actions"].cuda()); pp.post_physics_step(); pp.post_physics_step()
        pp.compute_observations(); pp.reset_idx(torch.arange(min(n, 5), device="cuda:0")); pp.episode_statistics()
        _ = pp.buffer_obs
    an, _, anst = U.make_anymal_pair(n, seed=n)
    an.pre_physics_step(anst["actions"].cuda()); an.post_physics_step(); an.post_physics_step()
    an.compute_observations(); an.compute_reward(); an.reset_idx(torch.arange(min(n, 5), device="cuda:0")); an.episode_statistics()
    op, _, opst = U.make_ma_op3_pair(n, seed=n)
    op.pre_physics_step(opst["actions"].cuda()); op.post_physics_step(); op.post_physics_step()
    op.compute_observations(); op.compute_reward(); op.reset_idx(torch.arange(min(n, 5), device="cuda:0")); op.episode_statistics()
    # VecTask.step epilogue (noise lambda + clamp kernel) through the public step() of a task with a finite clip
    obs_dict, rew, done, info = g.step(gst["actions"].cuda())
    torch.cuda.synchronize()
print("sanitize smoke done")
