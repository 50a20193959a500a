"""GPU parity of the 8-channel x 20-frame Gogoro variant against the CPU oracle
(oracle/gogoro_paper_oracle.py, pinned bit-for-bit to gogoro_realistic_turning_sim_paper.py's JIT functions).

Bar: reset / progress / timeout masks and the steer-delay integers bit-exact; reset-state writes exact; frames that are
only moved (the 19 old frames of every env) bit-exact; fp32 observations and rewards within 1e-5 relative
(tests/parity_util.py).
"""
import math

import pytest
import torch

import parity_util as U
from thormang_isaacgym_b200 import synthetic as syn

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _hit_sets(orc, max_len_freq=300):
    """Ids the next post_physics_step will reset / resample (needed to hand the oracle its compact draws)."""
    reset_ids = orc.reset_buf.nonzero(as_tuple=False).squeeze(-1)
    prog = orc.progress_buf + 1
    prog[reset_ids] = 0
    speed_ids = (prog == max_len_freq).nonzero(as_tuple=False).flatten()
    return reset_ids, speed_ids


def _post_both(env, orc, nz):
    reset_ids, speed_ids = _hit_sets(orc)
    env.post_physics_step(nz["sensor_unit"].to(DEV), nz["reset_unit"].to(DEV), nz["speed_unit"].to(DEV), nz["yaw_unit"].to(DEV))
    orc.post_physics_step(nz["sensor_unit"], nz["reset_unit"][reset_ids], nz["speed_unit"][speed_ids], nz["yaw_unit"])
    orc.timeout_buf = (orc.progress_buf >= orc.max_episode_length - 1) & (orc.reset_buf != 0)
    return reset_ids, speed_ids


def _compare_step(env, orc, what):
    n = orc.n_envs
    torch.cuda.synchronize()
    U.exact(env.reset_buf, orc.reset_buf, what + " reset_buf")
    U.exact(env.progress_buf, orc.progress_buf, what + " progress_buf")
    U.exact(env.timeout_buf, orc.timeout_buf, what + " timeout_buf")
    U.exact(env.steer_delay, orc.steer_delay, what + " steer_delay")
    got = env.obs_buf.view(n, 20, 8).cpu()
    want = orc.obs_buf.view(n, 20, 8)
    U.exact(got[:, :-1], want[:, :-1], what + " moved noisy frames")
    U.gogoro_paper_frame_close(got[:, -1], want[:, -1], what + " fresh noisy frame", noisy=True)
    clean = env.buffer_obs.cpu()
    U.exact(clean[:, :-1], orc.buffer_obs[:, :-1], what + " old clean frames")
    U.gogoro_paper_frame_close(clean[:, -1], orc.buffer_obs[:, -1], what + " fresh clean frame")
    assert bool(torch.isfinite(env.rew_buf).all()), what + ": the fall sentinel must not survive the finalize kernel"
    U.close(env.rew_buf, orc.rew_buf, what=what + " rew_buf")
    U.exact(env.rew_buf.cpu() == -1.0, orc.rew_buf == -1.0, what + " fall penalty mask")
    U.close(env.speed_no_noise, orc.speed_no_noise, what=what + " speed_no_noise")
    U.exact(env.root_tensor, orc.root_tensor, what + " root_tensor")
    U.exact(env.state_dof, orc.state_dof, what + " state_dof")
    for name in ("curent_speed", "steer_offsets", "curent_speed_offset", "curent_imu_x_offset", "curent_command",
                 "command_history"):
        U.exact(getattr(env, name), getattr(orc, name), what + " " + name)
    U.angle_close(env.yaw_command, orc.yaw_command, what=what + " yaw_command")


@pytest.mark.parametrize("n", [1, 31, 32, 33, 1000, 4096])
def test_observations(n):
    from oracle import gogoro_paper_oracle as gp
    env, orc, st = U.make_gogoro_paper_pair(n, seed=300 + n)
    got = env.compute_observations()
    want = gp.paper_observations(st["root_tensor"].clone(), st["yaw_command"].clone(), st["curent_command"].clone(),
                                 (st["steer_delay"].clone() - 0) / (5 - 0))
    U.gogoro_paper_frame_close(got, want, f"paper obs n={n}")
    # nothing else may have been written
    U.exact(env.obs_buf.view(n, 20, 8), st["buffer_obs_noisy"], "obs_buf untouched")


def test_golden_vectors(golden_dir):
    import os
    import numpy as np
    z = np.load(os.path.join(golden_dir, "gogoro_paper_jit.npz"))
    g = {k: torch.from_numpy(z[k]) for k in z.files}
    n = g["root_tensor"].shape[0]
    st = syn.gogoro_paper_state(n, seed=1)
    st.update(root_tensor=g["root_tensor"], yaw_command=g["desired_yaw"], curent_command=g["curent_command"],
              steer_delay=g["steer_delay"], progress_buf=g["progress_buf"], reset_buf=torch.zeros(n, dtype=torch.long))
    # the golden frames already contain the fresh frame: put frames[:, :-1] where the step will find them (t = 1..19)
    st["buffer_obs"] = torch.cat([torch.zeros(n, 1, 8), g["frames"][:, :-1]], dim=1)
    env, orc, _ = U.make_gogoro_paper_pair(n, seed=1, state=st)
    U.gogoro_paper_frame_close(env.compute_observations(), g["obs"], "golden obs")
    env.compute_obs_rwd(torch.full((n, 6), 0.5))
    torch.cuda.synchronize()
    U.exact(env.reset_buf, g["reset"], "golden reset")
    U.close(env.rew_buf, g["rew"], what="golden reward")


@pytest.mark.parametrize("use_delay", [False, True])
def test_pre_physics(use_delay):
    n = 1000
    env, orc, st = U.make_gogoro_paper_pair(n, seed=41, useSteerDelay=use_delay)
    a = st["actions"]
    env.pre_physics_step(a.to(DEV))
    orc.pre_physics_step(a)
    torch.cuda.synchronize()
    U.exact(env.curent_command, orc.curent_command, "curent_command")
    U.exact(env.command_history, orc.command_history, "command_history")
    U.exact(env.dof_pos_command, orc.dof_pos_command, "dof_pos_command")
    U.exact(env.dof_speed_command, orc.dof_speed_command, "dof_speed_command")
    if use_delay:
        assert len(set(st["steer_delay"].tolist())) == 5     # taps 0..4 all exercised, incl. the "-0 = column 0" quirk


def test_reset_idx():
    n = 500
    env, orc, st = U.make_gogoro_paper_pair(n, seed=77)
    ids = torch.tensor([0, 3, 31, 32, 33, 250, 499])
    unit = torch.rand((len(ids), 5), generator=torch.Generator().manual_seed(2))
    env.reset_idx(ids.to(DEV), unit.to(DEV))
    orc.reset_idx(ids, unit)
    torch.cuda.synchronize()
    n_ = orc.n_envs
    U.exact(env.obs_buf.view(n_, 20, 8), orc.buffer_obs_noisy, "noisy frames")
    U.exact(env.buffer_obs, orc.buffer_obs, "clean frames")
    for name in ("root_tensor", "state_dof", "progress_buf", "reset_buf", "steer_delay", "curent_speed", "steer_offsets",
                 "curent_speed_offset", "curent_imu_x_offset", "curent_command", "yaw_command", "command_history",
                 "speed_no_noise"):
        U.exact(getattr(env, name), getattr(orc, name), name)


@pytest.mark.parametrize("n", [1, 33, 1000, 4096, 65536])
def test_fused_post_physics(n):
    env, orc, st = U.make_gogoro_paper_pair(n, seed=900 + n)
    nz = syn.gogoro_paper_noise(n, seed=n)
    reset_ids, speed_ids = _post_both(env, orc, nz)
    if n >= 1000:
        assert len(reset_ids) > 0 and len(speed_ids) > 0
    _compare_step(env, orc, f"paper post n={n}")


def test_compute_obs_rwd_method():
    n = 777
    env, orc, st = U.make_gogoro_paper_pair(n, seed=12)
    nz = syn.gogoro_paper_noise(n, seed=3)
    env.compute_obs_rwd(nz["sensor_unit"].to(DEV))
    orc.compute_obs_rwd(nz["sensor_unit"])
    torch.cuda.synchronize()
    U.exact(env.progress_buf, st["progress_buf"], "progress untouched")
    got, want = env.obs_buf.view(n, 20, 8).cpu(), orc.obs_buf.view(n, 20, 8)
    U.exact(got[:, :-1], want[:, :-1], "moved frames")
    U.gogoro_paper_frame_close(got[:, -1], want[:, -1], "fresh frame", noisy=True)
    U.close(env.rew_buf, orc.rew_buf, what="rew")
    U.exact(env.reset_buf, orc.reset_buf, "reset")


def test_rollout_wraps_the_ring():
    """30 steps of pre -> (identity physics) -> post with fresh noise: the clean ring wraps (20 slots), resets and
    resamples happen along the way, every step is compared."""
    n = 256
    st = syn.gogoro_paper_state(n, seed=5, max_steps=40)
    env, orc, st = U.make_gogoro_paper_pair(n, seed=5, state=st, max_steps=40)
    assert orc.max_episode_length == 40
    g = torch.Generator().manual_seed(9)
    for step in range(30):
        a = torch.rand((n, 1), generator=g) * 3 - 1.5
        env.pre_physics_step(a.to(DEV))
        orc.pre_physics_step(a)
        # a little motion so that the frames differ from step to step
        drift = torch.randn((n, 3), generator=g) * 0.05
        orc.root_tensor[:, 10:13] += drift
        env.root_tensor[:, 10:13] += drift.to(DEV)
        nz = syn.gogoro_paper_noise(n, seed=1000 + step)
        _post_both(env, orc, nz)
        _compare_step(env, orc, f"rollout step {step}")
        # the fresh frames agree within tolerance, not bit for bit; they become MOVED frames (compared exactly) from
        # the next step on, so the oracle continues from the device's values of what was just checked
        orc.buffer_obs_noisy[:, -1] = env.obs_buf.view(n, 20, 8)[:, -1].cpu()
        orc.buffer_obs[:, -1] = env.buffer_obs[:, -1].cpu()
        orc.yaw_command = env.yaw_command.cpu().clone()
    assert int(orc.progress_buf.max()) < 40


def test_philox_noise_mode():
    n = 2000
    env, orc, st = U.make_gogoro_paper_pair(n, seed=23)
    env.step_count = 7
    nz = {k: v.cpu() for k, v in env.philox_noise().items()}
    for v in nz.values():
        assert float(v.min()) >= 0.0 and float(v.max()) < 1.0
    assert abs(float(nz["sensor_unit"].mean()) - 0.5) < 0.02
    reset_ids, speed_ids = _hit_sets(orc)
    env.post_physics_step()                        # in-kernel Philox
    orc.post_physics_step(nz["sensor_unit"], nz["reset_unit"][reset_ids], nz["speed_unit"][speed_ids], nz["yaw_unit"])
    orc.timeout_buf = (orc.progress_buf >= orc.max_episode_length - 1) & (orc.reset_buf != 0)
    _compare_step(env, orc, "paper philox")
    # the next step draws different noise
    again = env.philox_noise()
    assert not torch.equal(again["sensor_unit"].cpu(), nz["sensor_unit"])


def test_sharded_equals_unsharded_with_summed_diff_sums():
    """reward7 couples every env: two half-size tasks whose difference sums are added (what the all-reduce does)
    must give exactly the rewards of the unsharded task."""
    from thormang_isaacgym_b200 import _native as nat
    n = 2048
    st = syn.gogoro_paper_state(n, seed=61)
    nz = syn.gogoro_paper_noise(n, seed=62)
    full, _, _ = U.make_gogoro_paper_pair(n, seed=61, state=st)
    full.post_physics_step(nz["sensor_unit"].to(DEV), nz["reset_unit"].to(DEV), nz["speed_unit"].to(DEV), nz["yaw_unit"].to(DEV))
    halves, sums = [], torch.zeros(2, device=DEV, dtype=torch.float64)
    for r in range(2):
        sl = slice(r * n // 2, (r + 1) * n // 2)
        sub = {k: v[sl].clone() if v.shape[0] == n else v[sl.start * 39:sl.stop * 39].clone() for k, v in st.items()}
        env, _, _ = U.make_gogoro_paper_pair(n // 2, seed=61, state=sub)
        keep = []
        dz = env._noise(nz["sensor_unit"][sl], nz["reset_unit"][sl], nz["speed_unit"][sl], nz["yaw_unit"][sl], keep)
        nat.check(env._lib.tk_gogoro_paper_post_physics(env._handle, dz, env._stream()))
        nat.check(env._lib.tk_gogoro_paper_diff_sums(env._handle, nat.ptr(env.diff_sums), env._stream()))
        sums += env.diff_sums
        halves.append(env)
    for env in halves:
        nat.check(env._lib.tk_gogoro_paper_finalize_reward(env._handle, nat.ptr(sums), env._stream()))
    torch.cuda.synchronize()
    assert float(sums[1]) == n * 19
    got = torch.cat([h.rew_buf for h in halves])
    U.close(got, full.rew_buf, rtol=0, atol=3e-7, what="sharded reward")       # one fp32 rounding of the global mean
    U.exact(torch.cat([h.obs_buf for h in halves]), full.obs_buf, "sharded obs")
    U.exact(torch.cat([h.reset_buf for h in halves]), full.reset_buf, "sharded reset")


def test_world_reduce_path_single_rank_nccl():
    import torch.distributed as dist
    if not dist.is_initialized():
        dist.init_process_group("nccl", init_method="tcp://127.0.0.1:29631", rank=0, world_size=1,
                                device_id=torch.device(DEV))
    try:
        n = 1000
        st = syn.gogoro_paper_state(n, seed=71)
        nz = syn.gogoro_paper_noise(n, seed=72)
        a, _, _ = U.make_gogoro_paper_pair(n, seed=71, state=st)
        b, _, _ = U.make_gogoro_paper_pair(n, seed=71, state=st, worldReduce=True)
        for env in (a, b):
            env.post_physics_step(nz["sensor_unit"].to(DEV), nz["reset_unit"].to(DEV), nz["speed_unit"].to(DEV),
                                  nz["yaw_unit"].to(DEV))
        torch.cuda.synchronize()
        U.exact(a.rew_buf, b.rew_buf, "all-reduced (1 rank) reward")
    finally:
        dist.destroy_process_group()


def test_cuda_graph_replay_matches_eager():
    """The frame ring position is device-side state: a captured step replays correctly."""
    n = 512
    st = syn.gogoro_paper_state(n, seed=81)
    nz = {k: v.to(DEV) for k, v in syn.gogoro_paper_noise(n, seed=82).items()}
    act = st["actions"].to(DEV)
    eager, _, _ = U.make_gogoro_paper_pair(n, seed=81, state=st)
    graph_env, _, _ = U.make_gogoro_paper_pair(n, seed=81, state=st)
    steps = 23

    def one(env):
        env.pre_physics_step(act)
        env.post_physics_step(nz["sensor_unit"], nz["reset_unit"], nz["speed_unit"], nz["yaw_unit"])

    for _ in range(steps):
        one(eager)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        one(graph_env)                     # warm-up step outside the capture
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            one(graph_env)
        for _ in range(steps - 1):         # the capture itself executes nothing
            g.replay()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    U.exact(graph_env.obs_buf, eager.obs_buf, "graph obs_buf")
    U.exact(graph_env.rew_buf, eager.rew_buf, "graph rew_buf")
    U.exact(graph_env.buffer_obs, eager.buffer_obs, "graph clean frames")
    U.exact(graph_env.progress_buf, eager.progress_buf, "graph progress")


def test_episode_statistics():
    n = 3000
    env, orc, st = U.make_gogoro_paper_pair(n, seed=91)
    nz = syn.gogoro_paper_noise(n, seed=92)
    _post_both(env, orc, nz)
    s = env.episode_statistics().cpu()
    assert abs(float(s[0]) - float(orc.rew_buf.double().sum())) < 1e-3 * n
    assert int(s[1]) == int((orc.reset_buf != 0).sum()) and int(s[2]) == int(orc.timeout_buf.sum()) and int(s[7]) == n
    assert abs(float(s[5]) - float(orc.buffer_obs[:, -1, 0].abs().double().sum())) < 1e-2


def test_error_behaviour():
    from thormang_isaacgym_b200 import _native as nat
    env, _, _ = U.make_gogoro_paper_pair(64, seed=3)
    with pytest.raises(ValueError):
        env.post_physics_step(torch.rand(10, 6, device=DEV))          # supplied draws must cover every env id
    bad = nat.GogoroPaperNoise()
    bad.mode = nat.NOISE_COMPACT
    import ctypes as C
    assert env._lib.tk_gogoro_paper_post_physics(env._handle, C.byref(bad), env._stream()) == -1
    assert b"TK_NOISE_DENSE" in env._lib.tk_last_error()


@pytest.mark.parametrize("turning", [False, True])
def test_cuda_graph_epoch_directly_in_front_of_the_step_many_replays(turning):
    """In-kernel Philox noise under graph replay: the fused kernel resolves the device-side epoch ahead of its dependency
    wait.  Epoch kernel directly in front of one post_physics step (two kernels), 25 replays, against eager launches keyed
    (epoch << 32) | step."""
    n, replays = 2048, 25
    st = syn.gogoro_paper_state(n, seed=83)
    env_a, _, _ = U.make_gogoro_paper_pair(n, seed=83, state={k: v.clone() for k, v in st.items()}, turning=turning)
    env_b, _, _ = U.make_gogoro_paper_pair(n, seed=83, state={k: v.clone() for k, v in st.items()}, turning=turning)
    for e in (env_a, env_b):
        e.reset_buf[::7] = 1
    for r in range(replays):
        env_a.step_count = (r + 1) << 32
        env_a.post_physics_step()
        env_a.reset_buf[::7] = 1
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    env_b.step_count = 0
    with torch.cuda.graph(g):
        env_b.advance_noise_epoch()
        env_b.post_physics_step()
        env_b.reset_buf[::7] = 1
    for _ in range(replays):
        g.replay()
    torch.cuda.synchronize()
    for name in ("obs_buf", "rew_buf", "progress_buf", "root_tensor", "yaw_command", "curent_speed", "steer_delay"):
        U.exact(getattr(env_b, name), getattr(env_a, name), f"graph vs eager after {replays} replays: {name}")
    U.exact(env_b.buffer_obs, env_a.buffer_obs, "clean frames")
