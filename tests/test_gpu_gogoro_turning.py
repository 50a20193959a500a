"""GPU parity of the gogoro_realistic_turning.py sibling of the 8-channel x 20-frame Gogoro variant (SURVEY.md 8(f)
row N1, Appendix B) against the CPU oracle (oracle/gogoro_turning_oracle.py, pinned bit-for-bit to that file's JIT
functions).

Bar: reset / progress / timeout masks and the steer-delay integers bit-exact; moved frames bit-exact; fp32
observations and rewards within 1e-5 relative; the spawn quaternion / start velocity of reset envs (sin / cos of random
angles) within 1e-5 relative + 2e-6 absolute, everything else a reset writes exact (tests/parity_util.py).
"""
import pytest
import torch

import parity_util as U
from thormang_isaacgym_b200 import synthetic as syn

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
COLS = 8


def _hit_sets(orc, freq=300):
    reset_ids = orc.reset_buf.nonzero(as_tuple=False).squeeze(-1)
    prog = orc.progress_buf + 1
    prog[reset_ids] = 0
    return reset_ids, (prog == freq).nonzero(as_tuple=False).flatten()


def _post_both(env, orc, nz):
    reset_ids, speed_ids = _hit_sets(orc)
    env.post_physics_step(nz["sensor_unit"].to(DEV), nz["reset_unit"].to(DEV), nz["speed_unit"].to(DEV), nz["yaw_unit"].to(DEV))
    orc.post_physics_step(nz["sensor_unit"], nz["reset_unit"][reset_ids], nz["speed_unit"][speed_ids], nz["yaw_unit"])
    orc.timeout_buf = (orc.progress_buf >= orc.max_episode_length - 1) & (orc.reset_buf != 0)
    return reset_ids, speed_ids


def _root_close(env, orc, reset_ids, what):
    """Rows of envs that were not reset are untouched (exact); a reset row carries sin / cos of the random spawn angles."""
    n = orc.n_envs
    keep = torch.ones(n, dtype=torch.bool)
    keep[reset_ids] = False
    got = env.root_tensor.cpu()
    U.exact(got[keep], orc.root_tensor[keep], what + " root_tensor (kept rows)")
    if len(reset_ids):
        U.exact(got[reset_ids][:, [0, 1, 2, 9, 10, 11, 12]], orc.root_tensor[reset_ids][:, [0, 1, 2, 9, 10, 11, 12]],
                what + " root_tensor (reset rows, copied columns)")
        U.close(got[reset_ids][:, 3:9], orc.root_tensor[reset_ids][:, 3:9], atol=2e-6, what=what + " spawn quaternion / velocity")


def _compare_step(env, orc, reset_ids, what):
    n = orc.n_envs
    torch.cuda.synchronize()
    U.exact(env.reset_buf, orc.reset_buf, what + " reset_buf")
    U.exact(env.progress_buf, orc.progress_buf, what + " progress_buf")
    U.exact(env.timeout_buf, orc.timeout_buf, what + " timeout_buf")
    U.exact(env.steer_delay, orc.steer_delay, what + " steer_delay")
    got, want = env.obs_buf.view(n, 20, 8).cpu(), orc.obs_buf.view(n, 20, 8)
    U.exact(got[:, :-1], want[:, :-1], what + " moved noisy frames")
    U.gogoro_paper_frame_close(got[:, -1], want[:, -1], what + " fresh noisy frame", noisy=True)
    clean = env.buffer_obs.cpu()
    U.exact(clean[:, :-1], orc.buffer_obs[:, :-1], what + " old clean frames")
    U.gogoro_paper_frame_close(clean[:, -1], orc.buffer_obs[:, -1], what + " fresh clean frame")
    assert bool(torch.isfinite(env.rew_buf).all()), what + ": the fall sentinel must not survive the finalize kernel"
    U.close(env.rew_buf, orc.rew_buf, what=what + " rew_buf")
    U.exact(env.rew_buf.cpu() == -1.0, orc.rew_buf == -1.0, what + " fall penalty mask")
    U.close(env.speed_no_noise, orc.speed_no_noise, what=what + " speed_no_noise")
    _root_close(env, orc, reset_ids, what)
    U.exact(env.state_dof, orc.state_dof, what + " state_dof")
    for name in ("curent_speed", "steer_offsets", "curent_speed_offset", "curent_imu_x_offset", "curent_command",
                 "command_history"):
        U.exact(getattr(env, name), getattr(orc, name), what + " " + name)
    U.angle_close(env.yaw_command, orc.yaw_command, what=what + " yaw_command")


def test_registered_and_variant_fields():
    import thormang_isaacgym_b200 as pkg
    from thormang_isaacgym_b200 import _native as nat
    env = pkg.make(seed=3, task="GogoroTurning", num_envs=96, sim_device=DEV, rl_device=DEV)
    assert env.VARIANT == nat.GOGORO_PAPER_VARIANT_TURNING and env.use_steer_delay and env.RESET_NOISE_COLS == COLS
    obs_d, rew, reset, extras = env.step(torch.zeros((96, 1), device=DEV))
    assert obs_d["obs"].shape == (96, 160) and rew.shape == (96,) and reset.dtype == torch.int64
    # all-zero command history: reward5 = 1 exactly, reward3 = 1: every standing env earns at least 20
    ok = rew >= 0
    assert bool(ok.any()) and float(rew[ok].min()) >= 20.0 and float(rew[ok].max()) <= 26.0 + 1e-4


def test_golden_vectors(golden_dir):
    import os
    import numpy as np
    z = np.load(os.path.join(golden_dir, "gogoro_turning_jit.npz"))
    g = {k: torch.from_numpy(z[k]) for k in z.files}
    n = g["root_tensor"].shape[0]
    st = syn.gogoro_paper_state(n, seed=1)
    st.update(root_tensor=g["root_tensor"], yaw_command=g["desired_yaw"], curent_command=g["curent_command"],
              steer_delay=g["steer_delay"], progress_buf=g["progress_buf"], reset_buf=torch.zeros(n, dtype=torch.long))
    st["buffer_obs"] = torch.cat([torch.zeros(n, 1, 8), g["frames"][:, :-1]], dim=1)
    env, orc, _ = U.make_gogoro_paper_pair(n, seed=1, state=st, turning=True)
    U.gogoro_paper_frame_close(env.compute_observations(), g["obs"], "golden obs")
    env.compute_obs_rwd(torch.full((n, 6), 0.5))
    torch.cuda.synchronize()
    U.exact(env.reset_buf, g["reset"], "golden reset")
    U.close(env.rew_buf, g["rew"], what="golden reward")


def test_pre_physics_adds_the_steering_offset():
    n = 1000
    env, orc, st = U.make_gogoro_paper_pair(n, seed=41, turning=True)
    a = st["actions"]
    env.pre_physics_step(a.to(DEV))
    orc.pre_physics_step(a)
    torch.cuda.synchronize()
    U.exact(env.curent_command, orc.curent_command, "curent_command")
    U.exact(env.command_history, orc.command_history, "command_history")
    U.exact(env.dof_pos_command, orc.dof_pos_command, "dof_pos_command")
    U.exact(env.dof_speed_command, orc.dof_speed_command, "dof_speed_command")
    assert len(set(st["steer_delay"].tolist())) == 5 and float(st["steer_offsets"].abs().min()) > 0.0


def test_reset_idx():
    n = 500
    env, orc, st = U.make_gogoro_paper_pair(n, seed=77, turning=True)
    ids = torch.tensor([0, 3, 31, 32, 33, 250, 499])
    unit = torch.rand((len(ids), COLS), generator=torch.Generator().manual_seed(2))
    env.reset_idx(ids.to(DEV), unit.to(DEV))
    orc.reset_idx(ids, unit)
    torch.cuda.synchronize()
    U.exact(env.obs_buf.view(n, 20, 8), orc.buffer_obs_noisy, "noisy frames")
    U.exact(env.buffer_obs, orc.buffer_obs, "clean frames")
    _root_close(env, orc, ids, "reset_idx")
    for name in ("state_dof", "progress_buf", "reset_buf", "steer_delay", "curent_speed", "steer_offsets",
                 "curent_speed_offset", "curent_imu_x_offset", "curent_command", "yaw_command", "command_history",
                 "speed_no_noise"):
        U.exact(getattr(env, name), getattr(orc, name), name)
    q = env.root_tensor[ids.to(DEV), 3:7]
    assert float((q.norm(dim=-1) - 1).abs().max()) < 1e-6 and float(q[:, 1].abs().max()) > 0.0    # roll AND yaw are random


@pytest.mark.parametrize("n", [1, 33, 1000, 4096, 65536])
def test_fused_post_physics(n):
    env, orc, st = U.make_gogoro_paper_pair(n, seed=900 + n, turning=True)
    nz = syn.gogoro_paper_noise(n, seed=n, reset_cols=COLS)
    reset_ids, speed_ids = _post_both(env, orc, nz)
    if n >= 1000:
        assert len(reset_ids) > 0 and len(speed_ids) > 0
    _compare_step(env, orc, reset_ids, f"turning post n={n}")


def test_command_mean_away_from_zero():
    """reward5 = 1 - |mean(act_buff)| with a biased command history (the synthetic one is centred on zero)."""
    n = 3000
    st = syn.gogoro_paper_state(n, seed=15)
    st["buffer_obs"][:, :, 6] = st["buffer_obs"][:, :, 6] * 0.3 - 0.31
    st["curent_command"] = st["curent_command"] * 0.3 - 0.31
    env, orc, _ = U.make_gogoro_paper_pair(n, seed=15, state=st, turning=True)
    nz = syn.gogoro_paper_noise(n, seed=16, reset_cols=COLS)
    reset_ids, _ = _post_both(env, orc, nz)
    _compare_step(env, orc, reset_ids, "biased commands")
    mean = float(orc.buffer_obs[:, :, 6].mean())
    assert mean < -0.2
    stand = orc.rew_buf >= 0
    assert float(orc.rew_buf[stand].min()) >= 10.0 + 10.0 * (1.0 + mean) - 1e-3


def test_rollout_wraps_the_ring():
    n = 256
    st = syn.gogoro_paper_state(n, seed=5, max_steps=40)
    env, orc, st = U.make_gogoro_paper_pair(n, seed=5, state=st, max_steps=40, turning=True)
    g = torch.Generator().manual_seed(9)
    for step in range(30):
        a = torch.rand((n, 1), generator=g) * 3 - 1.5
        env.pre_physics_step(a.to(DEV))
        orc.pre_physics_step(a)
        U.exact(env.dof_pos_command, orc.dof_pos_command, f"rollout step {step} dof_pos_command")
        drift = torch.randn((n, 3), generator=g) * 0.05
        orc.root_tensor[:, 10:13] += drift
        env.root_tensor[:, 10:13] += drift.to(DEV)
        nz = syn.gogoro_paper_noise(n, seed=1000 + step, reset_cols=COLS)
        reset_ids, _ = _post_both(env, orc, nz)
        _compare_step(env, orc, reset_ids, f"rollout step {step}")
        # continue from the device's values of what was just checked within tolerance
        orc.buffer_obs_noisy[:, -1] = env.obs_buf.view(n, 20, 8)[:, -1].cpu()
        orc.buffer_obs[:, -1] = env.buffer_obs[:, -1].cpu()
        orc.yaw_command = env.yaw_command.cpu().clone()
        orc.root_tensor.copy_(env.root_tensor.cpu())
    assert int(orc.progress_buf.max()) < 40


def test_philox_noise_mode():
    n = 2000
    env, orc, st = U.make_gogoro_paper_pair(n, seed=23, turning=True)
    env.step_count = 7
    nz = {k: v.cpu() for k, v in env.philox_noise().items()}
    assert nz["reset_unit"].shape == (n, COLS)
    for v in nz.values():
        assert float(v.min()) >= 0.0 and float(v.max()) < 1.0
    assert abs(float(nz["reset_unit"][:, 5:].mean()) - 0.5) < 0.02
    reset_ids, speed_ids = _hit_sets(orc)
    env.post_physics_step()                        # in-kernel Philox
    orc.post_physics_step(nz["sensor_unit"], nz["reset_unit"][reset_ids], nz["speed_unit"][speed_ids], nz["yaw_unit"])
    orc.timeout_buf = (orc.progress_buf >= orc.max_episode_length - 1) & (orc.reset_buf != 0)
    _compare_step(env, orc, reset_ids, "turning philox")


def test_sharded_equals_unsharded_with_summed_command_sums():
    """reward5 couples every env: two half-size tasks whose command sums are added (what the all-reduce does) must give
    the rewards of the unsharded task."""
    from thormang_isaacgym_b200 import _native as nat
    n = 2048
    st = syn.gogoro_paper_state(n, seed=61)
    st["buffer_obs"][:, :, 6] += 0.11
    nz = syn.gogoro_paper_noise(n, seed=62, reset_cols=COLS)
    full, _, _ = U.make_gogoro_paper_pair(n, seed=61, state=st, turning=True)
    full.post_physics_step(nz["sensor_unit"].to(DEV), nz["reset_unit"].to(DEV), nz["speed_unit"].to(DEV), nz["yaw_unit"].to(DEV))
    halves, sums = [], torch.zeros(2, device=DEV, dtype=torch.float64)
    for r in range(2):
        sl = slice(r * n // 2, (r + 1) * n // 2)
        sub = {k: v[sl].clone() if v.shape[0] == n else v[sl.start * 39:sl.stop * 39].clone() for k, v in st.items()}
        env, _, _ = U.make_gogoro_paper_pair(n // 2, seed=61, state=sub, turning=True)
        keep = []
        dz = env._noise(nz["sensor_unit"][sl], nz["reset_unit"][sl], nz["speed_unit"][sl], nz["yaw_unit"][sl], keep)
        nat.check(env._lib.tk_gogoro_paper_post_physics(env._handle, dz, env._stream()))
        nat.check(env._lib.tk_gogoro_paper_diff_sums(env._handle, nat.ptr(env.diff_sums), env._stream()))
        sums += env.diff_sums
        halves.append(env)
    for env in halves:
        nat.check(env._lib.tk_gogoro_paper_finalize_reward(env._handle, nat.ptr(sums), env._stream()))
    torch.cuda.synchronize()
    assert float(sums[1]) == n * 20
    got = torch.cat([h.rew_buf for h in halves])
    U.close(got, full.rew_buf, rtol=0, atol=2e-5, what="sharded reward")       # 10 x one fp32 rounding of the global mean
    U.exact(torch.cat([h.obs_buf for h in halves]), full.obs_buf, "sharded obs")
    U.exact(torch.cat([h.reset_buf for h in halves]), full.reset_buf, "sharded reset")


def test_cuda_graph_replay_matches_eager():
    n = 512
    st = syn.gogoro_paper_state(n, seed=81)
    nz = {k: v.to(DEV) for k, v in syn.gogoro_paper_noise(n, seed=82, reset_cols=COLS).items()}
    act = st["actions"].to(DEV)
    eager, _, _ = U.make_gogoro_paper_pair(n, seed=81, state=st, turning=True)
    graph_env, _, _ = U.make_gogoro_paper_pair(n, seed=81, state=st, turning=True)
    steps = 23

    def one(env):
        env.pre_physics_step(act)
        env.post_physics_step(nz["sensor_unit"], nz["reset_unit"], nz["speed_unit"], nz["yaw_unit"])

    for _ in range(steps):
        one(eager)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        one(graph_env)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            one(graph_env)
        for _ in range(steps - 1):
            g.replay()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    U.exact(graph_env.obs_buf, eager.obs_buf, "graph obs_buf")
    U.exact(graph_env.rew_buf, eager.rew_buf, "graph rew_buf")
    U.exact(graph_env.buffer_obs, eager.buffer_obs, "graph clean frames")
    U.exact(graph_env.progress_buf, eager.progress_buf, "graph progress")


def test_bad_variant_is_rejected():
    import ctypes as C
    from thormang_isaacgym_b200 import _native as nat
    lib = nat.lib()
    p = nat.GogoroPaperParams()
    p.device, p.num_envs, p.num_dof, p.steering_joint, p.rear_wheel_joint = 0, 64, 39, 0, 1
    p.max_episode_length, p.buff_size, p.delay_width, p.variant = 3600, 20, 5, 7
    p.command_delay[0], p.command_delay[1] = 0.0, 5.0
    h = C.c_void_p()
    assert lib.tk_gogoro_paper_create(C.byref(p), C.byref(h)) == -1
    assert b"TK_GOGORO_PAPER_VARIANT" in lib.tk_last_error()


def test_supplied_draws_of_the_wrong_width_are_refused():
    """The TURNING variant reads 8 reset draws per row: a [N,5] tensor (the paper variant's shape) must be refused on the
    host instead of being read out of bounds on the device."""
    n = 64
    env, _, _ = U.make_gogoro_paper_pair(n, seed=5, turning=True)
    nz = syn.gogoro_paper_noise(n, seed=6)                      # reset_cols = 5
    with pytest.raises(ValueError, match="reset_unit"):
        env.post_physics_step(nz["sensor_unit"].to(DEV), nz["reset_unit"].to(DEV), nz["speed_unit"].to(DEV), nz["yaw_unit"].to(DEV))
    with pytest.raises(ValueError, match="sensor_unit"):
        env.post_physics_step(nz["sensor_unit"][:, :5].to(DEV), None, None, None)
    with pytest.raises(ValueError, match="unit"):
        env.reset_idx(torch.arange(4, device=DEV), torch.rand(4, 5))
