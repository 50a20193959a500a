"""GPU parity: Gogoro scooter CUDA path (through the C ABI) vs the CPU oracle on identical synthetic state."""
import os

import numpy as np
import pytest
import torch

import parity_util as U
from thormang_isaacgym_b200 import synthetic as syn

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("n", [1, 31, 33, 255, 256, 257, 4096, 65536])
def test_observations_and_reward(n):
    from oracle import gogoro_oracle as go
    env, orc, st = U.make_gogoro_pair(n, seed=300 + n)
    got = env.compute_observations()
    want = go.gogoro_observations(orc.root_tensor.clone(), orc.yaw_command.clone(), orc.curent_command.clone())
    U.gogoro_obs_close(got, want)
    # reward on identical observations: masks exact, reward within tolerance
    env.compute_reward(want.to(DEV))
    rew, reset = go.gogoro_reward(want.view(n, 1, 6), orc.progress_buf, orc.action_history, 1000)
    U.exact(env.reset_buf, reset, "reset_buf")
    U.close(env.rew_buf, rew, what="rew_buf")


def test_golden_vectors(golden_dir):
    z = np.load(os.path.join(golden_dir, "gogoro_jit.npz"))
    g = {k: torch.from_numpy(z[k]) for k in z.files}
    n = g["root_tensor"].shape[0]
    st = syn.gogoro_state(n, seed=1)
    st.update(root_tensor=g["root_tensor"], yaw_command=g["desired_yaw"], curent_command=g["last_command"],
              action_history=g["action_history"], progress_buf=g["progress_buf"])
    env, _, _ = U.make_gogoro_pair(n, state=st)
    U.gogoro_obs_close(env.compute_observations(), g["obs"], "obs vs golden")
    env.compute_reward(g["obs"].to(DEV))
    U.exact(env.reset_buf, g["reset"], "reset vs golden")
    U.close(env.rew_buf, g["rew"], what="rew vs golden")


def _compare_state(env, orc, quat_atol=2e-6):
    for name in ("curent_speed", "curent_speed_offset", "steer_offsets", "imu_offsets", "config_vector",
                 "curent_command", "action_history", "progress_buf", "reset_buf"):
        U.exact(getattr(env, name), getattr(orc, name), name)
    U.exact(env.state_dof, orc.state_dof, "state_dof")
    # the spawn quaternion goes through sin/cos (euler_to_quaternion): few-ulp libm difference, everything else exact
    cols = [0, 1, 2, 7, 8, 9, 10, 11, 12]
    U.exact(env.root_tensor[:, cols], orc.root_tensor[:, cols], "root_tensor (non-quaternion columns)")
    U.close(env.root_tensor[:, 3:7], orc.root_tensor[:, 3:7], atol=quat_atol, what="root quaternion")
    U.close(env.yaw_command, orc.yaw_command, atol=1e-6, what="yaw_command")


def test_pre_physics():
    n = 3000
    env, orc, st = U.make_gogoro_pair(n, seed=12)
    nz = syn.gogoro_noise(n, seed=5)
    env.pre_physics_step(st["actions"].to(DEV), nz["steer_unit"].to(DEV))
    orc.pre_physics_step(st["actions"], nz["steer_unit"])
    U.exact(env.action_history, orc.action_history, "action_history")
    U.exact(env.curent_command, orc.curent_command, "curent_command")
    U.exact(env.dof_pos_command, orc.dof_pos_command, "dof_pos_command")
    U.exact(env.dof_speed_command, orc.dof_speed_command, "dof_speed_command")
    assert env.sim.dof_position_target is env.dof_pos_command
    # full-rewrite mode (fresh zeros like gogoro_new.py:360,366)
    env.dof_pos_command.fill_(7.0); env.dof_speed_command.fill_(7.0)
    env.zero_fill_targets = True
    env.pre_physics_step(st["actions"].to(DEV), nz["steer_unit"].to(DEV))
    orc.pre_physics_step(st["actions"], nz["steer_unit"])
    U.exact(env.dof_pos_command, orc.dof_pos_command, "dof_pos_command (zero fill)")
    U.exact(env.dof_speed_command, orc.dof_speed_command, "dof_speed_command (zero fill)")


def test_reset_idx():
    n = 700
    env, orc, st = U.make_gogoro_pair(n, seed=13)
    ids = torch.tensor([0, 1, 31, 32, 255, 256, 699])
    nz = syn.gogoro_noise(n, seed=6)
    unit = nz["reset_unit"][:len(ids)]
    env.reset_idx(ids.to(DEV), unit.to(DEV))
    orc.reset_idx(ids, unit)
    _compare_state(env, orc)
    U.exact(env.buffer_obs, orc.buffer_obs, "buffer_obs")
    U.exact(env.obs_buf, orc.obs_buf, "obs_buf")


@pytest.mark.parametrize("n", [1, 257, 4096, 65536])
@pytest.mark.parametrize("mode", ["compact", "philox"])
def test_fused_post_physics(n, mode):
    env, orc, st = U.make_gogoro_pair(n, seed=40 + n)
    if n > 300:
        for i in (0, 31, 32, 255, 256, n - 1):
            st["reset_buf"][i] = 1
        st["progress_buf"][5] = 299; st["reset_buf"][5] = 0
        st["progress_buf"][256] = 299; st["reset_buf"][256] = 1   # would hit 300 but is reset first
        env.reset_buf.copy_(st["reset_buf"].to(DEV)); orc.reset_buf = st["reset_buf"].clone()
        env.progress_buf.copy_(st["progress_buf"].to(DEV)); orc.progress_buf = st["progress_buf"].clone()
    nz = syn.gogoro_noise(n, seed=8)
    ids = orc.reset_buf.nonzero().flatten()
    speed_ids = ((orc.reset_buf == 0) & (orc.progress_buf + 1 == 300)).nonzero().flatten()
    if mode == "compact":
        sensor, reset_unit, speed_unit, yaw_unit = nz["sensor_unit"], nz["reset_unit"][:max(len(ids), 1)], nz["speed_unit"][:max(len(speed_ids), 1)], nz["yaw_unit"]
        env.post_physics_step(sensor.to(DEV), reset_unit.to(DEV), speed_unit.to(DEV), yaw_unit.to(DEV))
    else:
        pn = {k: v.cpu() for k, v in env.philox_noise().items()}
        assert abs(float(pn["sensor_unit"].mean())) < 0.1 and abs(float(pn["sensor_unit"].std()) - 1.0) < 0.1 or n < 1000
        sensor, reset_unit, speed_unit, yaw_unit = pn["sensor_unit"], pn["reset_unit"][ids], pn["speed_unit"][speed_ids], pn["yaw_unit"]
        env.post_physics_step()
    orc.post_physics_step(sensor, reset_unit, speed_unit, yaw_unit)
    orc.timeout_buf = (orc.progress_buf >= 999) & (orc.reset_buf != 0)
    # reset rows start from a sin/cos-built quaternion: compare those obs with a looser angle tolerance
    U.gogoro_obs_close(env.buffer_obs[:, 0], orc.buffer_obs[:, 0], "buffer_obs")
    U.gogoro_obs_close(env.obs_buf, orc.obs_buf, "obs_buf (noisy)", noisy=True)
    U.close(env.rew_buf, orc.rew_buf, rtol=2e-5, atol=2e-5, what="rew_buf")
    U.exact(env.timeout_buf, orc.timeout_buf, "timeout_buf")
    _compare_state(env, orc)
    stats = env.episode_statistics().cpu()
    want = torch.tensor([orc.rew_buf.double().sum(), (orc.reset_buf != 0).sum(), orc.timeout_buf.sum(), 0.0,
                         orc.progress_buf[orc.reset_buf != 0].sum(), orc.buffer_obs[:, 0, 0].abs().double().sum(),
                         orc.buffer_obs[:, 0, 4].abs().double().sum(), n], dtype=torch.float64)
    keep = [0, 1, 2, 4, 5, 6, 7]
    U.close(stats[keep], want[keep], rtol=1e-5, atol=1e-2, what="episode statistics")


def test_compute_obs_rwd_method():
    n = 1025
    env, orc, st = U.make_gogoro_pair(n, seed=17)
    nz = syn.gogoro_noise(n, seed=9)
    env.compute_obs_rwd(nz["sensor_unit"].to(DEV))
    orc.compute_obs_rwd(nz["sensor_unit"])
    U.gogoro_obs_close(env.buffer_obs[:, 0], orc.buffer_obs[:, 0], "buffer_obs")
    U.gogoro_obs_close(env.obs_buf, orc.obs_buf, "obs_buf", noisy=True)
    U.exact(env.reset_buf, orc.reset_buf, "reset_buf")
    U.exact(env.progress_buf, orc.progress_buf, "progress_buf (untouched)")
    U.close(env.rew_buf, orc.rew_buf, rtol=2e-5, atol=2e-5, what="rew_buf")


def test_rollout():
    n, horizon = 1024, 10
    env, orc, st = U.make_gogoro_pair(n, seed=23)
    g = torch.Generator().manual_seed(3)

    def physics_t(root, dt):
        root[:, 0:3] += dt * root[:, 7:10]
        root[:, 10:13] *= 0.99

    for t in range(horizon):
        a = torch.rand(n, 1, generator=g) * 3 - 1.5
        nz = syn.gogoro_noise(n, seed=100 + t)
        ids = orc.reset_buf.nonzero().flatten()
        env.pre_physics_step(a.to(DEV), nz["steer_unit"].to(DEV))
        physics_t(env.root_tensor, 0.03)
        env.post_physics_step(nz["sensor_unit"].to(DEV), nz["reset_unit"].to(DEV), nz["speed_unit"].to(DEV), nz["yaw_unit"].to(DEV))
        orc.step(a, nz["steer_unit"], nz["sensor_unit"], nz["reset_unit"], nz["speed_unit"], nz["yaw_unit"],
                 physics=lambda o: physics_t(o.root_tensor, 0.03))
        U.exact(env.reset_buf, orc.reset_buf, f"reset t={t}")
        U.exact(env.progress_buf, orc.progress_buf, f"progress t={t}")
        U.gogoro_obs_close(env.obs_buf, orc.obs_buf, f"obs t={t}", noisy=True)
        U.close(env.rew_buf, orc.rew_buf, rtol=2e-5, atol=2e-5, what=f"rew t={t}")
        orc.root_tensor.copy_(env.root_tensor.cpu()); orc.yaw_command.copy_(env.yaw_command.cpu())


def test_sharded_equals_unsharded():
    n = 65536
    st = syn.gogoro_state(n, seed=1234)
    env, _, _ = U.make_gogoro_pair(n, state={k: v.clone() for k, v in st.items()})
    env.post_physics_step()
    half = n // 2
    for r in range(2):
        sl = slice(r * half, (r + 1) * half)
        sub = {k: (v[sl.start * 39:sl.stop * 39].clone() if k == "state_dof" else v[sl].clone()) for k, v in st.items()}
        e2, _, _ = U.make_gogoro_pair(half, state=sub, envIdOffset=r * half)
        e2.noise_seed = env.noise_seed
        e2.post_physics_step()
        for k in ("obs_buf", "buffer_obs", "rew_buf", "reset_buf", "progress_buf", "yaw_command", "curent_speed", "root_tensor"):
            U.exact(getattr(e2, k), getattr(env, k)[sl], f"shard {r} {k}")


def test_cuda_graph_replay_matches_eager():
    n = 4096
    env_a, _, st = U.make_gogoro_pair(n, seed=61)
    env_b, _, _ = U.make_gogoro_pair(n, seed=61, state={k: v.clone() for k, v in st.items()})
    a = st["actions"].to(DEV)

    def step(env, key=None):
        if key is not None:
            env.step_count = key
        env.pre_physics_step(a)
        env.post_physics_step()

    for r in range(2):
        for i in range(2):
            step(env_a, key=i + ((r + 1) << 32))    # captured step i of replay r: host step i, device epoch r + 1
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        env_b.advance_noise_epoch()
        step(env_b)
        step(env_b)
    g.replay(); g.replay()
    torch.cuda.synchronize()
    for name in ("obs_buf", "buffer_obs", "rew_buf", "reset_buf", "progress_buf", "root_tensor", "curent_command",
                 "action_history", "yaw_command", "curent_speed", "dof_pos_command"):
        U.exact(getattr(env_b, name), getattr(env_a, name), f"graph vs eager: {name}")


@pytest.mark.parametrize("n", [4099, 65536])
def test_speculative_reset_matches_lazy_reset(n, monkeypatch):
    """Thin grids draw every env's reset speculatively ahead of the dependency wait (gogoro_tile_kernel<0, true>); deep grids
    draw lazily, warp-cooperatively, for the flagged envs only.  Same Philox streams, same expressions: every state tensor
    must agree bit for bit after three steps with resets (also more than two per warp, and a ragged last tile)."""
    st = syn.gogoro_state(n, seed=77)
    for i in (0, 1, 2, 3, 31, 32, 127, 128, n - 1):
        st["reset_buf"][i] = 1
    outs = []
    for spec in ("0", "1"):
        monkeypatch.setenv("TK_GOGORO_SPEC", spec)          # read when the handle is created
        env, _, _ = U.make_gogoro_pair(n, state={k: v.clone() for k, v in st.items()})
        for _ in range(3):
            env.post_physics_step()
        torch.cuda.synchronize()
        outs.append({k: getattr(env, k).clone() for k in (
            "obs_buf", "buffer_obs", "rew_buf", "reset_buf", "progress_buf", "yaw_command", "curent_command", "curent_speed",
            "curent_speed_offset", "steer_offsets", "imu_offsets", "config_vector", "action_history", "root_tensor", "state_dof")})
    for k in outs[0]:
        assert torch.equal(outs[0][k], outs[1][k]), f"{k}: speculative reset differs from the lazy reset"


def test_config5_ppo_rollout_shape_with_in_kernel_noise():
    """The rl_games PPO shape of cfg/train/GogoroPPO.yaml:47,61 -- 4096 actors x horizon 64 -- with ALL noise drawn in the
    kernels (Philox: steering noise in pre_physics, sensor / reset / resample draws in the fused post_physics, speculative
    reset variant on this thin grid).  The draws of every step are materialised with philox_noise() and handed to the oracle,
    which steps alongside; command resampling (progress hitting the update frequencies) and resets occur on the way."""
    n, horizon = 4096, 64
    env, orc, st = U.make_gogoro_pair(n, seed=64)
    env.progress_buf[::7] = 280                      # some envs reach speed_freq_update (300) inside the horizon
    orc.progress_buf[::7] = 280
    g = torch.Generator().manual_seed(11)

    def physics_t(root, dt):
        root[:, 0:3] += dt * root[:, 7:10]
        root[:, 10:13] *= 0.99

    resets = resamples = 0
    for t in range(horizon):
        a = torch.rand(n, 1, generator=g) * 3 - 1.5
        pn = {k: v.cpu() for k, v in env.philox_noise().items()}
        ids = orc.reset_buf.nonzero().flatten()
        speed_ids = ((orc.reset_buf == 0) & (orc.progress_buf + 1 == 300)).nonzero().flatten()
        resets += len(ids); resamples += len(speed_ids)
        env.pre_physics_step(a.to(DEV))
        physics_t(env.root_tensor, 0.03)
        env.post_physics_step()
        orc.step(a, pn["steer_unit"], pn["sensor_unit"], pn["reset_unit"][ids] if len(ids) else pn["reset_unit"][:1],
                 pn["speed_unit"][speed_ids] if len(speed_ids) else pn["speed_unit"][:1], pn["yaw_unit"],
                 physics=lambda o: physics_t(o.root_tensor, 0.03))
        U.exact(env.reset_buf, orc.reset_buf, f"reset t={t}")
        U.exact(env.progress_buf, orc.progress_buf, f"progress t={t}")
        U.gogoro_obs_close(env.obs_buf, orc.obs_buf, f"obs t={t}", noisy=True)
        U.close(env.rew_buf, orc.rew_buf, rtol=2e-5, atol=2e-5, what=f"rew t={t}")
        U.close(env.curent_speed, orc.curent_speed, what=f"curent_speed t={t}")
        orc.root_tensor.copy_(env.root_tensor.cpu()); orc.yaw_command.copy_(env.yaw_command.cpu())
    assert resets > 0 and resamples > 0


@pytest.mark.parametrize("cuts", [(1,), (333,), (127, 128, 700), (512, 513)])
def test_ragged_shards_with_global_env_ids(cuts):
    """Env-range sharding at arbitrary cut points (offsets that are no multiple of the 128-env tile): all in-kernel draws are
    keyed by the global env id, so the shards reproduce the unsharded step bit for bit (resets and resampling included)."""
    n = 1000
    st = syn.gogoro_state(n, seed=78)
    st["reset_buf"][::9] = 1
    st["progress_buf"][5::11] = 299
    env, _, _ = U.make_gogoro_pair(n, state={k: v.clone() for k, v in st.items()})
    env.post_physics_step()
    keys = ("obs_buf", "buffer_obs", "rew_buf", "reset_buf", "progress_buf", "yaw_command", "curent_speed", "curent_command",
            "imu_offsets", "steer_offsets", "action_history", "root_tensor")
    full = {k: getattr(env, k).clone() for k in keys + ("state_dof",)}
    edges = (0,) + tuple(cuts) + (n,)
    for lo, hi in zip(edges[:-1], edges[1:]):
        sub = {k: (v[lo * 39:hi * 39].clone() if k == "state_dof" else v[lo:hi].clone()) for k, v in st.items()}
        e2, _, _ = U.make_gogoro_pair(hi - lo, state=sub, envIdOffset=lo)
        e2.noise_seed = env.noise_seed
        e2.post_physics_step()
        for k in keys:
            assert torch.equal(getattr(e2, k), full[k][lo:hi]), f"shard [{lo},{hi}) {k}"
        assert torch.equal(e2.state_dof, full["state_dof"][lo * 39:hi * 39]), f"shard [{lo},{hi}) state_dof"


def test_cuda_graph_epoch_directly_in_front_of_the_step_many_replays():
    """The fused kernel reads the device-side noise epoch AHEAD of its dependency wait.  Worst case for that: a graph whose
    epoch kernel sits directly in front of ONE post_physics launch, replayed 40 times back to back -- every replay must use
    the epoch bumped by its own graph (bit-identical to eager launches keyed (epoch << 32) | step), never a stale one."""
    n, replays = 4096, 40
    env_a, _, st = U.make_gogoro_pair(n, seed=62)
    env_b, _, _ = U.make_gogoro_pair(n, seed=62, state={k: v.clone() for k, v in st.items()})
    for e in (env_a, env_b):
        e.reset_buf[::5] = 1                       # plenty of resets: the reset draws depend on the epoch too
    for r in range(replays):
        env_a.step_count = (r + 1) << 32           # captured host step 0, device epoch r + 1
        env_a.post_physics_step()
        env_a.reset_buf[::5] = 1
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    env_b.step_count = 0
    with torch.cuda.graph(g):
        env_b.advance_noise_epoch()
        env_b.post_physics_step()
        env_b.reset_buf[::5] = 1
    for _ in range(replays):
        g.replay()
    torch.cuda.synchronize()
    for name in ("obs_buf", "buffer_obs", "rew_buf", "progress_buf", "root_tensor", "yaw_command", "curent_speed",
                 "imu_offsets", "steer_offsets", "config_vector"):
        U.exact(getattr(env_b, name), getattr(env_a, name), f"graph vs eager after {replays} replays: {name}")


def test_cuda_graph_epoch_bump_recorded_after_the_steps():
    """A graph [step, step, advance_epoch] on a handle whose epoch was never bumped before the capture: the captured
    launches must still read the device-side epoch word (a NULL baked in at capture time would make every replay draw
    the noise of replay 0).  Replay r runs host steps 0 and 1 under epoch r."""
    n, replays = 2048, 6
    env_a, _, st = U.make_gogoro_pair(n, seed=63)
    env_b, _, _ = U.make_gogoro_pair(n, seed=63, state={k: v.clone() for k, v in st.items()})
    for e in (env_a, env_b):
        e.reset_buf[::5] = 1
    for r in range(replays):
        for host_step in (0, 1):
            env_a.step_count = (r << 32) | host_step
            env_a.post_physics_step()
            env_a.reset_buf[::5] = 1
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for host_step in (0, 1):
            env_b.step_count = host_step
            env_b.post_physics_step()
            env_b.reset_buf[::5] = 1
        env_b.advance_noise_epoch()
    for _ in range(replays):
        g.replay()
    torch.cuda.synchronize()
    for name in ("obs_buf", "buffer_obs", "rew_buf", "progress_buf", "root_tensor", "yaw_command", "curent_speed",
                 "imu_offsets", "steer_offsets", "config_vector"):
        U.exact(getattr(env_b, name), getattr(env_a, name), f"[step, step, bump] graph vs eager after {replays} replays: {name}")
