"""GPU parity: Humanoid CUDA path (through the C ABI) vs the CPU oracle on identical synthetic state."""
import os

import numpy as np
import pytest
import torch

import parity_util as U
from thormang_isaacgym_b200 import synthetic as syn

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SIZES = [1, 31, 32, 33, 127, 128, 129, 1000, 4096]


def _check_obs_outputs(env, orc):
    U.humanoid_obs_close(env.obs_buf, orc.obs_buf)
    U.exact(env.obs_buf[:, 0], orc.obs_buf[:, 0], "obs[:,0] (height, raw copy)")
    U.exact(env.obs_buf[:, 87:108], orc.obs_buf[:, 87:108], "obs[:,87:108] (actions, raw copy)")
    U.close(env.potentials, orc.potentials, atol=0.0, what="potentials")
    U.close(env.prev_potentials, orc.prev_potentials, atol=0.0, what="prev_potentials")
    U.close(env.up_vec, orc.up_vec, what="up_vec")
    U.close(env.heading_vec, orc.heading_vec, what="heading_vec")


@pytest.mark.parametrize("n", SIZES)
def test_observations_and_reward_split(n):
    env, orc, _ = U.make_humanoid_pair(n, seed=100 + n)
    env.compute_observations()
    orc.compute_observations()
    _check_obs_outputs(env, orc)
    env.compute_reward(env.actions)
    orc.compute_reward()
    U.exact(env.reset_buf, orc.reset_buf, "reset_buf")
    U.close(env.rew_buf, orc.rew_buf, what="rew_buf")
    # the reward depends on obs only through two thresholds; feeding the oracle's obs must give identical bits
    env.obs_buf.copy_(orc.obs_buf.to(DEV))
    env.reset_buf.copy_(torch.zeros_like(env.reset_buf))
    orc.reset_buf = torch.zeros_like(orc.reset_buf)
    env.compute_reward(env.actions)
    orc.compute_reward()
    U.exact(env.rew_buf, orc.rew_buf, "rew_buf on identical obs")
    U.exact(env.reset_buf, orc.reset_buf, "reset_buf on identical obs")


def test_golden_vectors(golden_dir):
    """CUDA path against the vectors frozen from the reference's own JIT functions (incl. the edge rows)."""
    z = np.load(os.path.join(golden_dir, "humanoid_jit.npz"))
    g = {k: torch.from_numpy(z[k]) for k in z.files}
    n = g["root_states"].shape[0]
    asset = dict(dof_limits_lower=g["dof_limits_lower"], dof_limits_upper=g["dof_limits_upper"], motor_efforts=g["motor_efforts"])
    state = dict(root_states=g["root_states"], dof_state=g["dof_state"], dof_force=g["dof_force"], sensors=g["sensors"],
                 actions=g["actions"], progress_buf=g["progress_buf"], reset_buf=g["reset_buf"], potentials=g["potentials"])
    env, _, _ = U.make_humanoid_pair(n, state=state, asset=asset)
    env.compute_observations()
    U.humanoid_obs_close(env.obs_buf, g["obs"], "obs vs golden")
    U.close(env.up_vec, g["up_vec"], what="up_vec vs golden")
    U.close(env.heading_vec, g["heading_vec"], what="heading_vec vs golden")
    U.close(env.potentials, g["potentials_out"], atol=0.0, what="potentials vs golden")
    env.compute_reward(env.actions)
    U.exact(env.reset_buf, g["reset_out"], "reset vs golden")
    U.close(env.rew_buf, g["rew"], what="rew vs golden")
    wm = env.wm_reward(g["obs"].view(2, n // 2, 110))
    U.close(wm, g["wm_reward"], atol=1e-4, what="wm reward vs golden")   # potentials difference of ~6e4-sized terms / 6e4


@pytest.mark.parametrize("n", [1, 33, 128, 257, 4096, 16384, 131072])   # 16384 / 131072: BASELINE configs 2 and 4
@pytest.mark.parametrize("mode", ["compact", "dense", "philox"])
def test_fused_post_physics(n, mode):
    from thormang_isaacgym_b200 import _native as nat
    env, orc, st = U.make_humanoid_pair(n, seed=7 + n)
    if n >= 33:   # make sure resets exist and straddle warp/tile edges
        for i in (0, 31, 32, n - 1):
            st["reset_buf"][i] = 1
        env.reset_buf.copy_(st["reset_buf"].to(DEV)); orc.reset_buf = st["reset_buf"].clone()
    ids = orc.reset_buf.nonzero().flatten()
    k = len(ids)
    ua, ub = syn.humanoid_reset_noise(max(k, 1), seed=n)
    if mode == "compact":
        env.post_physics_step(ua.to(DEV), ub.to(DEV))
    elif mode == "dense":
        da, db = torch.zeros(n, 21), torch.zeros(n, 21)
        da[ids], db[ids] = ua[:k], ub[:k]
        env.post_physics_step(da.to(DEV), db.to(DEV), noise_mode=nat.NOISE_DENSE)
    else:
        pa, pb = env.philox_noise()
        assert 0.0 <= float(pa.min()) and float(pa.max()) < 1.0 and abs(float(pa.mean()) - 0.5) < 0.2
        ua, ub = pa.cpu()[ids], pb.cpu()[ids]
        env.post_physics_step()
    orc.post_physics_step(ua, ub)
    orc.timeout_buf = (orc.progress_buf >= 999) & (orc.reset_buf != 0)
    _check_obs_outputs(env, orc)
    U.exact(env.reset_buf, orc.reset_buf, "reset_buf")
    U.exact(env.progress_buf, orc.progress_buf, "progress_buf")
    U.exact(env.randomize_buf, orc.randomize_buf, "randomize_buf")
    U.exact(env.timeout_buf, orc.timeout_buf, "timeout_buf")
    U.close(env.rew_buf, orc.rew_buf, what="rew_buf")
    # reset-state writes: exact (add / mul / clamp / sqrt / div only)
    U.exact(env.dof_state, orc.dof_state, "dof_state after reset")
    U.exact(env.root_states, orc.root_states, "root_states after reset")
    # episode statistics vs the single-process oracle sums
    stats = env.episode_statistics().cpu()
    want = torch.tensor([orc.rew_buf.double().sum(), (orc.reset_buf != 0).sum(), orc.timeout_buf.sum(), 0.0,
                         orc.progress_buf[orc.reset_buf != 0].sum(), orc.obs_buf[:, 10].double().sum(),
                         orc.obs_buf[:, 11].double().sum(), n], dtype=torch.float64)
    keep = [0, 1, 2, 4, 5, 6, 7]
    U.close(stats[keep], want[keep], rtol=1e-5, atol=1e-3, what="episode statistics")


@pytest.mark.parametrize("mode", ["compact", "philox"])
def test_fused_every_env_of_a_tile_resets(mode):
    """All 32 envs of tile 1 (and a few of tiles 0 and 3) reset in one step: the kernel writes reset rows to global memory
    while the bulk load of the same rows is in flight and patches the staged copy afterwards (tk_humanoid_tile.cuh); the
    result must be the reference's reset -> observe order for every one of them, noise row k <-> k-th reset env."""
    n = 160
    env, orc, st = U.make_humanoid_pair(n, seed=404)
    st["reset_buf"][:] = 0
    st["reset_buf"][32:64] = 1
    for i in (0, 5, 31, 96, 97, 98, 99, 100, 127):
        st["reset_buf"][i] = 1
    env.reset_buf.copy_(st["reset_buf"].to(DEV)); orc.reset_buf = st["reset_buf"].clone()
    ids = orc.reset_buf.nonzero().flatten()
    if mode == "compact":
        ua, ub = syn.humanoid_reset_noise(len(ids), seed=11)
        env.post_physics_step(ua.to(DEV), ub.to(DEV))
    else:
        pa, pb = env.philox_noise()
        ua, ub = pa.cpu()[ids], pb.cpu()[ids]
        env.post_physics_step()
    orc.post_physics_step(ua, ub)
    _check_obs_outputs(env, orc)
    for name in ("reset_buf", "progress_buf", "dof_state", "root_states"):
        U.exact(getattr(env, name), getattr(orc, name), name)
    U.close(env.rew_buf, orc.rew_buf, what="rew_buf")
    assert int((env.progress_buf[32:64] == 0).sum()) == 32


def test_angle_to_target_beyond_pi():
    """angle_to_target = atan2(..) - yaw spans (-3*pi, pi] before normalize_angle (humanoid.py:400-405).  With the target
    BEHIND the robots (targets = (-1000, 0, 0), so atan2 ~ +-pi) every row has |angle_to_target| in (pi, 3*pi): the
    kernel's +-2*pi wrap must land in [-pi, pi] and agree with atan2(sin, cos) -- checked with the range assertion on."""
    n = 2048
    env, orc, st = U.make_humanoid_pair(n, seed=31)
    # pos.z above / below the target height so that atan2 sits on both sides of the cut, yaw uniform in [0, 2*pi)
    env.targets[:] = torch.tensor([-1000.0, 0.0, 1.0], device=DEV)
    orc.targets[:] = torch.tensor([-1000.0, 0.0, 1.0])
    env._lib.tk_humanoid_destroy(env._handle)
    env._create_native()
    env.compute_observations()
    orc.compute_observations()
    from oracle import isaac_helpers as ih
    q = ih.quat_mul(orc.root_states[:, 3:7], orc.inv_start_rot)
    _, _, yaw = ih.get_euler_xyz(q)
    raw = torch.atan2(orc.targets[:, 2] - orc.root_states[:, 2], orc.targets[:, 0] - orc.root_states[:, 0]) - yaw
    frac = float(((raw.abs() > np.pi) & (raw.abs() < 3 * np.pi)).float().mean())
    assert frac > 0.45, f"only {frac:.2f} of the rows exercise the wide range"
    _check_obs_outputs(env, orc)
    assert float(env.obs_buf[:, 9].abs().max()) <= U.PI_F32 * (1 + 2.0 ** -22)


def test_pre_physics_and_reset_idx():
    n = 1000
    env, orc, st = U.make_humanoid_pair(n, seed=5)
    a = (torch.rand(n, 21, generator=torch.Generator().manual_seed(1)) * 3 - 1.5)
    env.pre_physics_step(a.to(DEV))                 # direct call: no clamp (humanoid.py:281-285)
    orc.pre_physics_step(a, clip=False)
    U.exact(env.actions, orc.actions, "actions"); U.exact(env.forces, orc.forces, "forces")
    env._clip_in_kernel = True
    env.pre_physics_step(a.to(DEV))                 # as called from VecTask.step: clamp +-clipActions fused in
    env._clip_in_kernel = False
    orc.pre_physics_step(a, clip=True)
    U.exact(env.actions, orc.actions, "clamped actions"); U.exact(env.forces, orc.forces, "clamped forces")
    assert env.sim.actuation_force is env.forces
    ids = torch.tensor([0, 5, 31, 32, 33, 500, 999])
    ua, ub = syn.humanoid_reset_noise(len(ids), seed=3)
    env.reset_idx(ids.to(DEV), ua.to(DEV), ub.to(DEV))
    orc.reset_idx(ids, ua, ub)
    for name in ("dof_state", "root_states", "potentials", "prev_potentials", "progress_buf", "reset_buf"):
        U.exact(getattr(env, name), getattr(orc, name), name)
    # default noise path = the two torch.rand draws of torch_rand_float, reproducible from the torch seed
    torch.manual_seed(11); env.reset_idx(ids.to(DEV)); s1 = env.dof_state.clone()
    torch.manual_seed(11); ua2 = torch.rand(len(ids), 21, device=DEV); ub2 = torch.rand(len(ids), 21, device=DEV)
    orc.reset_idx(ids, ua2.cpu(), ub2.cpu())
    U.exact(s1, orc.dof_state, "dof_state, torch-seeded noise")


def test_reward_terms():
    n = 513
    env, orc, _ = U.make_humanoid_pair(n, seed=9)
    env.compute_observations(); orc.compute_observations()
    env.obs_buf.copy_(orc.obs_buf.to(DEV)); env.potentials.copy_(orc.potentials.to(DEV)); env.prev_potentials.copy_(orc.prev_potentials.to(DEV))
    from oracle import humanoid_oracle as ho
    t = ho.humanoid_reward_terms(orc.obs_buf, orc.actions, 0.1, 0.5, orc.potentials, orc.prev_potentials, 0.25,
                                 orc.max_motor_effort, orc.motor_efforts)
    got = env.reward_terms().cpu()
    for i, key in enumerate(["heading_reward", "up_reward", "actions_cost", "dof_at_limit_cost", "electricity_cost",
                             "alive_reward", "progress_reward"]):
        U.close(got[:, i], t[key], rtol=1e-5, atol=1e-5, what=key)
    assert float(t["dof_at_limit_cost"].max()) > 0, "synthetic state must exercise the joints-at-limit branch"


def test_rollout_with_fake_physics():
    """pre -> fake physics -> post for a PPO-rollout-shaped loop (BASELINE config 5), trajectories compared each step."""
    n, horizon = 512, 12
    env, orc, st = U.make_humanoid_pair(n, seed=21)

    def physics_t(root, dof_state, forces, dt):
        root[:, 0:3] += dt * root[:, 7:10]
        root[:, 2] -= 0.02
        dof_state[:, 0] += dt * dof_state[:, 1]
        dof_state[:, 1] += 0.001 * forces.reshape(-1)

    g = torch.Generator().manual_seed(5)
    for t in range(horizon):
        a = torch.rand(n, 21, generator=g) * 2.4 - 1.2
        k = int((orc.reset_buf != 0).sum())
        ua, ub = syn.humanoid_reset_noise(max(k, 1), seed=1000 + t)
        env.sim.physics = lambda s: physics_t(s.root_states, s.dof_state, env.forces, 0.0166)
        obs_d, rew, reset, extras = _step_env(env, a, ua, ub)
        o_obs, o_rew, o_reset, o_ex = orc.step(a, ua, ub, physics=lambda o: physics_t(o.root_states, o.dof_state, o.forces, 0.0166))
        U.exact(reset, o_reset, f"reset t={t}")
        U.exact(extras["time_outs"], o_ex["time_outs"], f"time_outs t={t}")
        U.exact(env.progress_buf, orc.progress_buf, f"progress t={t}")
        U.humanoid_obs_close(obs_d["obs"], o_obs, f"obs t={t}")
        U.close(rew, o_rew, what=f"rew t={t}")
        # keep the two simulators bit-identical so that threshold decisions cannot drift apart
        orc.root_states.copy_(env.root_states.cpu()); orc.dof_state.copy_(env.dof_state.cpu())
        orc.potentials.copy_(env.potentials.cpu()); orc.prev_potentials.copy_(env.prev_potentials.cpu())
    assert int(env.progress_buf.max()) > 0


def _step_env(env, a, ua, ub):
    env._clip_in_kernel = True
    env.pre_physics_step(a.to(DEV))
    env._clip_in_kernel = False
    env.sim.simulate()
    env.post_physics_step(ua.to(DEV), ub.to(DEV))
    env.extras["time_outs"] = env.timeout_buf
    return {"obs": env.obs_buf}, env.rew_buf, env.reset_buf, env.extras


def test_unaligned_base_pointers_take_scalar_path():
    """Buffers at a 4-byte (not 16-byte) aligned address must still be handled (generic path of the kernels)."""
    n = 300
    env, orc, st = U.make_humanoid_pair(n, seed=77)

    def shifted(t):
        raw = torch.empty(t.numel() + 1, device=DEV, dtype=t.dtype)
        v = raw[1:].view(t.shape)
        v.copy_(t)
        return v
    env.sim.root_states = env.root_states = shifted(env.root_states)
    env.sim.dof_force = env.dof_force_tensor = shifted(env.dof_force_tensor)
    env.obs_buf = shifted(env.obs_buf)
    env.actions = shifted(env.actions)
    env.rebind()
    ids = orc.reset_buf.nonzero().flatten()
    ua, ub = syn.humanoid_reset_noise(max(len(ids), 1), seed=4)
    env.post_physics_step(ua.to(DEV), ub.to(DEV))
    orc.post_physics_step(ua, ub)
    U.humanoid_obs_close(env.obs_buf, orc.obs_buf)
    U.exact(env.reset_buf, orc.reset_buf, "reset_buf")
    U.exact(env.root_states, orc.root_states, "root_states")


@pytest.mark.parametrize("n", [16384, 131072])
def test_full_size_properties(n):
    """BASELINE sizes: sharded == unsharded bit-for-bit, idempotent observations, mask invariants, oracle spot check."""
    from thormang_isaacgym_b200 import _native as nat
    asset = syn.humanoid_asset()
    st = syn.humanoid_state(n, seed=1234, asset=asset)
    env, _, _ = U.make_humanoid_pair(n, state={k: v.clone() for k, v in st.items()}, asset=asset)
    env.post_physics_step()                      # philox resets
    full = {k: getattr(env, k).clone() for k in ("obs_buf", "rew_buf", "reset_buf", "progress_buf", "potentials", "dof_state", "root_states")}
    # two range shards with global env-id offsets reproduce the unsharded run exactly (no cross-env dependence)
    half = n // 2
    for r in range(2):
        sl = slice(r * half, (r + 1) * half)
        sub = {k: (v[sl.start * 21:sl.stop * 21].clone() if k == "dof_state" else v[sl].clone()) for k, v in st.items()}
        e2, _, _ = U.make_humanoid_pair(half, state=sub, asset=asset, envIdOffset=r * half)
        e2.noise_seed = env.noise_seed
        e2.post_physics_step()
        for k in ("obs_buf", "rew_buf", "reset_buf", "progress_buf", "potentials", "root_states"):
            U.exact(getattr(e2, k), full[k][sl], f"shard {r} {k}")
        U.exact(e2.dof_state, full["dof_state"][sl.start * 21:sl.stop * 21], f"shard {r} dof_state")
    # invariants
    obs, rew, reset = full["obs_buf"], full["rew_buf"], full["reset_buf"]
    assert bool(torch.isfinite(obs).all())
    fallen = obs[:, 0] < 0.8
    assert bool((reset[fallen] == 1).all())
    assert bool((rew[fallen] == -1.0).all())
    assert set(torch.unique(rew).tolist()) <= {-1.0, 2.0, float(np.float32(2.0) + np.float32(0.1))}
    assert bool(((reset == 0) | fallen | (full["progress_buf"] >= 999)).all())
    # observations are a pure function of state: recomputing changes nothing but prev_potentials
    env.compute_observations()
    U.exact(env.obs_buf[:, :109], obs[:, :109], "idempotent obs")
    # oracle spot check on a strided sample of rows (obs only: same state, no reset involved)
    idx = torch.arange(0, n, max(1, n // 2048))
    from oracle import humanoid_oracle as ho
    root = full["root_states"].cpu()[idx]
    dof = full["dof_state"].cpu().view(n, 21, 2)[idx]
    o, *_ = ho.humanoid_observations(root, torch.tensor([1000.0, 0, 0]).repeat(len(idx), 1), torch.zeros(len(idx)),
                                     torch.tensor([0.0, 0, 0, 1]).repeat(len(idx), 1), dof[..., 0], dof[..., 1],
                                     st["dof_force"][idx], asset["dof_limits_lower"], asset["dof_limits_upper"], 0.1,
                                     st["sensors"][idx], st["actions"][idx], 0.0166, 0.01, 0.25,
                                     torch.tensor([1.0, 0, 0]).repeat(len(idx), 1), torch.tensor([0.0, 0, 1]).repeat(len(idx), 1))
    U.humanoid_obs_close(obs.cpu()[idx][:, :109], o[:, :109], "spot check vs oracle")


def test_error_behaviour():
    import ctypes as C
    from thormang_isaacgym_b200 import _native as nat
    lib = nat.lib()
    p = nat.HumanoidParams()
    p.num_envs, p.num_dof, p.dt = 0, 21, 0.0166
    h = C.c_void_p()
    assert lib.tk_humanoid_create(C.byref(p), C.byref(h)) == -1 and b"num_envs" in lib.tk_last_error()
    p.num_envs, p.num_dof = 8, 8
    assert lib.tk_humanoid_create(C.byref(p), C.byref(h)) == -1 and b"num_dof" in lib.tk_last_error()
    p.num_dof = 21
    assert lib.tk_humanoid_create(C.byref(p), C.byref(h)) == 0
    assert lib.tk_humanoid_post_physics(h, None, None) == -3      # not bound
    with pytest.raises(RuntimeError):
        nat.check(lib.tk_humanoid_compute_observations(h, None), "unbound")
    b = nat.HumanoidBuffers()
    assert lib.tk_humanoid_bind(h, C.byref(b)) == -3 and b"root_states" in lib.tk_last_error()
    assert lib.tk_humanoid_destroy(h) == 0


def test_cuda_graph_replay_matches_eager():
    """The C-ABI calls allocate nothing and never synchronise, so a rollout can be captured in a CUDA graph.  The host
    step counter is frozen in the captured launches; the device-side noise epoch (bumped once per replay) supplies the
    upper 32 bits of the Philox step, so every replay draws fresh reset noise -- reproduced here by an eager run."""
    n = 4096
    env_a, _, st = U.make_humanoid_pair(n, seed=31)
    env_b, _, _ = U.make_humanoid_pair(n, seed=31, state={k: v.clone() for k, v in st.items()})
    acts = [(torch.rand(n, 21, generator=torch.Generator().manual_seed(i)) * 2 - 1).to(DEV) for i in range(2)]

    def two_steps(env, epoch=None):
        for i, a in enumerate(acts):
            env._clip_in_kernel = True
            env.pre_physics_step(a)
            env._clip_in_kernel = False
            if epoch is not None:
                env.step_count = i + (epoch << 32)      # what the captured launch + device epoch amount to
            env.post_physics_step()

    for r in range(3):
        two_steps(env_a, epoch=r + 1)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        env_b.advance_noise_epoch()
        two_steps(env_b)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    for name in ("obs_buf", "rew_buf", "reset_buf", "progress_buf", "dof_state", "root_states", "potentials", "forces"):
        U.exact(getattr(env_b, name), getattr(env_a, name), f"graph vs eager: {name}")
    assert int(env_a.progress_buf.max()) >= 5


def test_host_bridge_slabs_match_plain_task():
    """host_io: slab-backed task (one H2D, one D2H per step) == plain task on the same host state."""
    from thormang_isaacgym_b200 import config as tcfg
    from thormang_isaacgym_b200.host_io import ChunkedHostBridge, HostBridge, make_humanoid_for_host
    from thormang_isaacgym_b200.sim.fake_sim import FakeSim
    from thormang_isaacgym_b200.tasks.humanoid import Humanoid
    n = 2048
    asset = syn.humanoid_asset()
    st = syn.humanoid_state(n, seed=5, asset=asset)
    cfg = tcfg.humanoid_cfg(n, seed=0)
    cfg["env"]["episodeStatistics"] = False
    cfg["sim_backend"] = FakeSim(n, 21, DEV, num_sensor_floats=12, asset=asset)
    plain = Humanoid(cfg, DEV, DEV, -1, True)
    chunked = ChunkedHostBridge(lambda m, off: make_humanoid_for_host(m, DEV, off, seed=0), n, 4)
    assert chunked.bridges[0].in_slab is not None and chunked.bridges[0].out_slab is not None
    whole = HostBridge(plain)
    for b in (chunked, whole):
        b.fill_host(st) if isinstance(b, HostBridge) else b.fill_host(st)
    for _ in range(3):      # step 0 resets every env (reset_buf starts at ones), later steps reset the fallen ones
        chunked.step(); whole.step()
    chunked.synchronize(); whole.synchronize()
    per = n // 4
    for c, b in enumerate(chunked.bridges):
        sl = slice(c * per, (c + 1) * per)
        for k in ("obs_buf", "rew_buf", "reset_buf"):
            U.exact(b.host_out[k], whole.host_out[k][sl], f"chunk {c} {k}")
            U.exact(b.host_out[k], getattr(b.env, k).cpu(), f"chunk {c} {k} read-back")
    assert chunked.h2d_bytes >= 436 * n and chunked.d2h_bytes >= 452 * n


def test_stats_reducer_cuda_path_single_rank_nccl():
    """EpisodeStatsReducer on the GPU path (side stream + NCCL all-reduce), world_size 1: result == local statistics."""
    import torch.distributed as dist
    from thormang_isaacgym_b200.parallel import EpisodeStatsReducer
    if not dist.is_initialized():
        dist.init_process_group("nccl", init_method="tcp://127.0.0.1:29533", rank=0, world_size=1,
                                device_id=torch.device(DEV))
    try:
        n = 1024
        env, orc, _ = U.make_humanoid_pair(n, seed=8)
        env.post_physics_step()
        red = EpisodeStatsReducer(DEV)
        red.launch(env.episode_statistics())
        assert red._work is not None, "the all-reduce must be issued when a process group exists"
        out = red.result()
        assert out["env_steps"] == n
        assert abs(out["sum_reward"] - float(env.rew_buf.double().sum())) < 1e-3
        assert out["num_reset"] == int((env.reset_buf != 0).sum())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [4103, 16384])
def test_quad_kernel_matches_warp_kernel(n, monkeypatch):
    """The two tile kernels of the fused step (one warp per tile | four warps per tile, picked by grid size) are the same
    arithmetic cut differently: every output must agree bit for bit, resets and Philox noise included (3 steps)."""
    outs = []
    for quad in ("0", "1"):
        monkeypatch.setenv("TK_HUMANOID_QUAD", quad)        # read when the handle is created
        env, _, st = U.make_humanoid_pair(n, seed=99)
        for i in (0, 31, 32, 33, n - 1):
            st["reset_buf"][i] = 1
        env.reset_buf.copy_(st["reset_buf"].to(DEV))
        for _ in range(3):
            env.post_physics_step()
        torch.cuda.synchronize()
        outs.append({k: getattr(env, k).clone() for k in ("obs_buf", "rew_buf", "reset_buf", "progress_buf", "timeout_buf",
                                                           "potentials", "prev_potentials", "dof_state", "root_states")})
    for k in outs[0]:
        assert torch.equal(outs[0][k], outs[1][k]), f"{k}: four-warp kernel differs from the one-warp kernel"


def test_config4_one_million_envs_in_eight_shards():
    """BASELINE config 4 exactly: 1,048,576 Humanoid envs as 8 range shards of 131072 with global env-id offsets (what 8
    ranks run) against the unsharded run on one GPU: every output bit for bit, and the per-shard episode statistics (the
    all-reduce payload) sum to the single-process vector."""
    n, shards = 1048576, 8
    asset = syn.humanoid_asset()
    st = syn.humanoid_state(n, seed=4, asset=asset)
    env, _, _ = U.make_humanoid_pair(n, state={k: v.clone() for k, v in st.items()}, asset=asset)
    env.post_physics_step()                      # in-kernel Philox resets keyed by the global env id
    keys = ("obs_buf", "rew_buf", "reset_buf", "progress_buf", "timeout_buf", "potentials", "root_states")
    full = {k: getattr(env, k).clone() for k in keys + ("dof_state",)}
    total = env.episode_statistics().double().cpu()
    del env
    per = n // shards
    acc = torch.zeros_like(total)
    for r in range(shards):
        sl = slice(r * per, (r + 1) * per)
        sub = {k: (v[sl.start * 21:sl.stop * 21].clone() if k == "dof_state" else v[sl].clone()) for k, v in st.items()}
        e2, _, _ = U.make_humanoid_pair(per, state=sub, asset=asset, envIdOffset=r * per)
        e2.post_physics_step()
        for k in keys:
            assert torch.equal(getattr(e2, k), full[k][sl]), f"shard {r} {k}"
        assert torch.equal(e2.dof_state, full["dof_state"][sl.start * 21:sl.stop * 21]), f"shard {r} dof_state"
        acc += e2.episode_statistics().double().cpu()
        del e2
    U.close(acc, total, rtol=1e-9, atol=1e-3, what="sum of the shard statistics vs the single-process vector")
    assert float(total[7]) == n


def test_config5_ppo_rollout_shape_through_step():
    """BASELINE config 5 at the rl_games PPO shape of cfg/train/HumanoidPPO.yaml:47,66 -- 4096 actors x horizon 32 -- through
    the public VecTask.step (clamp + pre_physics -> fake physics -> fused post_physics with in-kernel Philox resets), obs
    delivered into the [horizon, actors, obs] experience tensor by the kernel; every step compared with the oracle."""
    n, horizon = 4096, 32
    env, orc, st = U.make_humanoid_pair(n, seed=32)

    def physics_t(root, dof_state, forces, dt):
        root[:, 0:3] += dt * root[:, 7:10]
        root[:, 2] -= 0.01
        dof_state[:, 0] += dt * dof_state[:, 1]
        dof_state[:, 1] += 0.001 * forces.reshape(-1)

    env.sim.physics = lambda s: physics_t(s.root_states, s.dof_state, env.forces, 0.0166)
    buf = torch.full((horizon, n, 110), float("nan"), device=DEV)
    env.set_rollout_buffer(buf)
    g = torch.Generator().manual_seed(9)
    resets = 0
    for t in range(horizon):
        a = torch.rand(n, 21, generator=g) * 2.4 - 1.2          # beyond the clip range: the clamp of vec_task.py:327 is live
        ids = orc.reset_buf.nonzero().flatten()
        resets += len(ids)
        pa, pb = env.philox_noise()                              # the draws the next post-physics launch will use
        ua, ub = (pa.cpu()[ids], pb.cpu()[ids]) if len(ids) else syn.humanoid_reset_noise(1, seed=1)
        obs_d, rew, reset, extras = env.step(a.to(DEV))
        o_obs, o_rew, o_reset, o_ex = orc.step(a, ua, ub, physics=lambda o: physics_t(o.root_states, o.dof_state, o.forces, 0.0166))
        assert obs_d["obs"].data_ptr() == buf[t].data_ptr()
        U.exact(reset, o_reset, f"reset t={t}")
        U.exact(extras["time_outs"], o_ex["time_outs"], f"time_outs t={t}")
        U.humanoid_obs_close(buf[t], o_obs, f"rollout slot t={t}")
        U.close(rew, o_rew, what=f"rew t={t}")
        orc.root_states.copy_(env.root_states.cpu()); orc.dof_state.copy_(env.dof_state.cpu())
        orc.potentials.copy_(env.potentials.cpu()); orc.prev_potentials.copy_(env.prev_potentials.cpu())
    assert resets > 0 and bool(torch.isfinite(buf).all())


@pytest.mark.parametrize("cuts", [(1,), (333,), (31, 64, 700), (512, 513)])
def test_ragged_shards_with_global_env_ids(cuts):
    """Env-range sharding at ARBITRARY cut points (shard sizes 1, 31, 333 ...; offsets that are no multiple of the 32-env tile):
    in-kernel Philox noise is keyed by the global env id, so the shards must reproduce the unsharded step bit for bit."""
    n = 1000
    asset = syn.humanoid_asset()
    st = syn.humanoid_state(n, seed=77, asset=asset)
    st["reset_buf"][::9] = 1
    env, _, _ = U.make_humanoid_pair(n, state={k: v.clone() for k, v in st.items()}, asset=asset)
    env.post_physics_step()
    keys = ("obs_buf", "rew_buf", "reset_buf", "progress_buf", "potentials", "root_states")
    full = {k: getattr(env, k).clone() for k in keys + ("dof_state",)}
    edges = (0,) + tuple(cuts) + (n,)
    for lo, hi in zip(edges[:-1], edges[1:]):
        sub = {k: (v[lo * 21:hi * 21].clone() if k == "dof_state" else v[lo:hi].clone()) for k, v in st.items()}
        e2, _, _ = U.make_humanoid_pair(hi - lo, state=sub, asset=asset, envIdOffset=lo)
        e2.post_physics_step()
        for k in keys:
            assert torch.equal(getattr(e2, k), full[k][lo:hi]), f"shard [{lo},{hi}) {k}"
        assert torch.equal(e2.dof_state, full["dof_state"][lo * 21:hi * 21]), f"shard [{lo},{hi}) dof_state"


def test_write_debug_vectors_switch():
    """cfg env.writeDebugVectors = False: up_vec / heading_vec (read by the reference's debug drawing only, humanoid.py:298-316)
    keep their initial values, every other output of the fused step is bit-identical to the default (which refreshes them)."""
    n = 4099
    outs = []
    for flag in (True, False):
        env, _, st = U.make_humanoid_pair(n, seed=5, writeDebugVectors=flag)
        up0, head0 = env.up_vec.clone(), env.heading_vec.clone()
        env.post_physics_step()
        torch.cuda.synchronize()
        if flag:
            assert not torch.equal(env.up_vec, up0) and not torch.equal(env.heading_vec, head0)
        else:
            assert torch.equal(env.up_vec, up0) and torch.equal(env.heading_vec, head0)
        outs.append({k: getattr(env, k).clone() for k in ("obs_buf", "rew_buf", "reset_buf", "progress_buf", "timeout_buf",
                                                           "potentials", "prev_potentials", "dof_state", "root_states")})
    for k in outs[0]:
        assert torch.equal(outs[0][k], outs[1][k]), k


@pytest.mark.parametrize("n", [4096, 65536])      # four-warp kernel | one-warp kernel
def test_cuda_graph_epoch_directly_in_front_of_the_step_many_replays(n):
    """Both tile kernels resolve the device-side noise epoch AHEAD of their dependency wait.  Worst case: a graph whose epoch
    kernel sits directly in front of ONE post_physics launch, replayed 40 times back to back; the reset noise of every replay
    must be that of ITS epoch (bit-identical to eager launches keyed (epoch << 32) | step)."""
    replays = 40
    env_a, _, st = U.make_humanoid_pair(n, seed=33)
    env_b, _, _ = U.make_humanoid_pair(n, seed=33, state={k: v.clone() for k, v in st.items()})
    for e in (env_a, env_b):
        e.reset_buf[::5] = 1
    for r in range(replays):
        env_a.step_count = (r + 1) << 32
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
    for name in ("obs_buf", "rew_buf", "progress_buf", "dof_state", "root_states", "potentials", "prev_potentials"):
        U.exact(getattr(env_b, name), getattr(env_a, name), f"graph vs eager after {replays} replays: {name}")
