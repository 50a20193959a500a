"""GPU arm of bench.py (see bench.py for the contract).  One process per GPU; env ranges sharded per rank."""
import json
import math
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import bench as B   # noqa: E402  (constants, ClockSampler, CPU arm)

HORIZON = 32        # PPO horizon_length of cfg/train/HumanoidPPO.yaml:66: statistics are all-reduced once per horizon


def _make_humanoid(n, seed, device, env_off=0, stats=True, debug_vectors=True):
    import torch
    from thormang_isaacgym_b200 import config as tcfg
    from thormang_isaacgym_b200 import synthetic as syn
    from thormang_isaacgym_b200.sim.fake_sim import FakeSim
    from thormang_isaacgym_b200.tasks.humanoid import Humanoid
    asset = syn.humanoid_asset()
    st = syn.humanoid_state(n, seed=seed, asset=asset, fallen_frac=0.01)
    cfg = tcfg.humanoid_cfg(n, seed=1234)
    cfg["env"]["envIdOffset"] = env_off
    cfg["env"]["episodeStatistics"] = stats
    # up_vec / heading_vec (humanoid.py:246) are read by the viewer's debug drawing only (:298-316) and are not part of the
    # 924 B per env-step of SURVEY 8(a) ("+24 B if written").  The reference assigns them every step, so the headline and
    # every default line refresh them too (the task's default); extras *_no_debug_vectors time the same launches without
    cfg["env"]["writeDebugVectors"] = bool(debug_vectors)
    sim = FakeSim(n, 21, device, num_sensor_floats=12, asset=asset)
    sim.load_state(st)
    cfg["sim_backend"] = sim
    dev_index = torch.device(device).index or 0
    env = Humanoid(cfg, device, f"cuda:{dev_index}", -1, True)
    env.actions.copy_(st["actions"].to(device))
    env.progress_buf.copy_(st["progress_buf"].to(device))
    env.reset_buf.copy_(st["reset_buf"].to(device))
    env.potentials.copy_(st["potentials"].to(device))
    return env, st


def _make_gogoro(n, seed, device, env_off=0):
    import torch
    from thormang_isaacgym_b200 import config as tcfg
    from thormang_isaacgym_b200 import synthetic as syn
    from thormang_isaacgym_b200.sim.fake_sim import FakeSim
    from thormang_isaacgym_b200.tasks.gogoro_new import Gogoro
    st = syn.gogoro_state(n, seed=seed)
    cfg = tcfg.gogoro_cfg(n, seed=1234)
    cfg["env"]["envIdOffset"] = env_off
    cfg["env"]["resetOnInit"] = False
    names = ["steering_joint", "rear_wheel_joint"] + [f"j{i}" for i in range(37)]
    sim = FakeSim(n, 39, device, num_sensor_floats=0, dof_names=names)
    sim.load_state(st)
    cfg["sim_backend"] = sim
    dev_index = torch.device(device).index or 0
    env = Gogoro(cfg, device, f"cuda:{dev_index}", -1, True)
    for name in ("yaw_command", "curent_command", "curent_speed", "steer_offsets", "imu_offsets",
                 "curent_speed_offset", "action_history", "progress_buf", "reset_buf", "thormang_pose"):
        getattr(env, name).copy_(st[name].to(device))
    return env, st


def _make_gogoro_paper(n, seed, device, env_off=0, world_reduce=False):
    """8-channel x 20-frame Gogoro variant; ~2 % of the envs beyond the fall threshold per step."""
    import torch
    from thormang_isaacgym_b200 import config as tcfg
    from thormang_isaacgym_b200 import synthetic as syn
    from thormang_isaacgym_b200.sim.fake_sim import FakeSim
    from thormang_isaacgym_b200.tasks.gogoro_realistic_turning_sim_paper import Gogoro as GogoroPaper
    st = syn.gogoro_paper_state(n, seed=seed, fallen_frac=0.02)
    cfg = tcfg.gogoro_paper_cfg(n, seed=1234)
    cfg["env"]["envIdOffset"] = env_off
    cfg["env"]["resetOnInit"] = False
    if world_reduce:
        cfg["env"]["worldReduce"] = True      # :747 cross-env mean: diff sums -> NCCL all-reduce -> finalize, every step
    names = ["steering_joint", "rear_wheel_joint"] + [f"j{i}" for i in range(37)]
    sim = FakeSim(n, 39, device, num_sensor_floats=0, dof_names=names)
    sim.load_state(st)
    cfg["sim_backend"] = sim
    dev_index = torch.device(device).index or 0
    env = GogoroPaper(cfg, device, f"cuda:{dev_index}", -1, True)
    for name in ("yaw_command", "curent_command", "curent_speed", "steer_offsets", "curent_speed_offset",
                 "curent_imu_x_offset", "steer_delay", "command_history", "progress_buf", "reset_buf", "thormang_pose"):
        getattr(env, name).copy_(st[name].to(device))
    env.buffer_obs_ring.copy_(st["buffer_obs"].to(device))
    env.command_ring.copy_(st["buffer_obs"][:, :, 6].to(device))
    env.obs_buf.copy_(st["buffer_obs_noisy"].flatten(start_dim=1).to(device))
    return env, st


def _make_anymal(n, seed, device, env_off=0):
    import torch
    from thormang_isaacgym_b200 import config as tcfg
    from thormang_isaacgym_b200 import synthetic as syn
    from thormang_isaacgym_b200.sim.fake_sim import FakeSim
    from thormang_isaacgym_b200.tasks.anymal import Anymal
    st = syn.anymal_state(n, seed=seed, contact_frac=0.01)
    cfg = tcfg.anymal_cfg(n, seed=1234)
    cfg["env"]["envIdOffset"] = env_off
    cfg["env"]["resetOnInit"] = False
    sim = FakeSim(n, 12, device, num_sensor_floats=0, dof_names=list(tcfg.ANYMAL_DOF_NAMES), num_bodies=13)
    sim.load_state(st)
    cfg["sim_backend"] = sim
    dev_index = torch.device(device).index or 0
    env = Anymal(cfg, device, f"cuda:{dev_index}", -1, True)
    for name in ("commands", "actions", "progress_buf", "reset_buf"):
        getattr(env, name).copy_(st[name].to(device))
    return env, st


def _make_ant(n, seed, device, env_off=0):
    """Ant through the Humanoid kernel templates (AntSpec: D = 8, 60 obs, 24 sensor floats, every reward term live)."""
    import torch
    from thormang_isaacgym_b200 import config as tcfg
    from thormang_isaacgym_b200 import synthetic as syn
    from thormang_isaacgym_b200.sim.fake_sim import FakeSim
    from thormang_isaacgym_b200.tasks.ant import Ant
    asset = syn.ant_asset()
    st = syn.ant_state(n, seed=seed, asset=asset, fallen_frac=0.01)     # ~1 % of the envs below the termination height, like the Humanoid workload
    cfg = tcfg.ant_cfg(n, seed=1234)
    cfg["env"]["envIdOffset"] = env_off
    cfg["env"]["episodeStatistics"] = True
    sim = FakeSim(n, 8, device, num_sensor_floats=24, asset=asset)
    sim.load_state(st)
    cfg["sim_backend"] = sim
    dev_index = torch.device(device).index or 0
    env = Ant(cfg, device, f"cuda:{dev_index}", -1, True)
    for name in ("actions", "progress_buf", "reset_buf", "potentials", "prev_potentials"):
        getattr(env, name).copy_(st[name].to(device))
    return env, st


def _make_ma_op3(n, seed, device, env_off=0, grip_all_reduce=False):
    """MA_OP3: two OP3 + table per env, 51 rigid bodies; ~1 % of the envs in reset per step."""
    import torch
    from thormang_isaacgym_b200 import config as tcfg
    from thormang_isaacgym_b200 import synthetic as syn
    from thormang_isaacgym_b200.sim.fake_sim import FakeSim
    from thormang_isaacgym_b200.tasks.MA_OP3 import MA_OP3
    st = syn.ma_op3_state(n, seed=seed)
    cfg = tcfg.ma_op3_cfg(n, seed=1234)
    cfg["env"]["envIdOffset"] = env_off
    cfg["env"]["resetOnInit"] = False
    cfg["env"]["gripAllReduce"] = bool(grip_all_reduce)   # MA_OP3.py:346-358 batch-global counts: int64[4] all-reduce per step
    nb = st["contact_forces"].shape[1]
    sim = FakeSim(n, 44, device, num_sensor_floats=0, dof_names=list(tcfg.OP3_DOF_NAMES), num_bodies=nb, actors_per_env=3)
    sim.load_state(st)
    cfg["sim_backend"] = sim
    dev_index = torch.device(device).index or 0
    env = MA_OP3(cfg, device, f"cuda:{dev_index}", -1, True)
    for name in ("prev_torques", "commands", "actions", "last_actions", "feet_air_time", "last_contacts", "potentials",
                 "prev_potentials", "table_potentials", "progress_buf", "reset_buf"):
        getattr(env, name).copy_(st[name].to(device))
    return env, st


_MAKERS = {"ant": _make_ant, "ma_op3": _make_ma_op3, "anymal": _make_anymal, "humanoid": _make_humanoid, "gogoro": _make_gogoro, "gogoro_paper": _make_gogoro_paper}
_LAUNCHES_PER_STEP = {"ant": 1, "ma_op3": 2, "anymal": 1, "humanoid": 1, "gogoro": 1, "gogoro_paper": 2}   # the paper variant: fused kernel + reward finalize


def _rotating_sets(n, bytes_per_env):
    if os.environ.get("TK_BENCH_SETS"):           # developer knob (tools/kbench.py): L2-resident vs HBM-cold comparisons
        return int(os.environ["TK_BENCH_SETS"])
    per_set = n * bytes_per_env
    return max(3, int(math.ceil(2.2 * B.L2_BYTES / per_set)))


def _time_steps(torch, step_fn, steps, warmup, barrier=None, after_warmup=None):
    """W untimed steps, then exactly K steps between two events on the launching stream.
    step_fn(i) returns the number of kernels of ours it launched.  Returns (ms, launches, (t0, t1)).
    after_warmup() runs between the warm-up and the barrier (outside the timed region)."""
    for i in range(warmup):
        step_fn(i)
    if after_warmup is not None:
        after_warmup()
    torch.cuda.synchronize()
    if barrier:
        barrier()
        torch.cuda.synchronize()      # the barrier's own NCCL kernel has drained before the first event is recorded
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    ev0.record()
    launches = 0
    for i in range(steps):
        launches += step_fn(warmup + i)
    ev1.record()
    _time_steps.last_host_enqueue_ms = (time.perf_counter() - t0) * 1e3   # host time to enqueue the K steps (diagnostic)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    if barrier:
        barrier()
    return ev0.elapsed_time(ev1), launches, (t0, t1)


def _kernel_bench(torch, task, n, device, steps, warmup, seed=1234, env_off=0, barrier=None, horizon_hook=None,
                  graph=False, horizon=HORIZON, cfg_patch=None):
    """Fused post-physics kernel over rotating state sets.  Returns dict(ms_total, launches, sets, window).
    graph=True: one pass over the sets is captured in a CUDA graph and replayed (steps is rounded to whole passes).
    horizon_hook(env, steps_since_last_finalize) fires after every `horizon`-th TIMED step (warm-up steps do not count;
    the statistics partials of every set are discarded after the warm-up, outside the timed region)."""
    bpe = B.BYTES_PER_ENV[task]
    sets = _rotating_sets(n, bpe)
    make = _MAKERS[task]
    per_step = _LAUNCHES_PER_STEP[task]
    envs = [make(n, seed + 17 * s, device, env_off, **(cfg_patch or {}))[0] for s in range(sets)]

    if graph:
        for e in envs:
            e.post_physics_step()
        torch.cuda.synchronize()
        # one captured graph = one rollout horizon (>= 32 steps, whole passes over the state sets), like a training loop
        # that captures its horizon: ONE noise-epoch bump per task object per replay, all of them ahead of the steps so that
        # the step kernels follow one another directly (programmatic-launch edges between consecutive kernel nodes); every
        # captured launch carries its own host step number, the epoch makes every replay draw fresh noise
        per_graph = max(1, -(-HORIZON // sets))
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for e in envs:
                e.advance_noise_epoch()
            for _ in range(per_graph):
                for e in envs:
                    e.post_physics_step()
        per_replay = per_graph * sets
        replays = max(1, steps // per_replay)

        def replay(i):
            g.replay()
            return per_replay * per_step + sets
        ms, launches, window = _time_steps(torch, replay, replays, max(1, warmup // per_replay), barrier)
        del envs
        return {"ms_total": ms, "launches": launches, "sets": sets, "window": window, "steps": replays * per_replay}

    since = [0] * sets          # steps of each set since its statistics were last finalized

    def after_warmup():
        if horizon_hook is not None:
            horizon_hook(envs[0], 0, warm=True)   # first use of the side stream / the fp64 all-reduce happens untimed
            for e in envs:
                e.episode_statistics()      # finalize + zero the partials: the timed region starts from empty statistics
            for k in range(sets):
                since[k] = 0

    def step(i):
        k = i % sets
        envs[k].post_physics_step()
        since[k] += 1
        extra = 0
        t = i - warmup                      # index inside the timed region
        if horizon_hook is not None and t >= 0 and (t + 1) % horizon == 0:
            extra = horizon_hook(envs[k], since[k])
            since[k] = 0
        return per_step + extra

    ms, launches, window = _time_steps(torch, step, steps, warmup, barrier, after_warmup)
    del envs
    return {"ms_total": ms, "launches": launches, "sets": sets, "window": window, "steps": steps,
            "host_enqueue_ms": _time_steps.last_host_enqueue_ms}


def _e2e_bench(torch, n, device, steps, chunks, env_off, barrier=None):
    """Public-API path with HOST buffers: per step H2D(state) -> fused kernel -> D2H(obs, rew, reset), chunk-pipelined."""
    from thormang_isaacgym_b200 import synthetic as syn
    from thormang_isaacgym_b200.host_io import ChunkedHostBridge, make_humanoid_for_host
    while n % chunks:
        chunks //= 2
    chunks = max(1, chunks)
    bridge = ChunkedHostBridge(lambda m, off: make_humanoid_for_host(m, device, env_off + off, seed=1234), n, chunks)
    st = syn.humanoid_state(n, seed=4321, asset=syn.humanoid_asset(), fallen_frac=0.01)
    bridge.fill_host(st)
    for _ in range(3):
        bridge.step()
    bridge.synchronize()
    torch.cuda.synchronize()
    if barrier:
        barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        bridge.step()
    bridge.synchronize()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    if barrier:
        barrier()
    # sanity: what came back is what the kernel wrote
    b0 = bridge.bridges[0]
    assert torch.equal(b0.host_out["reset_buf"], b0.env.reset_buf.cpu()), "e2e read-back mismatch"
    return {"seconds": t1 - t0, "h2d": bridge.h2d_bytes, "d2h": bridge.d2h_bytes, "chunks": chunks,
            "launches": steps * chunks, "window": (t0, t1), "placement": bridge.placement}


def _pcie_probe(torch, device, mb=64):
    """Pinned-memory copy bandwidth of this box, for reading the e2e number (one direction at a time)."""
    host = torch.empty(mb * 1024 * 1024, dtype=torch.uint8, pin_memory=True)
    dev = torch.empty_like(host, device=device)
    out = {}
    for name, (dst, src) in (("h2d_GBps", (dev, host)), ("d2h_GBps", (host, dev))):
        dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5):
            dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        out[name] = 5 * host.numel() / (time.perf_counter() - t0) / 1e9
    # both directions at once on two streams: what a perfectly overlapped e2e step could reach per direction
    host2 = torch.empty_like(host).pin_memory()
    dev2 = torch.empty_like(dev)
    s1, s2 = torch.cuda.Stream(device=device), torch.cuda.Stream(device=device)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        with torch.cuda.stream(s1):
            dev.copy_(host, non_blocking=True)
        with torch.cuda.stream(s2):
            host2.copy_(dev2, non_blocking=True)
    torch.cuda.synchronize()
    out["duplex_GBps_per_direction"] = 5 * host.numel() / (time.perf_counter() - t0) / 1e9
    return out


def _pcie_probe_concurrent(torch, dist, device, world, barrier, mb=64, reps=8):
    """What the BOX gives when all N ranks copy at once (N > 1): every rank, after a barrier, drives both directions of its
    own PCIe link from pinned slabs placed like the e2e path places them (host_io.near_gpu).  The e2e line can at best
    move its h2d + d2h bytes at this rate: if the per-GPU figure drops as N grows the ceiling is the host side of the box
    (memory / root complex shared by the GPUs), not host_io.py."""
    from thormang_isaacgym_b200.host_io import near_gpu
    place = {}
    with near_gpu(torch.device(device).index or 0, place):
        host = torch.empty(mb * 1024 * 1024, dtype=torch.uint8, pin_memory=True)
        host2 = torch.empty(mb * 1024 * 1024, dtype=torch.uint8, pin_memory=True)
    dev, dev2 = torch.empty_like(host, device=device), torch.empty_like(host, device=device)
    s1, s2 = torch.cuda.Stream(device=device), torch.cuda.Stream(device=device)
    for _ in range(2):
        with torch.cuda.stream(s1):
            dev.copy_(host, non_blocking=True)
        with torch.cuda.stream(s2):
            host2.copy_(dev2, non_blocking=True)
    torch.cuda.synchronize()
    if barrier:
        barrier()
        torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        with torch.cuda.stream(s1):
            dev.copy_(host, non_blocking=True)
        with torch.cuda.stream(s2):
            host2.copy_(dev2, non_blocking=True)
    torch.cuda.synchronize()
    mine = reps * host.numel() / (time.perf_counter() - t0) / 1e9
    t = torch.tensor([mine], device=device, dtype=torch.float64)
    every = [torch.zeros_like(t) for _ in range(world)]
    if dist:
        dist.all_gather(every, t)
    else:
        every = [t]
    per = [float(x.item()) for x in every]
    return {"duplex_GBps_per_direction_per_rank": [round(v, 2) for v in per], "sum_over_ranks_GBps_per_direction": round(sum(per), 2),
            "ranks_copying_at_once": world, "mb_per_copy": mb, "pinned_placement": place}


def _launch_floor(torch, device):
    """Launch + drain floor of this box next to the small-grid numbers: an EMPTY kernel with the launch shape of the fused
    kernels (programmatic dependent launch, same CTA count / size / shared memory), back to back on one stream -- through
    the same ctypes call path (host-bound: ~3 us per call) and replayed from a CUDA graph (the GPU-side floor, which
    grows with the CTA count: a B200 dispatches ~0.85 CTA per ns)."""
    import ctypes as C
    from thormang_isaacgym_b200 import _native as nat
    fn = nat.lib().tk_debug_launch_floor
    fn.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    fn.restype = C.c_int
    out = {}
    for label, grid, block, smem in (("humanoid_16384_quad_512x128", 512, 128, 28 * 1024), ("humanoid_65536_2048x32", 2048, 32, 14 * 1024),
                                     ("humanoid_131072_4096x32", 4096, 32, 14 * 1024), ("gogoro_65536_512x128", 512, 128, 18 * 1024)):
        sink = torch.zeros(grid, dtype=torch.int32, device=device)

        def eager(i):
            fn(grid, block, smem, 1, sink.data_ptr(), nat.current_stream_ptr(device))
            return 1
        ms, _, _ = _time_steps(torch, eager, 1000, 50)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(50):
                fn(grid, block, smem, 1, sink.data_ptr(), nat.current_stream_ptr(device))
        msg, _, _ = _time_steps(torch, lambda i: (g.replay(), 50)[1], 20, 2)
        out[label] = {"us_per_launch_eager": 1e3 * ms / 1000, "us_per_launch_cuda_graph": 1e3 * msg / (20 * 50)}
    return out


def _l2_resident(torch, task, m, device, peak):
    """The same kernel over TWO state sets only (both L2-resident, translations hot): what a training loop that steps ONE
    env object sees between launches.  The headline extras rotate 20-26 sets to force every byte from HBM, which also makes
    every launch start from cold TLB / L2 state; the difference between the two lines is that artefact."""
    os.environ["TK_BENCH_SETS"] = "2"
    try:
        r = _kernel_bench(torch, task, m, device, 300, 20, seed=78)
    finally:
        del os.environ["TK_BENCH_SETS"]
    us = 1e3 * r["ms_total"] / r["steps"]
    gbs = m * B.BYTES_PER_ENV[task] / (us * 1e-6) / 1e9
    return {"us_per_step": us, "env_steps_per_s": m / (us * 1e-6), "achieved_GBps": gbs, "frac_of_peak": gbs / peak,
            "rotating_sets": 2, "note": "L2-resident: not an HBM number"}


def _config1(torch, device):
    """BASELINE config 1 (Humanoid compute_observations + compute_reward, 4096 envs): the CPU torch path as SURVEY 8(d)
    specifies it (all threads and one thread, 5 warm-ups, median of 50) and the same two method calls on the GPU."""
    out = {"cpu": B.cpu_config1(4096, 5, 50)}
    env, _ = _make_humanoid(4096, 4096, device)

    def pair(i):
        env.compute_observations()
        env.compute_reward(env.actions)
        return 2
    ms, _, _ = _time_steps(torch, pair, 400, 20)
    us = 1e3 * ms / 400
    out["gpu_split_methods"] = {"us_per_call_pair": us, "env_steps_per_s": 4096 / (us * 1e-6), "launches": 2,
                                "note": "compute_observations() + compute_reward() through the method-level C-ABI entry "
                                        "points (two launches, one L2-resident state set: launch-bound at this size)"}
    out["gpu_over_cpu_all_threads"] = out["gpu_split_methods"]["env_steps_per_s"] / out["cpu"]["all_threads"]["env_steps_per_s"]
    return out


def _step_loop_bench(torch, n, device, horizon, reps, graph=False):
    """BASELINE config 5: pre_physics (clamp + force write) -> fake physics (identity) -> fused post_physics, PPO horizon.
    graph=True captures one horizon (2 x horizon kernels) in a CUDA graph and replays it."""
    env, st = _make_humanoid(n, 999, device)
    acts = [torch.rand((n, 21), device=device) * 2.4 - 1.2 for _ in range(4)]

    def one(i):
        env._clip_in_kernel = True
        env.pre_physics_step(acts[i % 4])
        env._clip_in_kernel = False
        env.sim.simulate()
        env.post_physics_step()
        return 2
    if graph:
        for i in range(4):
            one(i)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            env.advance_noise_epoch()
            for i in range(horizon):
                one(i)
        ms, launches, _ = _time_steps(torch, lambda i: (g.replay(), 2 * horizon)[1], reps, 2)
    else:
        ms, launches, _ = _time_steps(torch, one, horizon * reps, horizon)
    return {"envs": n, "horizon": horizon, "env_steps_per_s": n * horizon * reps / (ms * 1e-3),
            "us_per_step": 1e3 * ms / (horizon * reps), "launches_per_step": 2, "cuda_graph": graph,
            "note": "pre (252 B/env) + post (924 B/env) on ONE state set: L2-resident at this size, not an HBM number"}


def _multi_gpu_diag(torch, dist, task, n, device, rank, world, local_rank, barrier, steps):
    """N > 1 only, outside the headline timing.  (1) per-step CUDA-event times of the first `steps` steps after a barrier,
    per rank: separates a barrier tail / clock ramp at the head of the window from a steady-state loss.  (2) the two
    IN-STEP exchanges of the path family, timed with their NCCL all-reduce inside every step and checked against the
    sum of the per-rank values: the paper variant's cross-env reward mean (gogoro_realistic_turning_sim_paper.py:747) and
    MA_OP3's batch-global gripper counts (MA_OP3.py:346-358)."""
    out = {}
    # ---- (1) per-step profile of the headline workload
    bpe = B.BYTES_PER_ENV[task]
    sets = _rotating_sets(n, bpe)
    envs = [_MAKERS[task](n, 99 + 17 * k + 1000 * rank, device, rank * n)[0] for k in range(sets)]
    for i in range(6):
        envs[i % sets].post_physics_step()
    torch.cuda.synchronize()
    if barrier:
        barrier()
        torch.cuda.synchronize()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    evs[0].record()
    for i in range(steps):
        envs[i % sets].post_physics_step()
        evs[i + 1].record()
    torch.cuda.synchronize()
    per = torch.tensor([evs[i].elapsed_time(evs[i + 1]) * 1e3 for i in range(steps)], device=device, dtype=torch.float64)
    allp = [torch.zeros_like(per) for _ in range(world)]
    if dist:
        dist.all_gather(allp, per)
    else:
        allp = [per]
    out["per_step_us_after_barrier"] = {
        "note": "an event after EVERY step (this serialises consecutive launches: slower than the headline loop, "
                "which records two events in all); first rows = head of the window",
        "per_rank": [[round(float(v), 2) for v in t.tolist()] for t in allp]}
    del envs
    if not dist:
        return out

    # ---- (2) in-step exchanges
    def timed(make_kwargs, tk, m, k_steps):
        r = _kernel_bench(torch, tk, m, device, k_steps, 10, seed=555 + 1000 * rank, env_off=rank * m, barrier=barrier,
                          cfg_patch=make_kwargs)
        t = torch.tensor([r["ms_total"]], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        return {"us_per_step": 1e3 * ms / k_steps, "env_steps_per_s": m * world * k_steps / (ms * 1e-3), "envs_per_gpu": m}

    m = 65536
    try:
        with_x = timed({"world_reduce": True}, "gogoro_paper", m, 60)
        no_x = timed({}, "gogoro_paper", m, 60)
        env = _make_gogoro_paper(m, 4242 + rank, device, rank * m, world_reduce=True)[0]
        env.post_physics_step()
        torch.cuda.synchronize()
        mine = env.diff_sums.clone()
        every = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(every, mine)
        same = all(torch.equal(every[0], e) for e in every)
        out["paper_cross_env_mean"] = {**with_x, "us_per_step_without_exchange": no_x["us_per_step"],
                                       "collective": "NCCL SUM all-reduce of 2 x fp64 between the fused kernel and the finalize grid, every step",
                                       "reduced_sums_identical_on_all_ranks": bool(same), "reduced_sums": every[0].tolist()}
        del env
    except Exception as exc:
        out["paper_cross_env_mean"] = {"error": repr(exc)}
    try:
        with_x = timed({"grip_all_reduce": True}, "ma_op3", m, 60)
        no_x = timed({}, "ma_op3", m, 60)
        twin = _make_ma_op3(m, 4343 + rank, device, rank * m, grip_all_reduce=False)[0]
        twin.launch_post_physics()                  # same state, no exchange: this rank's own counts
        local = twin.grip_count_buf.clone()
        want = local.clone()
        dist.all_reduce(want)                       # what the step's exchange must produce
        env = _make_ma_op3(m, 4343 + rank, device, rank * m, grip_all_reduce=True)[0]
        env.post_physics_step()
        torch.cuda.synchronize()
        del twin
        ok = torch.equal(env.grip_count_buf.view(-1), want.view(-1))
        okt = torch.tensor([1 if ok else 0], device=device, dtype=torch.int32)
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
        out["ma_op3_grip_counts"] = {**with_x, "us_per_step_without_exchange": no_x["us_per_step"],
                                     "collective": "NCCL SUM all-reduce of int64[4] between the fused kernel and the finalize grid, every step",
                                     "global_counts_equal_sum_of_rank_counts": bool(okt.item()),
                                     "global_counts": want.view(-1).tolist(), "rank0_counts": local.view(-1).tolist()}
        del env
    except Exception as exc:
        out["ma_op3_grip_counts"] = {"error": repr(exc)}
    return out


def run_gpu_arm(args, rank, world):
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl=ours) needs a CUDA device: the env-step path has no CPU fallback")
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    device = f"cuda:{local_rank}"
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(device))
    barrier = (lambda: dist.barrier(device_ids=[local_rank])) if dist else None

    from thormang_isaacgym_b200 import _native
    _native.lib()   # fail loudly if the extension is missing

    task, n = args.task, args.envs_per_gpu
    bpe = B.HUMANOID_BYTES_PER_ENV if task == "humanoid" else B.GOGORO_BYTES_PER_ENV
    sampler = B.ClockSampler(local_rank)
    if not os.environ.get("TK_BENCH_NO_SAMPLER"):
        sampler.start()
    windows = []

    # ---- episode statistics: finalize on device, one 8 x fp64 NCCL all-reduce per horizon (side stream).
    # The hook is keyed on the TIMED step index and horizon = min(32, steps), so it fires inside every timed region
    # (the driver's --steps 20 included); the partials of every state set are discarded after the warm-up.
    from thormang_isaacgym_b200.parallel import EpisodeStatsReducer
    reducer = EpisodeStatsReducer(device)
    horizon = max(1, min(HORIZON, args.steps))
    hook_log = {"launches": 0, "expected_env_steps": None}

    def horizon_hook(env, steps_since, warm=False):
        # what a training loop does once per rollout horizon: finalize the episode statistics of the env on device
        # (one tiny kernel) and all-reduce the 8 doubles on a side stream.  The rotating state sets exist only to defeat
        # the L2: the set that has just stepped is finalized; it has run `steps_since` steps since its last finalize,
        # identically on every rank, which fixes the env-step count the reduced vector must carry.
        reducer.launch(env.episode_statistics())      # side stream; NCCL all-reduce when world > 1
        if warm:
            reducer.result()
            return 1
        hook_log["launches"] += 1
        hook_log["expected_env_steps"] = float(world) * n * steps_since
        return 1

    res = _kernel_bench(torch, task, n, device, args.steps, max(args.warmup, 3), seed=1234 + 1000 * rank, env_off=rank * n,
                        barrier=barrier, horizon_hook=None if os.environ.get("TK_BENCH_NO_HOOK") else horizon_hook,
                        horizon=horizon)
    windows.append(res["window"])
    ms_local = res["ms_total"]
    ms_t = torch.tensor([ms_local], device=device, dtype=torch.float64)
    per_rank_ms = [ms_local]
    if dist:
        gathered = [torch.zeros_like(ms_t) for _ in range(world)]
        dist.all_gather(gathered, ms_t)
        per_rank_ms = [float(g.item()) for g in gathered]
        dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
    ms_total = float(ms_t.item())
    value = n * world * args.steps / (ms_total * 1e-3)
    ms_per_step = ms_total / args.steps

    # the collective must have run, and the reduced vector must carry world x n x (steps since the set's last finalize)
    stats_last = reducer.result()
    stats_ran = hook_log["launches"]
    stats_ok = stats_ran > 0 and abs(stats_last["env_steps"] - hook_log["expected_env_steps"]) < 0.5
    if dist:
        ok_t = torch.tensor([1 if stats_ok else 0], device=device, dtype=torch.int32)
        dist.all_reduce(ok_t, op=dist.ReduceOp.MIN)
        stats_ok = bool(ok_t.item())

    multi = None
    if not os.environ.get("TK_BENCH_NO_DIAG"):
        multi = _multi_gpu_diag(torch, dist, task, n, device, rank, world, local_rank, barrier, args.steps)
        multi["per_rank_ms_total"] = per_rank_ms
        if dist:
            try:
                multi["pcie_concurrent"] = _pcie_probe_concurrent(torch, dist, device, world, barrier)
            except Exception as exc:
                multi["pcie_concurrent"] = {"error": repr(exc)}
        multi["rank0_host_enqueue_ms"] = res.get("host_enqueue_ms")

    # ---- end-to-end through the public API with host buffers
    e2e = None
    if task == "humanoid" and args.e2e_steps > 0:
        r = _e2e_bench(torch, n, device, args.e2e_steps, args.e2e_chunks, rank * n, barrier)
        windows.append(r["window"])
        sec = torch.tensor([r["seconds"]], device=device, dtype=torch.float64)
        if dist:
            dist.all_reduce(sec, op=dist.ReduceOp.MAX)
        e2e = {"value": n * world * args.e2e_steps / float(sec.item()), "unit": "env-steps/s",
               "h2d_bytes_per_step": r["h2d"] * world, "d2h_bytes_per_step": r["d2h"] * world,
               "steps": args.e2e_steps, "chunks_per_gpu": r["chunks"], "pinned_placement_rank0": r["placement"],
               "GBps_per_direction_per_gpu": (r["h2d"] + r["d2h"]) / 2 * args.e2e_steps / float(sec.item()) / 1e9,
               "path": "HostBridge: pinned host state slab -> one H2D -> Humanoid.post_physics_step() -> one D2H of the "
                       "obs|rew|reset slab, range-sharded into chunks on separate streams so PCIe copies overlap the kernels"}

    if dist and multi is not None and e2e is not None:
        # the same e2e path with 1 and 4 chunks per GPU (the headline uses --e2e-chunks): does finer pipelining help or hurt
        # once N ranks share the host side of the box?
        sweep = {}
        for ch in (1, 4):
            try:
                r2 = _e2e_bench(torch, n, device, max(10, args.e2e_steps // 2), ch, rank * n, barrier)
                sec2 = torch.tensor([r2["seconds"]], device=device, dtype=torch.float64)
                dist.all_reduce(sec2, op=dist.ReduceOp.MAX)
                sweep[f"chunks_{r2['chunks']}"] = n * world * max(10, args.e2e_steps // 2) / float(sec2.item())
            except Exception as exc:
                sweep[f"chunks_{ch}"] = repr(exc)
        sweep[f"chunks_{e2e['chunks_per_gpu']}_headline"] = e2e["value"]
        multi["e2e_env_steps_per_s_by_chunks"] = sweep
    sampler.stop()
    clocks = sampler.summary(windows)
    if rank != 0:
        if dist:
            dist.barrier(device_ids=[local_rank])
            dist.destroy_process_group()
        return

    peak, peak_src = B.measured_peak_gbs()
    achieved = n * bpe / (ms_local / args.steps * 1e-3) / 1e9
    line = {
        "metric": B.METRIC, "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": B.workload_config(args, res["sets"]),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": B.recorded_traffic(task)[0], "traffic_note": B.recorded_traffic(task)[1],
                     "peak_source": peak_src,
                     "kernel": "humanoid_warp_kernel<Spec<21,12,1,0>,true>" if task == "humanoid" else "gogoro_tile_kernel<0>",
                     "algorithmic_bytes_per_env_step": bpe,
                     "how": "algorithmic bytes x envs per launch / (CUDA-event time of the timed region / launches), rank 0",
                     "not_counted": "the launch also writes up_vec / heading_vec (24 B per env: SURVEY 8(a) H2 '+24 B if written'; the "
                                    "reference assigns them every step, only its debug drawing reads them), the timeout mask (1 B) and "
                                    "reads + writes the running episode return (8 B) for the statistics: 33 B per env-step on top of "
                                    "the 924 the fraction is computed from; extras *_no_debug_vectors time the launches without the 24"},
        "e2e": e2e,
        "gpu_launches": res["launches"],
        "clocks": clocks,
        "stats_allreduce": {"per": f"{horizon} timed steps", "payload_bytes": 64, "launched_in_timed_region": stats_ran,
                            "backend": "nccl" if dist else "none (1 GPU: finalize kernel + side-stream copy, no collective)",
                            "expected_env_steps": hook_log["expected_env_steps"],
                            "sum_over_ranks_ok": bool(stats_ok) if dist else None,
                            "local_sum_ok": bool(stats_ok) if not dist else None,
                            "check": "reduced env_steps == world x envs_per_gpu x steps since the finalized set's last "
                                     "finalize, on every rank (MIN over ranks); fails when no all-reduce ran",
                            "last": stats_last},
    }
    if multi is not None:
        line["multi_gpu"] = multi

    if world == 1 and not args.no_cpu_baseline:
        v, ms_cpu, done, threads = B.time_cpu(task, n, 1234, 400, 3, budget_s=args.cpu_seconds)
        line["cpu_baseline"] = {"value": v, "unit": "env-steps/s", "cores": threads, "kind": "port",
                                "ms_per_step": ms_cpu,
                                "sample": f"{done} steps x {n} envs of the same post-physics path (reset+obs+reward), "
                                          "oracle/ torch port TorchScript-compiled like the reference, "
                                          f"torch.set_num_threads({threads})"}
    if world == 1 and not args.no_extras:
        extras = {}
        try:
            for label, t, m, gr in (("config2_humanoid_16384", "humanoid", 16384, False),
                                    ("config2_humanoid_16384_cuda_graph", "humanoid", 16384, True),
                                    ("humanoid_32768", "humanoid", 32768, False),
                                    ("humanoid_65536", "humanoid", 65536, False),
                                    ("humanoid_65536_cuda_graph", "humanoid", 65536, True),
                                    ("humanoid_65536_no_debug_vectors", "humanoid", 65536, "vec"),
                                    ("humanoid_131072_no_debug_vectors", "humanoid", 131072, "vec"),
                                    ("humanoid_262144", "humanoid", 262144, False),
                                    ("humanoid_1048576", "humanoid", 1048576, False),
                                    ("config3_gogoro_65536", "gogoro", 65536, False),
                                    ("config3_gogoro_65536_cuda_graph", "gogoro", 65536, True),
                                    ("gogoro_4194304", "gogoro", 4194304, False),
                                    ("n2_ant_65536", "ant", 65536, False),
                                    ("n2_ant_1048576", "ant", 1048576, False),
                                    ("n2_anymal_65536", "anymal", 65536, False),
                                    ("n2_anymal_1048576", "anymal", 1048576, False),
                                    ("n3_ma_op3_65536", "ma_op3", 65536, False),
                                    ("n3_ma_op3_524288", "ma_op3", 524288, False),
                                    ("n1_gogoro_paper_4096", "gogoro_paper", 4096, False),
                                    ("n1_gogoro_paper_65536", "gogoro_paper", 65536, False),
                                    ("n1_gogoro_paper_262144", "gogoro_paper", 262144, False),
                                    ("n1_gogoro_paper_1048576", "gogoro_paper", 1048576, False)):
                k = 300 if m <= 262144 else 60
                if gr == "vec":
                    r = _kernel_bench(torch, t, m, device, k, 20, seed=77, cfg_patch={"debug_vectors": False})
                else:
                    r = _kernel_bench(torch, t, m, device, k, 20, seed=77, graph=gr)
                k = r["steps"]
                b = B.BYTES_PER_ENV[t]
                gbs = m * b / (r["ms_total"] / k * 1e-3) / 1e9
                extras[label] = {"env_steps_per_s": m * k / (r["ms_total"] * 1e-3), "us_per_step": 1e3 * r["ms_total"] / k,
                                 "achieved_GBps": gbs, "frac_of_peak": gbs / peak, "rotating_sets": r["sets"]}
            extras["launch_floor"] = _launch_floor(torch, device)
            extras["config3_gogoro_65536_l2_resident"] = _l2_resident(torch, "gogoro", 65536, device, peak)
            extras["config2_humanoid_16384_l2_resident"] = _l2_resident(torch, "humanoid", 16384, device, peak)
            extras["config1_humanoid_4096"] = _config1(torch, device)
            extras["config5_step_loop_131072x32"] = _step_loop_bench(torch, 131072, device, 32, 8)
            extras["config5_step_loop_4096x32_cuda_graph"] = _step_loop_bench(torch, 4096, device, 32, 20, graph=True)
            extras["config5_step_loop_4096x32"] = _step_loop_bench(torch, 4096, device, 32, 20)
            extras["pcie"] = _pcie_probe(torch, device)
        except Exception as exc:   # extras must never cost the headline line
            extras["error"] = repr(exc)
        line["extras"] = extras
    print(json.dumps(line), flush=True)
    if dist:
        dist.barrier(device_ids=[local_rank])
        dist.destroy_process_group()
