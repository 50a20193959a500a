"""world_size-2 gloo test (CPU) of the N > 1 host logic: env-range sharding, global-env-id offsets and the
episode-statistics all-reduce.  The per-rank math is the CPU oracle (this is a test; the product kernels need a GPU):
sharded outputs concatenated == unsharded outputs bit for bit, reduced statistics == single-process sums."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from thormang_isaacgym_b200 import parallel as par


def test_shard_range_partitions_exactly():
    for total in (1, 7, 4096, 1_000_003):
        for world in (1, 2, 3, 8):
            edges = [par.shard_range(total, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == total
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in edges]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        par.shard_range(10, 2, 2)
    cfg = {"env": {"numEnvs": 1048576}}
    par.shard_cfg(cfg, 3, 8)
    # world > 1 also switches the two in-step exchanges on (MA_OP3 gripper counts, paper variants' cross-env mean)
    assert cfg["env"] == {"numEnvs": 131072, "envIdOffset": 393216, "gripAllReduce": True, "worldReduce": True}
    one = par.shard_cfg({"env": {"numEnvs": 4096}}, 0, 1)
    assert one["env"] == {"numEnvs": 4096, "envIdOffset": 0}


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _humanoid_step(state, asset, noise, lo, hi):
    from oracle import humanoid_oracle as ho
    from thormang_isaacgym_b200 import config as tcfg
    n = hi - lo
    d = 21
    sub = {k: (v[lo * d:hi * d].clone() if k == "dof_state" else v[lo:hi].clone()) for k, v in state.items()}
    orc = ho.HumanoidOracle(ho.HumanoidConsts.from_cfg(tcfg.humanoid_cfg(n)), sub["root_states"], sub["dof_state"],
                            sub["dof_force"], sub["sensors"], asset["dof_limits_lower"], asset["dof_limits_upper"],
                            asset["motor_efforts"])
    orc.actions, orc.progress_buf, orc.reset_buf, orc.potentials = (sub["actions"], sub["progress_buf"], sub["reset_buf"],
                                                                    sub["potentials"])
    # dense noise rows by GLOBAL env id -> compact rows for this shard (what Philox keyed by global id gives each env)
    ids = orc.reset_buf.nonzero().flatten()
    ua, ub = noise[0][lo:hi][ids], noise[1][lo:hi][ids]
    orc.post_physics_step(ua if len(ids) else noise[0][:1], ub if len(ids) else noise[1][:1])
    timeout = (orc.progress_buf >= 999) & (orc.reset_buf != 0)
    stats = torch.tensor([orc.rew_buf.double().sum(), (orc.reset_buf != 0).sum(), timeout.sum(), 0.0,
                          orc.progress_buf[orc.reset_buf != 0].sum(), orc.obs_buf[:, 10].double().sum(),
                          orc.obs_buf[:, 11].double().sum(), n], dtype=torch.float64)
    return orc, stats


def _worker(rank, world, port, total, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    from thormang_isaacgym_b200 import synthetic as syn
    asset = syn.humanoid_asset()
    state = syn.humanoid_state(total, seed=1234, asset=asset)
    noise = syn.humanoid_reset_noise(total, seed=5)
    lo, hi = par.shard_range(total, rank, world)
    orc, stats = _humanoid_step(state, asset, noise, lo, hi)
    red = par.EpisodeStatsReducer("cpu")
    red.launch(stats)
    res = red.result()
    torch.save({"obs": orc.obs_buf, "rew": orc.rew_buf, "reset": orc.reset_buf, "dof": orc.dof_state, "stats": res,
                "range": (lo, hi)}, os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_and_allreduce(tmp_path):
    total, world = 1001, 2          # odd total: ragged shards
    port = _free_port()
    mp.spawn(_worker, args=(world, port, total, str(tmp_path)), nprocs=world, join=True)
    from thormang_isaacgym_b200 import synthetic as syn
    asset = syn.humanoid_asset()
    state = syn.humanoid_state(total, seed=1234, asset=asset)
    noise = syn.humanoid_reset_noise(total, seed=5)
    full, full_stats = _humanoid_step(state, asset, noise, 0, total)
    parts = [torch.load(os.path.join(str(tmp_path), f"rank{r}.pt"), weights_only=False) for r in range(world)]
    assert parts[0]["range"] == (0, 501) and parts[1]["range"] == (501, 1001)
    # torch's CPU kernels take a vector (SLEEF) or a scalar (libm) path depending on where a row falls in the
    # tensor, so even the reference is only ulp-level shard-invariant on the CPU; the CUDA kernels are exactly
    # shard-invariant (tests/test_gpu_humanoid.py::test_full_size_properties).
    import parity_util as U
    U.humanoid_obs_close(torch.cat([p["obs"] for p in parts]), full.obs_buf, "sharded obs")
    assert torch.equal(torch.cat([p["rew"] for p in parts]), full.rew_buf)
    assert torch.equal(torch.cat([p["reset"] for p in parts]), full.reset_buf)
    assert torch.equal(torch.cat([p["dof"] for p in parts]), full.dof_state)
    for p in parts:   # every rank holds the same, global, statistics
        got = torch.tensor([p["stats"][k] for k in par.STAT_NAMES], dtype=torch.float64)
        torch.testing.assert_close(got, full_stats, rtol=1e-9, atol=1e-4)
        assert got[1] == full_stats[1] and got[2] == full_stats[2] and got[4] == full_stats[4]
        assert p["stats"]["env_steps"] == total
        assert abs(p["stats"]["mean_reward"] - float(full.rew_buf.double().mean())) < 1e-9


def _op3_state(total):
    """One env of the whole batch presses each gripper of agent 0 (one per shard): only the all-reduced counts give 1."""
    from thormang_isaacgym_b200 import synthetic as syn
    st = syn.ma_op3_state(total, seed=4321)
    _, grip = syn.ma_op3_body_indices()
    cf = st["contact_forces"]
    cf[:, grip.flatten(), 0] = 0.0
    cf[3, grip[0, 0], 0] = 0.5                 # rank 0's shard
    cf[total - 2, grip[0, 1], 0] = 0.7         # rank 1's shard
    cf[5, grip[1, 0], 0] = 0.5
    cf[total - 5, grip[1, 0], 0] = 0.5         # agent 1: two left presses over the two shards -> 0
    cf[6, grip[1, 1], 0] = 0.5
    return st


def _op3_shard(st, total, lo, hi):
    from oracle import ma_op3_oracle as mo
    sub = {}
    for k, v in st.items():
        per = v.shape[0] // total
        sub[k] = v[lo * per:hi * per].clone()
    return mo.oracle_from_state(sub)


def _op3_worker(rank, world, port, total, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    from oracle import ma_op3_oracle as mo
    lo, hi = par.shard_range(total, rank, world)
    orc = _op3_shard(_op3_state(total), total, lo, hi)
    counts = mo.grip_counts(orc.contact_forces, orc.gripper_indices)       # what the fused kernel counts on this rank
    local = counts.clone()
    dist.all_reduce(counts)                                                 # MA_OP3._all_reduce_counts (SUM, 32 bytes)
    orc.global_grip_counts = counts
    orc.compute_observations()
    orc.compute_reward()
    torch.save({"rew": orc.rew_buf, "reset": orc.reset_buf, "local": local, "global": counts},
               os.path.join(out_dir, f"op3_rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_ma_op3_grip_count_exchange(tmp_path):
    """MA_OP3 is the one path with a real exchange step: _reward_gripper_contact_hold reduces over all
    envs (MA_OP3.py:346-358), so a sharded run must all-reduce the four counts before the reward."""
    total, world = 257, 2
    port = _free_port()
    mp.spawn(_op3_worker, args=(world, port, total, str(tmp_path)), nprocs=world, join=True)
    full = _op3_shard(_op3_state(total), total, 0, total)
    full.compute_observations()
    full.compute_reward()
    parts = [torch.load(os.path.join(str(tmp_path), f"op3_rank{r}.pt"), weights_only=False) for r in range(world)]
    assert parts[0]["local"].tolist() == [[1, 0], [1, 1]] and parts[1]["local"].tolist() == [[0, 1], [1, 0]]
    for p in parts:
        assert p["global"].tolist() == [[1, 1], [2, 1]]
    torch.testing.assert_close(torch.cat([p["rew"] for p in parts]), full.rew_buf, rtol=1e-6, atol=1e-6)
    assert torch.equal(torch.cat([p["reset"] for p in parts]), full.reset_buf)
    # without the exchange rank-local counts would give agent 0 no hold term at all: the +1 is visible in the rewards
    assert float((full.rew_buf[:, 0] > 0).float().mean()) > 0.5


def _init_worker(rank, world, port, out_dir):
    os.environ.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank), "WORLD_SIZE": str(world),
                       "LOCAL_RANK": str(rank)})
    r, w, device, seed = par.init_multi_gpu(seed=40)         # what make(multi_gpu=True) calls (train.py:74-82)
    from thormang_isaacgym_b200 import config as tcfg
    cfg = par.shard_cfg(tcfg.humanoid_cfg(1001, seed=40), r, w)
    t = torch.tensor([cfg["env"]["numEnvs"]], dtype=torch.int64)
    dist.all_reduce(t)                                       # the group init_multi_gpu created is usable
    torch.save({"rank": r, "world": w, "device": device, "seed": seed, "env": {k: cfg["env"][k] for k in
               ("numEnvs", "envIdOffset", "gripAllReduce", "worldReduce")}, "total": int(t)},
               os.path.join(out_dir, f"init{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_make_multi_gpu_host_logic(tmp_path):
    """make(multi_gpu=True) = LOCAL_RANK -> device, seed + rank (train.py:74-82) + range sharding by global env id."""
    world = 2
    port = _free_port()
    mp.spawn(_init_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    got = [torch.load(os.path.join(str(tmp_path), f"init{r}.pt"), weights_only=False) for r in range(world)]
    assert [g["device"] for g in got] == ["cuda:0", "cuda:1"]
    assert [g["seed"] for g in got] == [40, 41]
    assert got[0]["env"] == {"numEnvs": 501, "envIdOffset": 0, "gripAllReduce": True, "worldReduce": True}
    assert got[1]["env"] == {"numEnvs": 500, "envIdOffset": 501, "gripAllReduce": True, "worldReduce": True}
    assert all(g["total"] == 1001 and g["world"] == 2 for g in got)
