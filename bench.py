#!/usr/bin/env python
"""Benchmark of the env-step hot path (BASELINE.json: env-steps/sec of obs+reward+reset, % of HBM peak).

    python bench.py [--gpus N --steps K --warmup W] [--impl reference] [--envs-per-gpu E] [--task humanoid|gogoro]

One "step" = one pass of the fused post-physics kernel (humanoid.py:287-296 + vec_task.py:345: progress++, reset of
flagged envs with in-kernel Philox noise, observations, reward, reset/timeout masks, statistics partials) over one batch
of synthetic env states.  Default workload: the per-GPU shard of BASELINE config 4 (Humanoid, 131072 envs per GPU; N = 8
is exactly the 1M-env configuration the metric is quoted on, N = 1 the same shard on one B200).  Timed steps rotate over
disjoint state sets whose total exceeds twice the 126 MB L2, so every step streams from HBM.

Prints ONE JSON line (contract in the task description): value = whole-job env-steps/s with inputs resident in HBM;
e2e = the same metric through the public task API with HOST buffers (H2D of the state, kernel, D2H of obs/reward/reset
inside the timed region); roofline for the fused kernel; cpu_baseline = the CPU oracle port timed on this box's cores.
``--impl reference`` times the reference's own CPU path (the torch restatement in oracle/, TorchScript-compiled like the
reference) on all host threads.
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

HUMANOID_BYTES_PER_ENV = 924     # 456 B read + 468 B written per env-step, fused post-physics
GOGORO_BYTES_PER_ENV = 172       # 192 B with supplied noise, 172 B with in-kernel RNG
# SURVEY.md 8(f) N1, 8-channel x 20-frame Gogoro variant: fused kernel 780 B read (root 52, 6 scalars 32,
# progress/reset 16, command ring 80, 19 noisy frames 608) + 705 B written (obs row 640, clean frame 32, command 4, reward 4,
# reset/progress 16, timeout 1, speed 4, yaw command 4); finalize kernel 16 B read + 8 B written
GOGORO_PAPER_BYTES_PER_ENV = 1509
# SURVEY.md 8(f) N2 Anymal: reads root 52 + dof_state 96 + torques 48 + contact 156 + commands 12 + progress 8 +
# reset 8; writes obs 244 + reward 4 + reset 8 + progress 8 + timeout 1
ANYMAL_BYTES_PER_ENV = 645
# SURVEY 8(f) N2 Ant (ant.py:213-287,318-408): reads root 52 + dof_state 64 + sensors 96 + actions 32 + potentials 8 + progress 8 +
# reset 8 = 268; writes obs 240 + potentials 8 + reward 4 + reset 8 + progress 8 = 268 (+ timeout 1, episode return 8, up / heading 24)
ANT_BYTES_PER_ENV = 536
MA_OP3_BYTES_PER_ENV = 2545      # csrc/tk_ma_op3.cu header: 1404 B read + 1141 B written per env-step (two robots + table)
BYTES_PER_ENV = {"ant": ANT_BYTES_PER_ENV, "ma_op3": MA_OP3_BYTES_PER_ENV, "anymal": ANYMAL_BYTES_PER_ENV, "humanoid": HUMANOID_BYTES_PER_ENV, "gogoro": GOGORO_BYTES_PER_ENV, "gogoro_paper": GOGORO_PAPER_BYTES_PER_ENV}
L2_BYTES = 126 * 1024 * 1024
METRIC = "env-steps/sec (obs+reward+reset)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=600)
    ap.add_argument("--warmup", type=int, default=30)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--task", default="humanoid", choices=["humanoid", "gogoro"])
    ap.add_argument("--envs-per-gpu", type=int, default=131072)
    ap.add_argument("--e2e-steps", type=int, default=40)
    ap.add_argument("--e2e-chunks", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the extra (non-headline) configurations")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    return ap.parse_args()


def measured_peak_gbs():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md: 6.65 TB/s)"


def recorded_traffic(task):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu capture (profiles/), or None."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(path) as fh:
            doc = json.load(fh)
            return doc.get(task), doc.get(task + "_note")
    except Exception:
        return None, None


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed regions run."""

    def __init__(self, device_index: int, period_s: float = 0.004):
        self.samples = []
        self.period = period_s
        self._stop = threading.Event()
        self._thread = None
        self.window = None
        self.handle = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            # NVML indices follow PCI order; CUDA_VISIBLE_DEVICES may remap
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = device_index
            if vis:
                try:
                    idx = int(vis.split(",")[device_index])
                except Exception:
                    idx = device_index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.handle = None
            self.max_mhz = None

    def _reasons(self):
        nv = self.nv
        try:
            mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
        except Exception:
            mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
        names = []
        table = (("hw_slowdown", 0x8), ("sw_power_cap", 0x4), ("hw_thermal_slowdown", 0x40),
                 ("sw_thermal_slowdown", 0x20), ("hw_power_brake_slowdown", 0x80), ("sync_boost", 0x10),
                 ("applications_clocks_setting", 0x2))
        for name, bit in table:
            if mask & bit:
                names.append(name)
        return names

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM)
                self.samples.append((time.perf_counter(), mhz, self._reasons()))
            except Exception:
                pass
            time.sleep(self.period)

    def start(self):
        if self.handle is not None:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2.0)

    def summary(self, windows):
        """windows: list of (t0, t1) perf_counter intervals that were timed."""
        if self.handle is None or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0, "source": "nvml unavailable"}
        inside = [s for s in self.samples if any(a <= s[0] <= b for a, b in windows)]
        used, where = (inside, "timed regions") if len(inside) >= 3 else (self.samples, "whole run (timed region shorter than the sampling period)")
        reasons = sorted({r for s in used for r in s[2]})
        return {"sm_mhz": statistics.median(s[1] for s in used), "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(used), "source": "nvml, " + where}


# ------------------------------------------------------------------------------------------------------------------
# CPU arm: the reference's torch CPU path (oracle port, TorchScript-compiled like the reference), all host threads
# ------------------------------------------------------------------------------------------------------------------
def build_cpu_humanoid(n, seed):
    import torch
    from oracle import humanoid_oracle as ho
    from thormang_isaacgym_b200 import config as tcfg
    from thormang_isaacgym_b200 import synthetic as syn
    asset = syn.humanoid_asset()
    st = syn.humanoid_state(n, seed=seed, asset=asset, fallen_frac=0.01)
    cfg = tcfg.humanoid_cfg(n, seed=seed)
    orc = ho.HumanoidOracle(ho.HumanoidConsts.from_cfg(cfg), st["root_states"], st["dof_state"], st["dof_force"],
                            st["sensors"], asset["dof_limits_lower"], asset["dof_limits_upper"], asset["motor_efforts"],
                            jit=True)
    orc.actions = st["actions"]
    orc.progress_buf = st["progress_buf"]
    orc.reset_buf = st["reset_buf"]
    orc.potentials = st["potentials"]
    noise = syn.humanoid_reset_noise(n, seed=seed)

    def step():
        k = int((orc.reset_buf != 0).sum())
        orc.post_physics_step(noise[0][:max(k, 1)], noise[1][:max(k, 1)])
        # vec_task.py:345 timeout line, part of the fused kernel on the GPU side
        orc.timeout_buf = (orc.progress_buf >= orc.c.max_episode_length - 1) & (orc.reset_buf != 0)
    return step


def build_cpu_gogoro(n, seed):
    from oracle import gogoro_oracle as go
    from thormang_isaacgym_b200 import config as tcfg
    from thormang_isaacgym_b200 import synthetic as syn
    st = syn.gogoro_state(n, seed=seed)
    cfg = tcfg.gogoro_cfg(n, seed=seed)
    orc = go.GogoroOracle(go.GogoroConsts.from_cfg(cfg), st["root_tensor"], st["state_dof"], st["thormang_pose"])
    for name in ("yaw_command", "curent_command", "curent_speed", "steer_offsets", "imu_offsets",
                 "curent_speed_offset", "action_history", "progress_buf", "reset_buf"):
        setattr(orc, name, st[name])
    nz = syn.gogoro_noise(n, seed=seed)

    def step():
        orc.post_physics_step(nz["sensor_unit"], nz["reset_unit"], nz["speed_unit"], nz["yaw_unit"])
        orc.timeout_buf = (orc.progress_buf >= orc.max_episode_length - 1) & (orc.reset_buf != 0)
    return step


def time_cpu(task, n, seed, steps, warmup, budget_s=None):
    """Returns (env_steps_per_s, ms_per_step, steps_done, threads)."""
    import torch
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    step = (build_cpu_humanoid if task == "humanoid" else build_cpu_gogoro)(n, seed)
    for _ in range(warmup):
        step()
    times = []
    t_start = time.perf_counter()
    for _ in range(steps):
        t0 = time.perf_counter()
        step()
        times.append(time.perf_counter() - t0)
        if budget_s is not None and time.perf_counter() - t_start > budget_s and len(times) >= 5:
            break
    total = sum(times)
    return n * len(times) / total, 1e3 * total / len(times), len(times), torch.get_num_threads()


def cpu_config1(n=4096, warmup=5, reps=50):
    """BASELINE config 1 as SURVEY 8(d) specifies it: Humanoid compute_observations + compute_reward (humanoid.py:219-251,
    323-417; no reset_idx) on `n` synthetic env states, CPU torch, TorchScript-compiled like the reference; all host
    threads AND one thread; `warmup` warm-ups, median of `reps`.  The functions are the pinned port in oracle/ (bit-for-bit
    equal to the reference's own JIT functions, tests/test_oracle_vs_reference.py) -- /root/reference does not exist on
    the GPU box."""
    import torch
    from oracle import humanoid_oracle as ho
    from thormang_isaacgym_b200 import config as tcfg
    from thormang_isaacgym_b200 import synthetic as syn
    asset = syn.humanoid_asset()
    st = syn.humanoid_state(n, seed=1234, asset=asset)
    orc = ho.HumanoidOracle(ho.HumanoidConsts.from_cfg(tcfg.humanoid_cfg(n, seed=1234)), st["root_states"], st["dof_state"],
                            st["dof_force"], st["sensors"], asset["dof_limits_lower"], asset["dof_limits_upper"],
                            asset["motor_efforts"], jit=True)
    orc.actions, orc.progress_buf, orc.reset_buf, orc.potentials = st["actions"], st["progress_buf"], st["reset_buf"], st["potentials"]
    out = {"envs": n, "what": "compute_observations + compute_reward (no reset), CPU torch " + torch.__version__,
           "kind": "port (oracle/, pinned bit-for-bit to the reference's JIT functions; the isaacgym.torch_utils helpers under them are an un-vendored dependency restated in oracle/isaac_helpers.py: parity unpinned there)", "warmup": warmup, "reps": reps,
           "torch": torch.__version__, "parallel_info": torch.__config__.parallel_info().strip().splitlines()[:8]}
    before = torch.get_num_threads()
    for label, threads in (("all_threads", os.cpu_count() or 1), ("single_thread", 1)):
        torch.set_num_threads(threads)
        for _ in range(warmup):
            orc.compute_observations(); orc.compute_reward()
        times = []
        for _ in range(reps):
            t0 = time.perf_counter()
            orc.compute_observations(); orc.compute_reward()
            times.append(time.perf_counter() - t0)
        med = statistics.median(times)
        out[label] = {"threads": torch.get_num_threads(), "median_ms": 1e3 * med, "env_steps_per_s": n / med,
                      "min_ms": 1e3 * min(times)}
    torch.set_num_threads(before)
    return out


def run_reference_arm(args, rank, world):
    """--impl reference: rank 0 alone times the CPU path on the same workload, bounded sample per step."""
    if rank != 0:
        return
    n = args.envs_per_gpu
    steps = max(1, args.steps)
    # bounded sample: each step is one pass over the rank-0 shard; cap the run at ~2 minutes of CPU work
    value, ms, done, threads = time_cpu(args.task, n, 1234, steps, max(1, min(args.warmup, 5)), budget_s=120.0)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": args.gpus,
        "steps": done, "warmup": max(1, min(args.warmup, 5)), "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, rotating_sets=max(3, -(-int(2.2 * L2_BYTES) // (n * (HUMANOID_BYTES_PER_ENV if args.task == "humanoid" else GOGORO_BYTES_PER_ENV))))),   # key for key the GPU arm's config
        "cpu_baseline": {"value": value, "unit": "env-steps/s", "cores": threads, "kind": "port",
                         "sample": f"{done} steps x {n} envs of the {args.task} post-physics path (reset+obs+reward), "
                                   "oracle/ torch port TorchScript-compiled like the reference, "
                                   f"torch.set_num_threads({threads})"},
        "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "the reference is Python/TorchScript and cannot travel to this box (isaacgym missing, sources not "
                "copied): its CPU path is timed through the oracle port, pinned bit-for-bit to the reference's own JIT "
                "functions in the build container (tests/golden).  The CPU arm steps ONE state set held in host memory; at "
                "N > 1 rank 0 alone times its own shard (envs_per_gpu envs per step) on all host threads: the ratio against "
                "the N-GPU line compares N GPUs with this one CPU pass",
    }
    print(json.dumps(line), flush=True)


def workload_config(args, rotating_sets, note=None):
    if args.task == "humanoid":
        wl = (f"Humanoid fused post_physics_step (reset+obs+reward+masks), {args.envs_per_gpu} envs per GPU "
              "(per-GPU shard of BASELINE config 4: 1M envs / 8 GPUs), D=21, 110 obs")
    else:
        wl = f"Gogoro scooter fused post_physics_step, {args.envs_per_gpu} envs per GPU (BASELINE config 3 shape), 6 obs"
    cfg = {"workload": wl, "envs_per_gpu": args.envs_per_gpu, "global_envs": args.envs_per_gpu * args.gpus,
           "parallelism": f"env-range sharding x{args.gpus}, no data-path collective; NCCL all-reduce of 8 fp64 episode statistics per horizon",
           "reset_noise": "in-kernel Philox4x32-10", "fallen_fraction": 0.01,
           "cache": f"{rotating_sets} disjoint state sets rotated per step (> 2x the 126 MB L2): every step streams from HBM"}
    if note:
        cfg["note"] = note
    return cfg


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    from bench_gpu import run_gpu_arm   # GPU arm lives next to this file (kept separate to keep both readable)
    run_gpu_arm(args, rank, world)


if __name__ == "__main__":
    main()
