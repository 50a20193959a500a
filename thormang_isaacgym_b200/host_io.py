"""Host <-> device staging for callers whose simulator state lives in host memory.

The reference's CPU pipeline (``sim_device="cpu"``, vec_task.py:338-339 ``fetch_results``) keeps the
state tensors on the host.  ``HostBridge`` is the public way to drive the CUDA env-step path from
such a caller: pinned staging buffers, asynchronous copies on the task's stream, no allocation per
step.  It is also the path ``bench.py`` times for its end-to-end (``e2e``) number: every step copies
the step's inputs host->device, runs the fused post-physics kernel and copies the results back.

Two things keep the PCIe link busy instead of waiting on per-copy overheads:

* slabs -- ``make_humanoid_for_host`` builds the task over a slab-backed fake sim
  (root | dof_state | dof_force | sensors | actions in ONE device allocation) and an output slab
  (obs | rew | reset), so a step is one H2D copy, one kernel, one D2H copy;
* chunks -- the env range is split into range shards, each a task object on its own stream, so the
  upload of shard c+1 overlaps the kernel and the download of shard c (the path has no cross-env
  dependence: range shards reproduce the unsharded results bit for bit, tests/test_gpu_humanoid.py).
"""
from typing import Callable, Dict, List

import torch

import contextlib
import os

HUMANOID_INPUTS = ("root_states", "dof_state", "dof_force", "sensors", "actions")
HUMANOID_OUTPUTS = ("obs_buf", "rew_buf", "reset_buf")


def make_humanoid_for_host(n: int, device, env_id_offset: int = 0, seed: int = 0, statistics: bool = False):
    """Humanoid task whose uploaded state and downloaded results each live in one device allocation."""
    from . import config as tcfg
    from .sim.fake_sim import FakeSim
    from .synthetic import humanoid_asset
    from .tasks.humanoid import Humanoid
    cfg = tcfg.humanoid_cfg(n, seed=seed)
    cfg["env"]["envIdOffset"] = env_id_offset
    cfg["env"]["episodeStatistics"] = statistics
    cfg["env"]["outputSlab"] = True
    cfg["sim_backend"] = FakeSim(n, 21, device, num_sensor_floats=12, asset=humanoid_asset(), slab=True)
    dev = torch.device(device)
    return Humanoid(cfg, str(dev), f"cuda:{dev.index or 0}", -1, True)


def gpu_local_cpus(device_index: int):
    """CPUs of the NUMA node the GPU hangs off (NVML), or None when the box does not say (no NVML, one node, a VM that
    hides the topology).  Returns (cpu list, numa node or None)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        words = (os.cpu_count() + 63) // 64
        try:      # node scope where the driver offers it, else the plain affinity mask
            mask = pynvml.nvmlDeviceGetCpuAffinityWithinScope(h, words, pynvml.NVML_AFFINITY_SCOPE_NODE)
        except Exception:
            mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = [64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1]
        node = None
        try:
            node = int(pynvml.nvmlDeviceGetNumaNodeId(h))
        except Exception:
            pass
        return (cpus or None), node
    except Exception:
        return None, None


@contextlib.contextmanager
def near_gpu(device_index: int, report: dict = None):
    """Run the enclosed allocations on the CPUs next to the GPU: pinned slabs are placed on the NUMA node of the thread
    that allocates them, and a slab on the far socket crosses the inter-socket link on every copy.  The affinity of the
    process is restored on exit (the env-step path itself is not CPU-bound)."""
    old = None
    cpus, node = gpu_local_cpus(device_index)
    try:
        if cpus and hasattr(os, "sched_setaffinity"):
            old = os.sched_getaffinity(0)
            usable = sorted(set(cpus) & set(old)) or None
            if usable and set(usable) != set(old):
                os.sched_setaffinity(0, usable)
            else:
                old = None
        if report is not None:
            report.update({"gpu": device_index, "numa_node": node, "gpu_local_cpus": len(cpus) if cpus else None,
                           "affinity_narrowed_for_allocation": old is not None})
        yield
    finally:
        if old is not None:
            os.sched_setaffinity(0, old)


class HostBridge:
    """Pinned staging + async copies for one Humanoid task object."""

    def __init__(self, env, stream: torch.cuda.Stream = None):
        self.env = env
        self.stream = stream or torch.cuda.Stream(device=env.device)
        sim = env.sim
        self.dev_in: Dict[str, torch.Tensor] = {
            "root_states": env.root_states, "dof_state": env.dof_state, "dof_force": env.dof_force_tensor,
            "sensors": env.vec_sensor_tensor, "actions": env.actions}
        self.dev_out: Dict[str, torch.Tensor] = {k: getattr(env, k) for k in HUMANOID_OUTPUTS}
        self.in_slab = getattr(sim, "state_slab", None)
        self.out_slab = getattr(env, "output_slab", None)
        if self.in_slab is not None:
            self.host_in_slab = torch.empty(self.in_slab.shape, dtype=self.in_slab.dtype, pin_memory=True)
            self.host_in = {k: self.host_in_slab[off:off + size].view(self.dev_in[k].shape)
                            for k, (off, size) in sim.slab_layout.items()}
            self.h2d_bytes = self.in_slab.numel() * self.in_slab.element_size()
        else:
            self.host_in = {k: torch.empty(v.shape, dtype=v.dtype, pin_memory=True) for k, v in self.dev_in.items()}
            self.h2d_bytes = sum(v.numel() * v.element_size() for v in self.host_in.values())
        if self.out_slab is not None:
            self.host_out_slab = torch.empty(self.out_slab.shape, dtype=torch.uint8, pin_memory=True)
            base = self.out_slab.data_ptr()
            self.host_out = {}
            for k, v in self.dev_out.items():
                off = v.data_ptr() - base
                nbytes = v.numel() * v.element_size()
                self.host_out[k] = self.host_out_slab[off:off + nbytes].view(v.dtype).view(v.shape)
            self.d2h_bytes = self.out_slab.numel()
        else:
            self.host_out = {k: torch.empty(v.shape, dtype=v.dtype, pin_memory=True) for k, v in self.dev_out.items()}
            self.d2h_bytes = sum(v.numel() * v.element_size() for v in self.host_out.values())

    def fill_host(self, state: Dict[str, torch.Tensor]):
        """Copy a host-side snapshot (synthetic.humanoid_state keys) into the pinned input buffers."""
        for k in HUMANOID_INPUTS:
            self.host_in[k].copy_(state[k].view(self.host_in[k].shape))

    def step(self):
        """H2D of the inputs, fused post-physics kernel, D2H of obs / reward / reset -- all async on ``stream``."""
        with torch.cuda.stream(self.stream):
            if self.in_slab is not None:
                self.in_slab.copy_(self.host_in_slab, non_blocking=True)
            else:
                for k, dst in self.dev_in.items():
                    dst.copy_(self.host_in[k], non_blocking=True)
            self.env.post_physics_step()
            if self.out_slab is not None:
                self.host_out_slab.copy_(self.out_slab, non_blocking=True)
            else:
                for k, src in self.dev_out.items():
                    self.host_out[k].copy_(src, non_blocking=True)

    def synchronize(self):
        self.stream.synchronize()


class ChunkedHostBridge:
    """Range-shard ``n`` envs into ``chunks`` task objects, one stream each, so copies overlap the kernels."""

    def __init__(self, make_env: Callable[[int, int], object], n: int, chunks: int):
        assert n % chunks == 0, "env count must divide into the chunks"
        self.per = n // chunks
        self.placement: dict = {}
        envs = [make_env(self.per, c * self.per) for c in range(chunks)]
        dev = torch.device(envs[0].device)
        with near_gpu(dev.index or 0, self.placement):      # pinned slabs on the GPU's own NUMA node
            self.bridges: List[HostBridge] = [HostBridge(e) for e in envs]
        self.h2d_bytes = sum(b.h2d_bytes for b in self.bridges)
        self.d2h_bytes = sum(b.d2h_bytes for b in self.bridges)

    def fill_host(self, state: Dict[str, torch.Tensor], num_dof: int = 21):
        for c, b in enumerate(self.bridges):
            lo, hi = c * self.per, (c + 1) * self.per
            sub = {k: (state[k][lo * num_dof:hi * num_dof] if k == "dof_state" else state[k][lo:hi]) for k in HUMANOID_INPUTS}
            b.fill_host(sub)

    def step(self):
        for b in self.bridges:
            b.step()

    def synchronize(self):
        for b in self.bridges:
            b.synchronize()
