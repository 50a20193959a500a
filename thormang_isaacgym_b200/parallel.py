"""Multi-GPU plumbing: env-range sharding and the episode-statistics all-reduce.

The env-step path has no cross-env dependence: rank r of W owns the contiguous env-id
range ``[r * N/W, (r+1) * N/W)`` of every state and output tensor, task constants are replicated, the
in-kernel Philox noise is keyed by the GLOBAL env id (``cfg["env"]["envIdOffset"]``), so results do
not depend on W.  The only collective is a SUM all-reduce of the 8 x fp64 episode-statistics vector
(64 bytes, latency-bound), issued once per rollout horizon on a side stream so that it never sits on
the step's critical path.  Two members of the task family have a cross-env reward term and therefore one more,
IN-STEP, exchange each (a 16- / 32-byte all-reduce between the fused kernel and its finalize grid): the Gogoro paper
variants (``worldReduce``) and MA_OP3 (``gripAllReduce``); ``shard_cfg`` switches both on.  The reference has no
collective at all (``multi_gpu`` only picks the device
and offsets the seed: train.py:74-82); parity for the reduced vector is defined against the
single-process sum.

Works with any ``torch.distributed`` backend: ``nccl`` on the GPUs, ``gloo`` in the CPU tests.
"""
from typing import Optional, Tuple

import torch

STAT_NAMES = ("sum_reward", "num_reset", "num_timeout", "sum_return_of_finished", "sum_length_of_finished",
              "sum_aux0", "sum_aux1", "env_steps")


def shard_range(total_envs: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous env-id range [lo, hi) owned by ``rank``; sizes differ by at most one env."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of size {world}")
    base, rem = divmod(total_envs, world)
    lo = rank * base + min(rank, rem)
    hi = lo + base + (1 if rank < rem else 0)
    return lo, hi


def shard_cfg(cfg: dict, rank: int, world: int) -> dict:
    """Rewrite ``cfg["env"]["numEnvs"]`` / ``envIdOffset`` in place for this rank's shard (global numEnvs in).

    With ``world > 1`` the two in-step exchanges of the task family are switched on as well, so that rewards do not
    depend on ``world``: MA_OP3's batch-global gripper counts (``gripAllReduce``, MA_OP3.py:346-358) and the paper
    variants' cross-env reward mean (``worldReduce``, gogoro_realistic_turning_sim_paper.py:747).  Tasks without a
    cross-env term ignore the keys."""
    lo, hi = shard_range(cfg["env"]["numEnvs"], rank, world)
    cfg["env"]["numEnvs"] = hi - lo
    cfg["env"]["envIdOffset"] = lo
    if world > 1:
        cfg["env"].setdefault("gripAllReduce", True)
        cfg["env"].setdefault("worldReduce", True)
    return cfg


def init_multi_gpu(seed: int = 0):
    """What ``train.py:74-82`` does for ``multi_gpu=True``: pick ``cuda:LOCAL_RANK`` and offset the seed by the rank
    -- plus the process group the stats / in-step all-reduces need.  Returns (rank, world, device string, seed)."""
    import os
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    device = f"cuda:{local_rank}"
    if world > 1 and dist.is_available() and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if torch.cuda.is_available():
            torch.cuda.set_device(local_rank)
            dist.init_process_group("nccl", device_id=torch.device(device))
        else:                                   # CPU tests of the host logic (gloo); tasks themselves need CUDA
            dist.init_process_group("gloo")
    return rank, world, device, seed + rank     # train.py:80-82: cfg.seed += rank


class EpisodeStatsReducer:
    """SUM all-reduce of the per-rank episode statistics, asynchronous w.r.t. the compute stream."""

    def __init__(self, device, group=None):
        self.device = torch.device(device)
        self.group = group
        self.buffer = torch.zeros(len(STAT_NAMES), dtype=torch.float64, device=self.device)
        self._stream = torch.cuda.Stream(device=self.device) if self.device.type == "cuda" else None
        self._work = None
        self._copied = None

    @staticmethod
    def _dist():
        import torch.distributed as dist
        return dist if (dist.is_available() and dist.is_initialized()) else None

    def launch(self, local_stats: torch.Tensor):
        """Start reducing ``local_stats`` (fp64[8] on ``device``); returns immediately."""
        dist = self._dist()
        if self._stream is not None:
            # the caller's stream must not overwrite local_stats before the side stream has copied it: the copy of the
            # PREVIOUS launch is awaited here (one horizon later: never a real wait), this launch's copy is marked below
            if self._copied is not None:
                torch.cuda.current_stream(self.device).wait_event(self._copied)
            ready = torch.cuda.Event()
            ready.record()
            with torch.cuda.stream(self._stream):
                self._stream.wait_event(ready)
                self.buffer.copy_(local_stats)
                self._copied = torch.cuda.Event()
                self._copied.record()
                if dist is not None:
                    self._work = dist.all_reduce(self.buffer, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        else:
            self.buffer.copy_(local_stats)
            if dist is not None:
                self._work = dist.all_reduce(self.buffer, op=dist.ReduceOp.SUM, group=self.group, async_op=True)

    def result(self) -> dict:
        """Wait for the last ``launch`` and return the global statistics as a dict of Python floats."""
        if self._work is not None:
            self._work.wait()
            self._work = None
        if self._stream is not None:
            self._stream.synchronize()
        vals = self.buffer.cpu().tolist()
        out = dict(zip(STAT_NAMES, vals))
        steps = max(out["env_steps"], 1.0)
        done = max(out["num_reset"], 1.0)
        out["mean_reward"] = out["sum_reward"] / steps
        out["mean_episode_return"] = out["sum_return_of_finished"] / done
        out["mean_episode_length"] = out["sum_length_of_finished"] / done
        return out
