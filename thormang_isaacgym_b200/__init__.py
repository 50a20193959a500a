"""thormang_isaacgym_b200 -- B200-native env-step hot path of Yougo-robotics/thormang_isaacgym.

``make`` mirrors ``isaacgymenvs.make`` (isaacgymenvs/__init__.py:14-55) for the two tasks whose
post-physics / pre-physics path is implemented here as sm_100a CUDA kernels: ``Humanoid`` and the
registered ``Gogoro`` scooter.  Importing the package does not load the CUDA extension; creating a
task does, and raises if it is missing (no CPU fallback).
"""
from . import config

__all__ = ["make", "config"]


def make(seed: int, task: str, num_envs: int, sim_device: str = "cuda:0", rl_device: str = "cuda:0",
         graphics_device_id: int = -1, headless: bool = True, multi_gpu: bool = False,
         virtual_screen_capture: bool = False, force_render: bool = False, cfg: dict = None):
    """isaacgymenvs/__init__.py:14-55: build the named task with the stock cfg (or ``cfg``).

    ``multi_gpu=True`` (isaacgymenvs/__init__.py:22,51 + train.py:74-82): under ``torchrun`` the task is built on
    ``cuda:LOCAL_RANK`` with ``seed + rank`` like the reference, and -- what the reference leaves to PhysX running N
    independent copies -- ``num_envs`` is read as the GLOBAL env count and range-sharded: this rank owns env ids
    ``[lo, hi)`` (``parallel.shard_cfg``), the in-kernel noise is keyed by the global env id, and the in-step
    exchanges of the tasks that have a cross-env term are switched on.  The in-kernel noise seed stays the job-level
    ``seed`` so that results do not depend on the number of ranks."""
    from .tasks import isaacgym_task_map
    makers = {"Humanoid": config.humanoid_cfg, "Gogoro": config.gogoro_cfg, "Ant": config.ant_cfg,
              "GogoroPaper": config.gogoro_paper_cfg, "GogoroTurning": config.gogoro_turning_cfg,
              "Anymal": config.anymal_cfg, "MA_OP3": config.ma_op3_cfg}
    if cfg is None:
        cfg = makers[task](num_envs, seed)
    else:
        cfg["env"]["numEnvs"] = num_envs
    if multi_gpu:
        from . import parallel
        rank, world, device, rank_seed = parallel.init_multi_gpu(seed)
        sim_device = rl_device = device
        parallel.shard_cfg(cfg, rank, world)
        cfg["rank_seed"] = rank_seed               # train.py:80-82; cfg["seed"] keys the kernels
        # set_seed(seed + rank) of train.py:80-82: the host-side torch draws (reset_idx without supplied draws, the DR
        # `corr` tensor) must differ between ranks like the reference's
        import torch
        torch.manual_seed(rank_seed)
        if torch.cuda.is_available():
            torch.cuda.manual_seed_all(rank_seed)
        import logging
        logging.getLogger("thormang_isaacgym_b200").info(
            "multi_gpu: rank %d/%d on %s owns envs [%d, %d) of the GLOBAL num_envs=%d (the reference reads num_envs per rank); "
            "host RNG seed %d", rank, world, device, cfg["env"].get("envIdOffset", 0),
            cfg["env"].get("envIdOffset", 0) + cfg["env"]["numEnvs"], num_envs, rank_seed)
    return isaacgym_task_map[task](cfg, rl_device, sim_device, graphics_device_id, headless,
                                   virtual_screen_capture, force_render)
