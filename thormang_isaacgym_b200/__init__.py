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
    """isaacgymenvs/__init__.py:14-55: build the named task with the stock cfg (or ``cfg``)."""
    from .tasks import isaacgym_task_map
    if cfg is None:
        cfg = {"Humanoid": config.humanoid_cfg, "Gogoro": config.gogoro_cfg, "Ant": config.ant_cfg,
               "GogoroPaper": config.gogoro_paper_cfg, "Anymal": config.anymal_cfg, "MA_OP3": config.ma_op3_cfg}[task](num_envs, seed)
    else:
        cfg["env"]["numEnvs"] = num_envs
    return isaacgym_task_map[task](cfg, rl_device, sim_device, graphics_device_id, headless,
                                   virtual_screen_capture, force_render)
