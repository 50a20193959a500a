"""Task registry, same shape as ``isaacgymenvs/tasks/__init__.py:54-78`` for the two hot-path tasks."""
from .gogoro_new import Gogoro
from .humanoid import Humanoid

isaacgym_task_map = {
    "Humanoid": Humanoid,
    "Gogoro": Gogoro,
}
