"""Task registry, same shape as ``isaacgymenvs/tasks/__init__.py:54-78`` for the hot-path tasks."""
from .ant import Ant
from .anymal import Anymal
from .gogoro_new import Gogoro
from .gogoro_realistic_turning import Gogoro as GogoroTurning
from .gogoro_realistic_turning_sim_paper import Gogoro as GogoroPaper
from .humanoid import Humanoid
from .MA_OP3 import MA_OP3

isaacgym_task_map = {
    "Humanoid": Humanoid,
    "Ant": Ant,
    "Anymal": Anymal,
    "Gogoro": Gogoro,
    "MA_OP3": MA_OP3,
    "GogoroPaper": GogoroPaper,   # gogoro_realistic_turning_sim_paper.py::Gogoro (unregistered in the reference)
    "GogoroTurning": GogoroTurning,   # gogoro_realistic_turning.py::Gogoro (unregistered in the reference)
}
