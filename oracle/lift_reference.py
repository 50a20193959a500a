"""Lift the reference's own ``@torch.jit.script`` hot-path functions.  TEST INFRASTRUCTURE.

The reference task modules cannot be imported (top-level ``from isaacgym import
gymtorch, gymapi`` -- isaacgymenvs/tasks/humanoid.py:33-35, gogoro_new.py:7-9), but
their JIT functions are pure torch.  This module slices the decorated top-level
``def``s out of the reference files BY SOURCE at run time (nothing is copied into
the repository), writes them to a temporary module next to an import of the
restated helpers (``oracle.isaac_helpers``) and lets ``torch.jit.script`` compile
them -- TorchScript needs a real source file, hence the temp file.

Only usable where ``/root/reference`` exists (the build container).  It is used
(a) by ``tests/golden/make_golden.py`` to freeze golden vectors and (b) by
``tests/test_oracle_vs_reference.py`` (skipped when the reference is absent).
"""
import ast
import importlib.util
import os
import sys
import tempfile
from types import ModuleType
from typing import Dict, Iterable

REFERENCE_ROOT = os.environ.get("THORMANG_REFERENCE_ROOT", "/root/reference")

# file (relative to the reference root) -> top-level function names to lift
LIFT_TABLE: Dict[str, Iterable[str]] = {
    "isaacgymenvs/tasks/humanoid.py": (
        "compute_humanoid_reward",        # humanoid.py:323-376
        "compute_humanoid_observations",  # humanoid.py:379-417
        "compute_reward",                 # humanoid.py:425-472
    ),
    "isaacgymenvs/tasks/ant.py": (
        "compute_ant_reward",             # ant.py:318-364
        "compute_ant_observations",       # ant.py:367-408
    ),
    "isaacgymenvs/tasks/anymal.py": (
        "compute_anymal_reward",           # anymal.py:323-372
        "compute_reward",                  # anymal.py:375-396
        "compute_anymal_observations_V2",  # anymal.py:399-431
    ),
    "isaacgymenvs/tasks/MA_OP3.py": (
        "compute_objective_reward",        # MA_OP3.py:496-547
        "compute_op3_reward",              # MA_OP3.py:550-658
        "compute_op3_observations",        # MA_OP3.py:662-708
    ),
    "isaacgymenvs/tasks/gogoro_new.py": (
        "compute_gogoro_reward",          # gogoro_new.py:645-684
        "shortest_angle_distance",        # gogoro_new.py:687-689
        "compute_gogoro_observations",    # gogoro_new.py:692-723
    ),
    "isaacgymenvs/tasks/gogoro_realistic_turning_sim_paper.py": (   # SURVEY 8(f) N1, the 8-channel x 20-frame variant
        "compute_gogoro_reward",          # :713-762
        "shortest_angle_distance",        # :765-767
        "compute_gogoro_observations",    # :770-807
    ),
    "isaacgymenvs/tasks/gogoro_realistic_turning.py": (             # SURVEY 8(f) N1, the rational-reward sibling
        "compute_gogoro_reward",          # :713-765
        "shortest_angle_distance",        # :768-770
        "compute_gogoro_observations",    # :773-810
    ),
}

_PRELUDE = (
    "import math\n"
    "import numpy as np\n"
    "import torch\n"
    "import torch.nn.functional as F\n"
    "from torch import Tensor\n"
    "from typing import Tuple, Dict, List\n"
    "from oracle.isaac_helpers import *\n\n"
)


def reference_available() -> bool:
    return all(os.path.isfile(os.path.join(REFERENCE_ROOT, f)) for f in LIFT_TABLE)


def _slice_functions(path: str, names: Iterable[str]) -> str:
    with open(path, "r") as fh:
        text = fh.read()
    lines = text.splitlines(keepends=True)
    tree = ast.parse(text)
    wanted = set(names)
    chunks = []
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in wanted:
            first = min([node.lineno] + [d.lineno for d in node.decorator_list])
            chunks.append("".join(lines[first - 1:node.end_lineno]))
            wanted.discard(node.name)
    if wanted:
        raise RuntimeError(f"{path}: functions not found: {sorted(wanted)}")
    return "\n\n".join(chunks) + "\n"


_CACHE: Dict[str, ModuleType] = {}


def lift(rel_file: str) -> ModuleType:
    """Return a module holding the compiled reference functions of ``rel_file``."""
    if rel_file in _CACHE:
        return _CACHE[rel_file]
    if not reference_available():
        raise FileNotFoundError(f"reference tree not found under {REFERENCE_ROOT}")
    import torch
    # the reference forces the legacy executor (tasks/base/vec_task.py:198-200)
    torch._C._jit_set_profiling_mode(False)
    torch._C._jit_set_profiling_executor(False)

    body = _slice_functions(os.path.join(REFERENCE_ROOT, rel_file), LIFT_TABLE[rel_file])
    tmpdir = tempfile.mkdtemp(prefix="thormang_lift_")
    mod_name = "_lifted_" + os.path.basename(rel_file).replace(".py", "")
    mod_path = os.path.join(tmpdir, mod_name + ".py")
    with open(mod_path, "w") as fh:
        fh.write(_PRELUDE + body)
    repo_root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if repo_root not in sys.path:
        sys.path.insert(0, repo_root)
    spec = importlib.util.spec_from_file_location(mod_name, mod_path)
    module = importlib.util.module_from_spec(spec)
    sys.modules[mod_name] = module
    spec.loader.exec_module(module)
    _CACHE[rel_file] = module
    return module


def humanoid() -> ModuleType:
    return lift("isaacgymenvs/tasks/humanoid.py")


def ant() -> ModuleType:
    return lift("isaacgymenvs/tasks/ant.py")


def gogoro() -> ModuleType:
    return lift("isaacgymenvs/tasks/gogoro_new.py")


def gogoro_paper() -> ModuleType:
    return lift("isaacgymenvs/tasks/gogoro_realistic_turning_sim_paper.py")


def gogoro_turning() -> ModuleType:
    return lift("isaacgymenvs/tasks/gogoro_realistic_turning.py")


def anymal() -> ModuleType:
    return lift("isaacgymenvs/tasks/anymal.py")


def ma_op3() -> ModuleType:
    return lift("isaacgymenvs/tasks/MA_OP3.py")
