"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every declared symbol,
the ctypes structs match the header, host logic of the task classes that needs no GPU."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built_lib():
    from thormang_isaacgym_b200 import build
    return build.build()


def test_header_symbols_exported(built_lib):
    """Every function include/thormang_b200.h declares is exported by the .so and bound in _native.SYMBOLS."""
    from thormang_isaacgym_b200 import _native as nat
    header = open(os.path.join(ROOT, "include", "thormang_b200.h")).read()
    declared = set(re.findall(r"\b(tk_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 26
    assert declared == set(nat.SYMBOLS), declared ^ set(nat.SYMBOLS)
    lib = C.CDLL(built_lib)
    for name in declared:
        assert hasattr(lib, name), f"{name} not exported"
    assert nat.lib().tk_abi_version() == 1


def test_struct_layouts_match_header(built_lib):
    """Every struct of include/thormang_b200.h has a ctypes mirror in _native.py with the same sizeof and the same
    offsetof for every field (compiled with gcc from the header): all 16 structs of the boundary, found by parsing the
    header, so that a struct added there without a mirror fails here."""
    import subprocess
    import tempfile
    from thormang_isaacgym_b200 import _native as nat
    header = open(os.path.join(ROOT, "include", "thormang_b200.h")).read()
    names = re.findall(r"^typedef struct (Tk\w+) \{", header, re.M)
    assert len(names) == 16, names
    lines = []
    for name in names:
        mirror = getattr(nat, name[2:])                       # TkHumanoidParams -> nat.HumanoidParams
        lines.append(f'printf("{name} %zu", sizeof({name}));')
        for field, _ in mirror._fields_:
            lines.append(f'printf(" %zu", offsetof({name}, {field}));')
        lines.append('printf("\\n");')
    src = '#include <stdio.h>\n#include <stddef.h>\n#include "thormang_b200.h"\nint main(){\n' + "\n".join(lines) + "\nreturn 0;}\n"
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "s.c"), "w").write(src)
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), os.path.join(d, "s.c"), "-o", os.path.join(d, "s")])
        out = subprocess.check_output([os.path.join(d, "s")]).decode().splitlines()
    assert len(out) == len(names)
    for row in out:
        name, size, *offs = row.split()
        mirror = getattr(nat, name[2:])
        assert int(size) == C.sizeof(mirror), f"{name}: sizeof {size} in C vs {C.sizeof(mirror)} in ctypes"
        got = [getattr(mirror, f).offset for f, _ in mirror._fields_]
        assert [int(o) for o in offs] == got, f"{name}: field offsets differ"


def test_no_cpu_fallback():
    """A CPU sim_device is refused loudly; the product never routes through the oracle."""
    import thormang_isaacgym_b200 as pkg
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.make(0, "Humanoid", 8, sim_device="cpu", rl_device="cpu")
    for dirpath, _, files in os.walk(os.path.join(ROOT, "thormang_isaacgym_b200")):
        for f in files:
            if f.endswith(".py"):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, re.M), f"{f} imports the oracle"


def test_invalid_params_rejected_without_gpu(built_lib):
    from thormang_isaacgym_b200 import _native as nat
    lib = nat.lib()
    h = C.c_void_p()
    p = nat.HumanoidParams()
    p.num_envs, p.num_dof = -3, 21
    assert lib.tk_humanoid_create(C.byref(p), C.byref(h)) == -1
    assert b"num_envs" in lib.tk_last_error()
    assert lib.tk_humanoid_create(None, None) == -1
    g = nat.GogoroParams()
    g.num_envs, g.num_dof, g.steering_joint = 4, 39, 99
    assert lib.tk_gogoro_create(C.byref(g), C.byref(h)) == -1 and b"joint" in lib.tk_last_error()
    assert lib.tk_humanoid_post_physics(None, None, None) == -1
    assert lib.tk_gogoro_stats(None, None, 0, None) == -1


def test_configs_carry_reference_constants():
    from thormang_isaacgym_b200 import config
    h = config.humanoid_cfg(16)
    assert h["env"]["terminationHeight"] == 0.8 and h["env"]["episodeLength"] == 1000 and h["sim"]["dt"] == 0.0166
    g = config.gogoro_cfg(16)
    assert g["noises"]["speed_range"] == [4.0, 13.0] and g["env"]["max_steps"] == 1000 and len(g["joints_pos"]) == 31
