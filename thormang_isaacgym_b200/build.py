"""In-tree build of the CUDA extension (explicit nvcc, sm_100a only).

    python -m thormang_isaacgym_b200.build [--force]

Produces ``thormang_isaacgym_b200/libthormang_b200.so`` -- a plain C-ABI shared library
(include/thormang_b200.h) loaded with ctypes.  The .so is git-ignored but travels to the
GPU box with the repo snapshot.  -fmad=false: see csrc/tk_common.cuh (floating-point policy).
"""
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libthormang_b200.so")
SOURCES = ["tk_common.cu", "tk_humanoid.cu", "tk_gogoro.cu", "tk_gogoro_paper.cu", "tk_anymal.cu", "tk_ma_op3.cu", "tk_epilogue.cu"]
HEADERS = ["tk_common.cuh", "tk_humanoid_tile.cuh", "tk_humanoid_quad.cuh", os.path.join("..", "..", "include", "thormang_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17", "-fmad=false",
    "-Xcompiler", "-fPIC", 
    "-shared", "-cudart", "static",
]


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.isfile(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def _stale() -> bool:
    if not os.path.isfile(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES] + [os.path.normpath(os.path.join(CSRC, h)) for h in HEADERS]
    deps.append(os.path.abspath(__file__))
    return any(os.path.getmtime(d) > t for d in deps if os.path.isfile(d))


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile the extension if it is missing or older than its sources; return the .so path."""
    if not force and not _stale():
        return LIB
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.isfile(os.path.join(CSRC, s))]
    extra = os.environ.get("TK_NVCC_EXTRA", "").split()   # developer knob, e.g. -DTK_HUMANOID_MIN_CTAS=14
    cmd = [nvcc_path()] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + srcs
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + proc.stdout + proc.stderr)
    if verbose:
        print(proc.stdout + proc.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
