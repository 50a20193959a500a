"""Launch + drain floor of this box: an empty kernel with the launch shape of the fused kernels, back to back on one
stream (developer probe; bench.py prints the same numbers in extras.launch_floor).

    python tools/launch_floor.py
"""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from thormang_isaacgym_b200 import _native as nat  # noqa: E402


def launch_floor_graph(grid, block, smem, pdl, per_graph=100, replays=20, device="cuda:0"):
    """Same launches captured in a CUDA graph (no host launch cost): the GPU-side kernel-to-kernel floor."""
    lib = nat.lib()
    fn = lib.tk_debug_launch_floor
    fn.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    fn.restype = C.c_int
    buf = torch.zeros(max(grid, 1), dtype=torch.int32, device=device)
    side = torch.cuda.Stream(device=device)
    with torch.cuda.stream(side):
        for _ in range(3):
            fn(grid, block, smem, pdl, buf.data_ptr(), side.cuda_stream)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        st = torch.cuda.current_stream(torch.device(device)).cuda_stream
        for _ in range(per_graph):
            fn(grid, block, smem, pdl, buf.data_ptr(), st)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(replays):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / (replays * per_graph)


def launch_floor(grid, block, smem, pdl, steps=2000, sink=True, device="cuda:0"):
    lib = nat.lib()
    fn = lib.tk_debug_launch_floor
    fn.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    fn.restype = C.c_int
    buf = torch.zeros(max(grid, 1), dtype=torch.int32, device=device) if sink else None
    ptr = buf.data_ptr() if sink else None
    st = torch.cuda.current_stream(torch.device(device)).cuda_stream
    for _ in range(50):
        fn(grid, block, smem, pdl, ptr, st)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn(grid, block, smem, pdl, ptr, st)
    e1.record()
    torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / steps


if __name__ == "__main__":
    torch.zeros(1, device="cuda:0")
    for grid, block, smem in ((128, 32, 14 * 1024), (512, 32, 14 * 1024), (512, 128, 28 * 1024), (2048, 32, 14 * 1024),
                              (2048, 128, 28 * 1024), (4096, 32, 14 * 1024), (2048, 32, 4608)):
        for pdl in (0, 1):
            print(json.dumps({"grid": grid, "block": block, "smem": smem, "pdl": pdl,
                              "us_per_launch": round(launch_floor(grid, block, smem, pdl), 2),
                              "us_per_launch_cuda_graph": round(launch_floor_graph(grid, block, smem, pdl), 2)}), flush=True)
