"""Per-tile phase timeline of the Humanoid stream kernel (developer probe; globaltimer stamps in ns).

    python tools/tile_timeline.py --envs 16384
"""
import argparse
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench_gpu as G  # noqa: E402
from thormang_isaacgym_b200 import _native as nat  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--envs", type=int, default=16384)
ap.add_argument("--task", default="humanoid", choices=["humanoid", "gogoro"])
ap.add_argument("--sets", type=int, default=8, help="state sets rotated (8 x 15 MB stays L2-resident; bench.py rotates > 2x L2)")
a = ap.parse_args()
n = a.envs
envs = [G._MAKERS[a.task](n, 10 + s, "cuda:0")[0] for s in range(max(3, a.sets))]
lib = nat.lib()
probe = lib.tk_debug_humanoid_timeline if a.task == "humanoid" else lib.tk_debug_gogoro_timeline
probe.argtypes = [C.c_void_p, C.c_void_p]
tile_envs = 32 if a.task == "humanoid" else 128
tiles = (n + tile_envs - 1) // tile_envs
for _ in range(6):          # past the transient: the synthetic state's fallen envs are reset by the first steps
    for e in envs:
        e.post_physics_step()
torch.cuda.synchronize()
buf = torch.zeros((tiles, 8), dtype=torch.int64, device="cuda:0")
ia, ib = (3, 4) if len(envs) >= 6 else (0, 1)
e = envs[ia]
probe(e._handle, C.c_void_p(buf.data_ptr()))
others = [o for k, o in enumerate(envs) if k not in (ia, ib)]
# flush L2 with the other sets (when there are enough of them), then the probed launch
for o in others:
    o.post_physics_step()
torch.cuda.synchronize()
e_reset_before = e.reset_buf.clone().cpu()
# the probed launch sits in the middle of a back-to-back stream of launches (programmatic dependent launch in play);
# the launch behind it is probed as well: the distance between "every tile of A has finished" and "the first tile of B has
# passed griddepcontrol.wait" is the launch-to-launch gap that no tile stamp of a single launch can see
buf_b = torch.zeros((tiles, 8), dtype=torch.int64, device="cuda:0")
eb = envs[ib]
probe(eb._handle, C.c_void_p(buf_b.data_ptr()))
for o in others:       # the probed sets were touched longest ago: HBM-cold when --sets is large
    o.post_physics_step()
e.post_physics_step()
eb.post_physics_step()
others[0].post_physics_step()
torch.cuda.synchronize()
t = buf.cpu().double()
tb = buf_b.cpu().double()
print(f"launch A: first wait release -> last tile done {(t[:, 6].max() - t[:, 0].min()) / 1e3:.2f} us; "
      f"gap to launch B's first wait release {(tb[:, 0].min() - t[:, 6].max()) / 1e3:.2f} us; "
      f"B's first release -> last done {(tb[:, 6].max() - tb[:, 0].min()) / 1e3:.2f} us; "
      f"A start -> B start {(tb[:, 0].min() - t[:, 0].min()) / 1e3:.2f} us")
t0 = t[:, 0].min()
names = ["start", "scalars issued", "scalars landed", "bulk landed", "inputs in regs", "math+STS done", "store read done"]
phases = ["issue", "scalar wait", "bulk wait", "reset+reg read", "math+STS", "store"]
if a.task == "gogoro":      # tk_gogoro.cu stamps: 0 wait released, 1 scalars landed + ballots, 2 reset prologue, 3 tile landed + patched, 4 math done, 5 CTA barrier, 6 stores read
    names = ["start", "scalars landed", "reset prologue", "tile landed", "math done", "cta barrier", "store read done"]
    phases = ["scalar round trip", "reset prologue", "tile wait", "math", "barrier", "store"]
print(f"envs flagged for reset in the probed step: {int((e_reset_before != 0).sum())} of {n}")
print(f"tiles {tiles}; kernel span {(t[:, 6].max() - t0) / 1e3:.2f} us")
for i, nm in enumerate(names):
    col = t[:, i] - t0
    print(f"{nm:18s} min {col.min()/1e3:7.2f}  median {col.median()/1e3:7.2f}  max {col.max()/1e3:7.2f} us")
d = t[:, 1:7] - t[:, 0:6]
for i, nm in enumerate(phases):
    print(f"  phase {nm:14s} median {d[:, i].median()/1e3:6.2f}  mean {d[:, i].mean()/1e3:6.2f}  p95 {d[:, i].quantile(0.95)/1e3:6.2f} us")

# the stragglers: the ten tiles that finish last, with their stamps (us after the first start) and reset counts
rb = e_reset_before.view(-1, tile_envs).sum(1) if e_reset_before.numel() % tile_envs == 0 else None
order = torch.argsort(t[:, 6], descending=True)[:10]
print("last tiles: blockIdx  resets | start issue scal bulk regs math done")
for i in order.tolist():
    row = " ".join(f"{(t[i, k] - t0) / 1e3:6.2f}" for k in range(8))
    print(f"  {i:6d}  {int(rb[i]) if rb is not None else -1:3d} | {row}")
first = torch.argsort(t[:, 6])[:5]
for i in first.tolist():
    row = " ".join(f"{(t[i, k] - t0) / 1e3:6.2f}" for k in range(7))
    print(f"  first {i:6d}  {int(rb[i]) if rb is not None else -1:3d} | {row}")

if a.task != "humanoid":
    sys.exit(0)
# bandwidth over time: a tile's 13.9 KB of loads are counted when its bulk load has landed, its 14.1 KB obs tile when the
# bulk store has been read out of shared memory (the per-env scalars ride along): where is HBM idle -- head, middle, tail?
w = 1024.0                                   # ns per bucket
landed = ((t[:, 3] - t0) / w).long()
stored = ((t[:, 6] - t0) / w).long()
nb = int(stored.max()) + 1
ld = torch.bincount(landed, minlength=nb).double() * (109 * 32 * 4 + 32 * 28)
stv = torch.bincount(stored, minlength=nb).double() * (110 * 32 * 4 + 32 * 29)
print("bucket(us)  loaded MB  stored MB  GB/s")
for b in range(nb):
    print(f"{b * w / 1e3:6.2f}    {ld[b] / 1e6:8.2f}  {stv[b] / 1e6:8.2f}  {(ld[b] + stv[b]) / w:8.0f}")

