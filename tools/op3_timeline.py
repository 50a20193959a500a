"""Per-warp phase timeline of the fused MA_OP3 kernel (developer probe; globaltimer stamps in ns).

    python tools/op3_timeline.py --envs 262144
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
ap.add_argument("--envs", type=int, default=262144)
ap.add_argument("--tile", type=int, default=16, help="envs per CTA the library was built with (TK_OP3_TILE)")
a = ap.parse_args()
n = a.envs
envs = [G._MAKERS["ma_op3"](n, 10 + s, "cuda:0")[0] for s in range(3)]
lib = nat.lib()
probe = lib.tk_debug_ma_op3_timeline
probe.argtypes = [C.c_void_p, C.c_void_p]
tiles = (n + a.tile - 1) // a.tile
for _ in range(4):
    for e in envs:
        e.post_physics_step()
torch.cuda.synchronize()
buf = torch.zeros((tiles, 4, 8), dtype=torch.int64, device="cuda:0")
e = envs[1]
probe(e._handle, C.c_void_p(buf.data_ptr()))
envs[2].post_physics_step()
envs[0].post_physics_step()
e.post_physics_step()
envs[2].post_physics_step()
torch.cuda.synchronize()
t = buf.cpu().double()[:, :(4 if a.tile == 32 else 2)]
t0 = t[:, :, 0].min()
print(f"tiles {tiles}; kernel span {(t[:, :, 6].max() - t0) / 1e3:.2f} us")
names = ["start", "scalars issued", "barrier passed", "tiles landed", "row math done", "finals done", "exit"]
roles = ["obs 0", "obs 1", "rew 0", "rew 1"] if a.tile == 32 else ["obs", "rew"]
for w, role in enumerate(roles):
    d = t[:, w, 1:7] - t[:, w, 0:6]
    life = t[:, w, 6] - t[:, w, 0]
    print(f"warp {w} ({role}): lifetime median {life.median() / 1e3:.2f} mean {life.mean() / 1e3:.2f} p95 {life.quantile(0.95) / 1e3:.2f} us")
    for i in range(6):
        print(f"    {names[i]:>16s} -> {names[i + 1]:<16s} median {d[:, i].median() / 1e3:6.2f}  mean {d[:, i].mean() / 1e3:6.2f}  p95 {d[:, i].quantile(0.95) / 1e3:6.2f} us")
cta_life = t[:, :, 6].max(1).values - t[:, :, 0].min(1).values
print(f"CTA lifetime median {cta_life.median() / 1e3:.2f} mean {cta_life.mean() / 1e3:.2f} us; "
      f"tiles per SM {tiles / 148:.1f}; resident CTAs implied {cta_life.mean() * tiles / 148 / (t[:, :, 6].max() - t0):.2f}")
