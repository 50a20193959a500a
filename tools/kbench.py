"""Kernel-only sweep of the fused post-physics kernels (developer tool; bench.py is the contract benchmark).

    python tools/kbench.py [--task humanoid] [--envs 131072,65536] [--tiles 32,64,128] [--steps 300]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bench as B  # noqa: E402
import bench_gpu as G  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--task", default="humanoid")
    ap.add_argument("--envs", default="131072")
    ap.add_argument("--tiles", default="32")
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--pdl", default="1")
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--kernel", default="post", choices=["post", "pre"])
    ap.add_argument("--graph", action="store_true", help="capture one pass over the state sets in a CUDA graph and replay it")
    ap.add_argument("--host", action="store_true", help="host cost of one post_physics_step() call (no GPU wait in between)")
    ap.add_argument("--sweep", default="", help="ENV_VAR=v1,v2,... : one line per value (developer knobs of the kernels)")
    args = ap.parse_args()
    peak, _ = B.measured_peak_gbs()
    bpe = B.BYTES_PER_ENV[args.task]
    # clock ramp: the first measurement of a fresh process reads ~20 % slow (seen at 16384 envs: 9.9 vs 8.3 us)
    w = torch.empty(64 << 20, device="cuda:0")
    t_end = __import__("time").perf_counter() + 0.5
    while __import__("time").perf_counter() < t_end:
        w.mul_(1.0001)
        torch.cuda.synchronize()
    del w
    if args.kernel == "pre":
        pre_bytes = 252 if args.task == "humanoid" else 48
        for n in [int(x) for x in args.envs.split(",")]:
            sets = max(3, int(2.2 * B.L2_BYTES / (n * pre_bytes)) + 1)
            sets = min(sets, 64)
            make = G._make_humanoid if args.task == "humanoid" else G._make_gogoro
            envs = [make(n, 5 + s, "cuda:0")[0] for s in range(sets)]
            width = 21 if args.task == "humanoid" else 1
            acts = [torch.rand((n, width), device="cuda:0") * 2.4 - 1.2 for _ in range(sets)]

            def step(i):
                e = envs[i % sets]
                e._clip_in_kernel = True
                e.pre_physics_step(acts[i % sets])
                return 1
            ms, _, _ = G._time_steps(torch, step, args.steps, args.warmup)
            us = 1e3 * ms / args.steps
            gbs = n * pre_bytes / (us * 1e-6) / 1e9
            print(json.dumps({"task": args.task, "kernel": "pre_physics", "envs": n, "us_per_step": round(us, 2),
                              "GBps": round(gbs, 1), "frac": round(gbs / peak, 4), "sets": sets}), flush=True)
        return
    if args.host:
        import time
        env = G._MAKERS[args.task](32, 5, "cuda:0")[0]
        for _ in range(200):
            env.post_physics_step()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5000):
            env.post_physics_step()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        print(json.dumps({"task": args.task, "host_us_per_call": round((t1 - t0) / 5000 * 1e6, 2),
                          "note": "32 envs: the GPU side is the empty-launch floor, the loop is bound by the host call"}), flush=True)
        return
    if args.sweep:
        key, vals = args.sweep.split("=")
        for n in [int(x) for x in args.envs.split(",")]:
            for v in vals.split(","):
                os.environ[key] = v
                r = G._kernel_bench(torch, args.task, n, "cuda:0", args.steps, args.warmup, seed=5, graph=args.graph)
                us = 1e3 * r["ms_total"] / r["steps"]
                gbs = n * bpe / (us * 1e-6) / 1e9
                print(json.dumps({"task": args.task, "envs": n, "graph": args.graph, key: v, "us_per_step": round(us, 2),
                                  "GBps": round(gbs, 1), "frac": round(gbs / peak, 4), "sets": r["sets"]}), flush=True)
        return
    for n in [int(x) for x in args.envs.split(",")]:
        for tile in [int(x) for x in args.pdl.split(",")]:
            os.environ["TK_HUMANOID_PDL"] = str(tile)
            os.environ["TK_HUMANOID_TILE"] = str(tile)
            os.environ["TK_GOGORO_TILE"] = str(tile)
            r = G._kernel_bench(torch, args.task, n, "cuda:0", args.steps, args.warmup, seed=5)
            us = 1e3 * r["ms_total"] / args.steps
            gbs = n * bpe / (us * 1e-6) / 1e9
            print(json.dumps({"task": args.task, "envs": n, "pdl": tile, "us_per_step": round(us, 2),
                              "GBps": round(gbs, 1), "frac": round(gbs / peak, 4), "sets": r["sets"]}), flush=True)


if __name__ == "__main__":
    main()
