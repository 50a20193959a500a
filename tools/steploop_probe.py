import sys, os, json
sys.path.insert(0, os.getcwd())
import torch, bench_gpu as G
for n, h, r, gr in ((131072, 32, 8, False), (4096, 32, 20, False), (4096, 32, 20, True), (16384, 32, 20, False)):
    out = G._step_loop_bench(torch, n, "cuda:0", h, r, graph=gr)
    print(n, gr, round(out["us_per_step"], 2))
