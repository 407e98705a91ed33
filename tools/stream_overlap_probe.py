#!/usr/bin/env python
"""Does running two views' pipelines concurrently (two host threads, two CUDA streams, separate gradient buffers) beat
running them back to back?  The blend kernels are issue/latency-bound, the preprocess / sort kernels HBM-bound, so
co-scheduling them could raise utilisation.  Prints ms per view for 1 and 2 (and 3) concurrent pipelines."""
import os, sys, time, threading
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "priordepth-gaussiansplatting_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch
import helpers
from pdgs_b200 import scenes
from pdgs_b200.viewparallel import ViewParallelRasterizer
V = int(sys.argv[1]) if len(sys.argv) > 1 else 24
dev = torch.device("cuda:0")
sc = scenes.make_config("C4", n_views=V)
H, W = sc.views[0].image_height, sc.views[0].image_width
g_dev = {k: getattr(sc, k).contiguous().to(dev) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
tg = {i: torch.rand(3, H, W, device=dev) for i in range(V)}
views = [helpers.cuda_settings(sc, v, dev) for v in sc.views]

def run(nthreads, reps=3):
    vps = [ViewParallelRasterizer(sc.P, 16, dev) for _ in range(nthreads)]
    streams = [torch.cuda.Stream(dev) for _ in range(nthreads)]
    parts = [list(range(t, V, nthreads)) for t in range(nthreads)]
    def work(t):
        with torch.cuda.stream(streams[t]):
            sub_views = [views[i] for i in parts[t]]
            sub_tg = {k: tg[i] for k, i in enumerate(parts[t])}
            vps[t].step(g_dev, sub_views, sub_tg)
    def once():
        ths = [threading.Thread(target=work, args=(t,)) for t in range(nthreads)]
        for th in ths: th.start()
        for th in ths: th.join()
        torch.cuda.synchronize()
    once()
    t0 = time.perf_counter()
    for _ in range(reps): once()
    dt = (time.perf_counter() - t0) / reps
    print(f"{nthreads} concurrent pipelines: {dt * 1e3 / V:.3f} ms/view ({V / dt:.1f} views/s)")
for n in (1, 2, 3, 1):
    run(n)
