#!/usr/bin/env python
"""Where does the end-to-end overhead go?  Times the batch step with (a) everything resident, (b) host target images,
(c) + Gaussians uploaded on the compute stream, (d) + Gaussians staged one step ahead on the side stream."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "priordepth-gaussiansplatting_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch
import helpers
from pdgs_b200 import scenes
from pdgs_b200.viewparallel import ViewParallelRasterizer
V = int(sys.argv[1]) if len(sys.argv) > 1 else 16
dev = torch.device("cuda:0")
sc = scenes.make_config("C4", n_views=V)
H, W = sc.views[0].image_height, sc.views[0].image_width
host = {k: getattr(sc, k).contiguous().pin_memory() for k in ("means3D", "shs", "opacities", "scales", "rotations")}
g_dev = {k: v.to(dev) for k, v in host.items()}
tg_host = {i: torch.rand(3, H, W).pin_memory() for i in range(V)}
tg_dev = {i: t.to(dev) for i, t in tg_host.items()}
views = [helpers.cuda_settings(sc, v, dev) for v in sc.views]
vp = ViewParallelRasterizer(sc.P, 16, dev)
def run(name, fn, n=4):
    fn(); fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for k in range(n): fn(k, n)
    torch.cuda.synchronize(); print(f"{name:40s} {(time.perf_counter() - t0) / n * 1e3:8.2f} ms/step")
run("resident", lambda *a: vp.step(g_dev, views, tg_dev)["loss_sums"].cpu())
run("host targets", lambda *a: vp.step(g_dev, views, tg_host)["loss_sums"].cpu())
run("host targets + sync param upload", lambda *a: vp.step({k: v.to(dev, non_blocking=True) for k, v in host.items()}, views, tg_host)["loss_sums"].cpu())
staged = [None]
def piped(k=0, n=1):
    g = staged[0] if staged[0] is not None else vp.stage_gaussians(host)
    staged[0] = vp.stage_gaussians(host) if k + 1 < n else None
    return vp.step(g, views, tg_host)["loss_sums"].cpu()
run("host targets + staged param upload", piped)
def sliced(k=0, n=1):
    g = staged[0] if staged[0] is not None else vp.stage_gaussians(host)
    out = vp.step(g, views, tg_host, prefetch_gaussians=host if k + 1 < n else None)
    staged[0] = out["prefetched"]
    return out["loss_sums"].cpu()
staged[0] = None
run("host targets + sliced param prefetch", sliced)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s = torch.cuda.Stream(dev)
with torch.cuda.stream(s):
    e0.record(s); g = {k: v.to(dev, non_blocking=True) for k, v in host.items()}; e1.record(s)
torch.cuda.synchronize(); print("param upload alone", e0.elapsed_time(e1), "ms")
