#!/usr/bin/env python
"""SURVEY 8(f) rank 4 measurement: optimiser step on the reference's six parameter groups at P Gaussians:
torch.optim.Adam (the reference's default, dense) vs SparseGaussianAdam (one launch, visible rows only).
Usage: time_sparse_adam.py [P] [visible_fraction]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "priordepth-gaussiansplatting_b200")):
    sys.path.insert(0, p)
import torch

from pdgs_b200.optim import SparseGaussianAdam

P = int(sys.argv[1]) if len(sys.argv) > 1 else 3_000_000
frac = float(sys.argv[2]) if len(sys.argv) > 2 else 0.79
dev = torch.device("cuda:0")
shapes = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "opacity": (1,), "scaling": (3,), "rotation": (4,)}


def build(cls, **kw):
    ps = {k: torch.nn.Parameter(torch.randn((P,) + s, device=dev)) for k, s in shapes.items()}
    for t in ps.values():
        t.grad = torch.randn_like(t) * 1e-3
    return ps, cls([{"params": [t], "lr": 1e-3, "name": k} for k, t in ps.items()], lr=0.0, eps=1e-15, **kw)


def timed(fn, n=30):
    for _ in range(5):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


radii = ((torch.rand(P, device=dev) < frac) * 5).int()
_, dense = build(torch.optim.Adam)
ms_dense = timed(dense.step)
del dense
torch.cuda.empty_cache()
_, sparse = build(SparseGaussianAdam)
ms_sparse = timed(lambda: sparse.step(radii > 0, P))
ms_sparse_radii = timed(lambda: sparse.step(radii, P))
V = int((radii > 0).sum())
bytes_alg = V * 59 * 7 * 4 + P  # 4 loads + 3 stores per updated float, 1 B of visibility per Gaussian
print(json.dumps(dict(P=P, visible=V, torch_adam_ms=ms_dense, sparse_adam_ms=ms_sparse, sparse_adam_radii_ms=ms_sparse_radii,
                      algorithmic_bytes=bytes_alg, achieved_GBps=bytes_alg / ms_sparse_radii / 1e6)))
