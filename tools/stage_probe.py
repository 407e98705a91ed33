#!/usr/bin/env python
"""Per-stage CUDA-event times of single-view forward (+ backward) at a config, for A/B of library variants (PDGS_LIB).
Usage: stage_probe.py [C5|C4|...] [iters] [--bwd]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "priordepth-gaussiansplatting_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch

import helpers
from pdgs_b200 import rasterizer as rz
from pdgs_b200 import scenes

cfg = sys.argv[1] if len(sys.argv) > 1 else "C5"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 10
bwd = "--bwd" in sys.argv
dev = torch.device("cuda:0")
sc = scenes.make_config(cfg, n_views=1)
view = sc.views[0]
rs = helpers.cuda_settings(sc, view, dev)
g = {k: getattr(sc, k).to(dev) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
empty = torch.empty(0, device=dev)
H, W = view.image_height, view.image_width


def once():
    out = rz.rasterize_gaussians(rs.bg, g["means3D"], empty, g["opacities"], g["scales"], g["rotations"], 1.0, empty, rs.viewmatrix,
                                 rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, g["shs"], rs.sh_degree, rs.campos, False, False, False)
    if bwd:
        R, color, radii, geom, binning, img, invd = out
        rz.rasterize_gaussians_backward(rs.bg, g["means3D"], radii, empty, g["opacities"], g["scales"], g["rotations"], 1.0, empty,
                                        rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, torch.ones_like(color), None, g["shs"],
                                        rs.sh_degree, rs.campos, geom, R, binning, img, False, False)
    return out[0]


for _ in range(3):
    R = once()
torch.cuda.synchronize()
rz.set_stage_timing(dev, True)
rz.stage_times(dev)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(N):
    once()
e1.record()
torch.cuda.synchronize()
st = rz.stage_times(dev)
print(json.dumps(dict(config=cfg, lib=os.environ.get("PDGS_LIB", "libpdgs.so"), R=R, ms_per_iter=e0.elapsed_time(e1) / N,
                      stages_ms={k: round(v[0] / max(v[1], 1), 4) for k, v in st.items() if v[1]})))
