#!/usr/bin/env python
"""SURVEY 8(f) rank 2 measurement: one training-shaped iteration (render() marshalling + rasterizer fwd + L1 + bwd down to
the optimiser's raw parameters) with
  unfused: torch sigmoid / exp / normalize + torch.cat of features_dc / features_rest, autograd back through them
           (what /root/reference/gaussian_renderer/__init__.py:55-112 and scene/gaussian_model.py:103-131 do), and
  fused:   dc= / shs= / raw_params=True -- activations and the split SH read inside the kernels.
CUDA events, median of N.  Usage: time_fused_inputs.py [C4|C2|C3|C1] [N]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "priordepth-gaussiansplatting_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch

import helpers
from diff_gaussian_rasterization import GaussianRasterizer
from pdgs_b200 import rasterizer as rz
from pdgs_b200 import scenes

cfg = sys.argv[1] if len(sys.argv) > 1 else "C4"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 30
dev = torch.device("cuda:0")
sc = scenes.make_config(cfg)
view = sc.views[0]
rs = helpers.cuda_settings(sc, view, dev)
raw = dict(xyz=sc.means3D, op=torch.logit(sc.opacities.clamp(1e-6, 1 - 1e-6)), sc=torch.log(sc.scales), rot=sc.rotations * 1.7,
           dc=sc.shs[:, :1, :].contiguous(), rest=sc.shs[:, 1:, :].contiguous())
lv = {k: v.to(dev).clone().requires_grad_(True) for k, v in raw.items()}
target = torch.rand(3, view.image_height, view.image_width, device=dev)
means2D = torch.zeros(sc.P, 3, device=dev, requires_grad=True)


def iteration(fused: bool):
    for t in list(lv.values()) + [means2D]:
        t.grad = None
    r = GaussianRasterizer(rs)
    if fused:
        color, radii, invd = r(means3D=lv["xyz"], means2D=means2D, opacities=lv["op"], scales=lv["sc"], rotations=lv["rot"],
                               dc=lv["dc"], shs=lv["rest"], raw_params=True)
    else:
        color, radii, invd = r(means3D=lv["xyz"], means2D=means2D, opacities=torch.sigmoid(lv["op"]),
                               scales=torch.exp(lv["sc"]), rotations=torch.nn.functional.normalize(lv["rot"]),
                               shs=torch.cat((lv["dc"], lv["rest"]), dim=1))
    dcolor, _, _ = rz.image_loss(color, target)
    color.backward(dcolor)


def timed(fused):
    for _ in range(5):
        iteration(fused)
    ts = []
    for _ in range(N):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        iteration(fused)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2], ts[len(ts) // 10], ts[(9 * len(ts)) // 10]


out = dict(config=cfg, P=sc.P, image=[view.image_width, view.image_height], n=N)
for name, f in (("unfused", False), ("fused", True), ("unfused_again", False)):
    rz.set_stage_timing(dev, True)
    rz.stage_times(dev)
    med, p10, p90 = timed(f)
    st = rz.stage_times(dev)
    rz.set_stage_timing(dev, False)
    out[name] = dict(ms_median=med, ms_p10=p10, ms_p90=p90,
                     stages_ms={k: v[0] / max(v[1], 1) for k, v in st.items() if v[1]})
print(json.dumps(out))
