#!/usr/bin/env python
"""One full training iteration of the reference's loop (/root/reference/train.py:129-338, GPU part) at a config, two ways:

  reference-shaped: the torch ops the reference runs around the rasterizer -- activations + torch.cat in render()
      (gaussian_renderer/__init__.py:55-112, scene/gaussian_model.py:103-131), loss = (1-l)*L1 + l*(1-fused_ssim) + w*depth L1
      (train.py:153-174), loss.backward() through autograd, densification statistics (train.py:313-314,
      scene/gaussian_model.py:502-506), torch.optim.Adam.step() -- with THIS rasterizer and THIS fused_ssim underneath;
  fused: raw parameters + split SH straight into the kernels, pdgs_image_loss_ssim (loss + cotangents, 3 launches),
      backward with the densification statistics inside K9, visibility-sparse Adam (one launch).

CUDA events, median of N.  Usage: train_iter.py [C4|C3|C2|C1] [N]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "priordepth-gaussiansplatting_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch

import helpers
from diff_gaussian_rasterization import GaussianRasterizer, SparseGaussianAdam
from fused_ssim import fused_ssim
from pdgs_b200 import rasterizer as rz
from pdgs_b200 import scenes

cfg = sys.argv[1] if len(sys.argv) > 1 else "C4"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 30
dev = torch.device("cuda:0")
sc = scenes.make_config(cfg)
view = sc.views[0]
rs = helpers.cuda_settings(sc, view, dev)
H, W, P = view.image_height, view.image_width, sc.P
LAMBDA, W_DEPTH = 0.2, 0.5
target = torch.rand(3, H, W, device=dev)
prior = torch.rand(1, H, W, device=dev)
mask = (torch.rand(1, H, W, device=dev) > 0.3).float()
LRS = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=1.25e-4, opacity=0.025, scaling=5e-3, rotation=1e-3)


def make_params():
    raw = dict(xyz=sc.means3D, f_dc=sc.shs[:, :1, :].contiguous(), f_rest=sc.shs[:, 1:, :].contiguous(),
               opacity=torch.logit(sc.opacities.clamp(1e-6, 1 - 1e-6)), scaling=torch.log(sc.scales), rotation=sc.rotations * 1.7)
    return {k: torch.nn.Parameter(v.to(dev).clone()) for k, v in raw.items()}


def groups(p):
    return [{"params": [p[k]], "lr": LRS[k], "name": k} for k in p]


# ---------------- reference-shaped iteration ----------------
pa = make_params()
opt_a = torch.optim.Adam(groups(pa), lr=0.0, eps=1e-15)
accum_a, denom_a, maxr_a = torch.zeros(P, 1, device=dev), torch.zeros(P, 1, device=dev), torch.zeros(P, device=dev)


def iter_reference():
    means2D = torch.zeros(P, 3, device=dev, requires_grad=True)
    color, radii, invd = GaussianRasterizer(rs)(
        means3D=pa["xyz"], means2D=means2D, opacities=torch.sigmoid(pa["opacity"]), scales=torch.exp(pa["scaling"]),
        rotations=torch.nn.functional.normalize(pa["rotation"]), shs=torch.cat((pa["f_dc"], pa["f_rest"]), dim=1))
    image = color.clamp(0, 1)
    Ll1 = (image - target).abs().mean()
    loss = (1.0 - LAMBDA) * Ll1 + LAMBDA * (1.0 - fused_ssim(image.unsqueeze(0), target.unsqueeze(0)))
    loss = loss + W_DEPTH * ((invd - prior) * mask).abs().mean()
    loss.backward()
    with torch.no_grad():
        vis = radii > 0
        maxr_a[vis] = torch.max(maxr_a[vis], radii[vis].float())
        accum_a[vis] += torch.norm(means2D.grad[vis, :2], dim=-1, keepdim=True)
        denom_a[vis] += 1
    opt_a.step()
    opt_a.zero_grad(set_to_none=True)


# ---------------- fused iteration ----------------
pb = make_params()
opt_b = SparseGaussianAdam(groups(pb), lr=0.0, eps=1e-15)
f32 = dict(dtype=torch.float32, device=dev)
grads = dict(means3D=torch.empty(P, 3, **f32), means2D=torch.empty(P, 3, **f32), shs_dc=torch.empty(P, 1, 3, **f32),
             shs_rest=torch.empty(P, 15, 3, **f32), opacities=torch.empty(P, 1, **f32), scales=torch.empty(P, 3, **f32),
             rotations=torch.empty(P, 4, **f32), densify_grad_accum=torch.zeros(P, **f32), densify_denom=torch.zeros(P, **f32),
             max_radii2D=torch.zeros(P, dtype=torch.int32, device=dev), overwrite=True)  # K9 writes every gradient row
loss_acc = torch.zeros(3, **f32)
ssim_scratch = torch.empty(rz.ssim_scratch_bytes(W, H), dtype=torch.uint8, device=dev)
empty = torch.empty(0, device=dev)
NAMES = dict(xyz="means3D", f_dc="shs_dc", f_rest="shs_rest", opacity="opacities", scaling="scales", rotation="rotations")


def iter_fused():
    with torch.no_grad():
        R, color, radii, geom, binning, img, invd = rz.rasterize_gaussians(
            rs.bg, pb["xyz"], empty, pb["opacity"], pb["scaling"], pb["rotation"], 1.0, empty, rs.viewmatrix, rs.projmatrix,
            rs.tanfovx, rs.tanfovy, H, W, pb["f_rest"], rs.sh_degree, rs.campos, False, False, False, dc=pb["f_dc"], raw_params=True)
        # render() clamps the image before the loss (gaussian_renderer/__init__.py:121): same here, with the clamp's mask
        image = color.clamp(0, 1)
        dcolor, ddepth, _ = rz.image_loss_ssim(image, target, invd, prior, mask, LAMBDA, W_DEPTH, loss_acc, ssim_scratch)
        dcolor = dcolor * ((color >= 0) & (color <= 1))
        rz.rasterize_gaussians_backward(
            rs.bg, pb["xyz"], radii, empty, pb["opacity"], pb["scaling"], pb["rotation"], 1.0, empty, rs.viewmatrix, rs.projmatrix,
            rs.tanfovx, rs.tanfovy, dcolor, ddepth, pb["f_rest"], rs.sh_degree, rs.campos, geom, R, binning, img, False, False,
            accumulate_into=grads, dc=pb["f_dc"], raw_params=True)
        for k, g in NAMES.items():
            pb[k].grad = grads[g].view_as(pb[k])
        opt_b.step(radii, P)


def timed(fn):
    for _ in range(5):
        fn()
    ts = []
    for _ in range(N):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return dict(ms_median=ts[len(ts) // 2], ms_p10=ts[len(ts) // 10], ms_p90=ts[(9 * len(ts)) // 10])


out = dict(config=cfg, P=P, image=[W, H], n=N, lambda_dssim=LAMBDA, w_depth=W_DEPTH)
out["reference_shaped"] = timed(iter_reference)
out["fused"] = timed(iter_fused)
out["reference_shaped_again"] = timed(iter_reference)
print(json.dumps(out))
