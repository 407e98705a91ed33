"""Shared helpers: run the oracle / the CUDA path on a synthetic scene and compare."""
from __future__ import annotations

import numpy as np
import torch

from oracle import bitmodel
from oracle.raster_oracle import Oracle, Settings
from pdgs_b200 import scenes


def oracle_settings(scene: scenes.Scene, view: scenes.View, antialiasing=False, scale_modifier=1.0, sh_degree=None):
    return Settings(image_height=view.image_height, image_width=view.image_width, tanfovx=view.tanfovx,
                    tanfovy=view.tanfovy, bg=scene.bg, scale_modifier=scale_modifier,
                    viewmatrix=view.world_view_transform, projmatrix=view.full_proj_transform,
                    sh_degree=scene.sh_degree if sh_degree is None else sh_degree, campos=view.camera_center,
                    antialiasing=antialiasing)


def bit_structure(scene, view, scale_modifier=1.0, cov3d=None):
    pre = bitmodel.preprocess(scene.means3D, None if cov3d is not None else scene.scales, scale_modifier,
                              None if cov3d is not None else scene.rotations, cov3d, view.world_view_transform,
                              view.full_proj_transform, view.image_width, view.image_height,
                              np.float32(view.tanfovx), np.float32(view.tanfovy))
    return pre


def run_oracle(scene, view, need_grad=True, antialiasing=False, use_structure=True, scale_modifier=1.0,
               colors_precomp=None, cov3d=None, sh_degree=None, tile_window=None, round_pixel_centre=True):
    st = bit_structure(scene, view, scale_modifier, cov3d) if use_structure else None
    o = Oracle(oracle_settings(scene, view, antialiasing, scale_modifier, sh_degree), scene.means3D, scene.opacities,
               shs=None if colors_precomp is not None else scene.shs, colors_precomp=colors_precomp,
               scales=None if cov3d is not None else scene.scales,
               rotations=None if cov3d is not None else scene.rotations, cov3D_precomp=cov3d, need_grad=need_grad,
               structure=st, tile_window=tile_window, round_pixel_centre=round_pixel_centre)
    res = o.forward()
    return o, res, st


def cuda_settings(scene, view, device, antialiasing=False, scale_modifier=1.0, sh_degree=None, debug=False):
    from diff_gaussian_rasterization import GaussianRasterizationSettings
    return GaussianRasterizationSettings(
        image_height=view.image_height, image_width=view.image_width, tanfovx=view.tanfovx, tanfovy=view.tanfovy,
        bg=scene.bg.to(device), scale_modifier=scale_modifier, viewmatrix=view.world_view_transform.to(device),
        projmatrix=view.full_proj_transform.to(device),
        sh_degree=scene.sh_degree if sh_degree is None else sh_degree, campos=view.camera_center.to(device),
        prefiltered=False, debug=debug, antialiasing=antialiasing)


def cuda_leaves(scene, device, colors_precomp=None, cov3d=None):
    def leaf(t):
        return None if t is None else t.detach().to(device).clone().requires_grad_(True)
    d = dict(means3D=leaf(scene.means3D), opacities=leaf(scene.opacities))
    d["means2D"] = torch.zeros(scene.P, 3, device=device, requires_grad=True)
    d["shs"] = None if colors_precomp is not None else leaf(scene.shs)
    d["colors_precomp"] = leaf(colors_precomp)
    d["scales"] = None if cov3d is not None else leaf(scene.scales)
    d["rotations"] = None if cov3d is not None else leaf(scene.rotations)
    d["cov3D_precomp"] = leaf(cov3d)
    return d


def run_cuda(scene, view, device="cuda:0", antialiasing=False, scale_modifier=1.0, colors_precomp=None, cov3d=None,
             sh_degree=None, debug=False):
    from diff_gaussian_rasterization import GaussianRasterizer
    rs = cuda_settings(scene, view, device, antialiasing, scale_modifier, sh_degree, debug)
    lv = cuda_leaves(scene, device, colors_precomp, cov3d)
    color, radii, invd = GaussianRasterizer(rs)(**lv)
    return lv, color, radii, invd


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    """Norm-wise relative error of a w.r.t. reference b."""
    a, b = a.double().cpu().reshape(-1), b.double().cpu().reshape(-1)
    return float((a - b).norm() / max(float(b.norm()), 1e-30))


def elem_err(a, b, floor_frac=1e-3):
    """max |a-b| / (|b| + floor) with floor = floor_frac * max|b|."""
    a, b = a.double().cpu().reshape(-1), b.double().cpu().reshape(-1)
    floor = floor_frac * float(b.abs().max()) + 1e-30
    return float(((a - b).abs() / (b.abs() + floor)).max())
