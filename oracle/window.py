"""TEST INFRASTRUCTURE (oracle): the float64 oracle restricted to a window of tiles of a LARGE view.

The float64 oracle (oracle/raster_oracle.py) cannot blend a whole 1.6K-px view of millions of Gaussians in reasonable
time, so full-size parity checks and bench.py's CPU baseline run it on the Gaussians whose tile rect (from the fp32
bit-model, oracle/bitmodel.c) touches a window of tiles.  Inside the window the result is exactly the whole-view
result: the per-tile lists depend only on the Gaussians that touch the tile.  Gradients of a loss whose cotangent is
zero outside the window's pixels are exact too.

Imported only by tests/ and bench.py's CPU legs.
"""
from __future__ import annotations

import numpy as np
import torch

from . import bitmodel
from .raster_oracle import Oracle, Settings


def tile_grid(W: int, H: int):
    return (W + 15) // 16, (H + 15) // 16


def place_window(W: int, H: int, window_tiles, position=(0.5, 0.5)):
    """(x0, y0, x1, y1) in tiles: a window of `window_tiles` = (tw, th) tiles whose centre sits at the fractional
    `position` of the tile grid (clipped to the image)."""
    gx, gy = tile_grid(W, H)
    tw, th = min(int(window_tiles[0]), gx), min(int(window_tiles[1]), gy)
    x0 = int(round(position[0] * (gx - tw)))
    y0 = int(round(position[1] * (gy - th)))
    return (x0, y0, x0 + tw, y0 + th)


def window_pixels(win, W: int, H: int):
    """(ys, xs) slices of the pixels covered by a tile window."""
    x0, y0, x1, y1 = win
    return slice(16 * y0, min(16 * y1, H)), slice(16 * x0, min(16 * x1, W))


def structure(scene, view, scale_modifier: float = 1.0):
    """fp32 bit-model of the whole view (radii, rects, depth bits, pixel centres, tiles touched)."""
    return bitmodel.preprocess(scene.means3D, scene.scales, scale_modifier, scene.rotations, None,
                               view.world_view_transform, view.full_proj_transform, view.image_width,
                               view.image_height, np.float32(view.tanfovx), np.float32(view.tanfovy))


def window_oracle(scene, view, win, st=None, need_grad: bool = True, antialiasing: bool = False):
    """Oracle over the Gaussians touching tile window `win` of `view`.
    -> (oracle, ids [n] int64 indices into the scene, info dict(pairs_window, pairs_view, st))."""
    W, H = view.image_width, view.image_height
    if st is None:
        st = structure(scene, view)
    rect, vis = st["rect"], st["radii"] > 0
    hit = vis & (rect[:, 0] < win[2]) & (rect[:, 2] > win[0]) & (rect[:, 1] < win[3]) & (rect[:, 3] > win[1])
    ids = torch.from_numpy(np.nonzero(hit)[0])
    cx0, cx1 = np.clip(rect[hit, 0], win[0], win[2]), np.clip(rect[hit, 2], win[0], win[2])
    cy0, cy1 = np.clip(rect[hit, 1], win[1], win[3]), np.clip(rect[hit, 3], win[1], win[3])
    pairs_window = int(((cx1 - cx0) * (cy1 - cy0)).sum())
    sub = {k: (v[hit] if hasattr(v, "shape") and v.shape[:1] == hit.shape else v) for k, v in st.items()}
    sub["num_rendered"] = int(sub["tiles"].sum())
    s = Settings(image_height=H, image_width=W, tanfovx=view.tanfovx, tanfovy=view.tanfovy, bg=scene.bg,
                 scale_modifier=1.0, viewmatrix=view.world_view_transform, projmatrix=view.full_proj_transform,
                 sh_degree=scene.sh_degree, campos=view.camera_center, antialiasing=antialiasing)
    o = Oracle(s, scene.means3D[ids], scene.opacities[ids], shs=scene.shs[ids], scales=scene.scales[ids],
               rotations=scene.rotations[ids], need_grad=need_grad, structure=sub, tile_window=win)
    return o, ids, dict(pairs_window=pairs_window, pairs_view=int(st["num_rendered"]), st=st)
