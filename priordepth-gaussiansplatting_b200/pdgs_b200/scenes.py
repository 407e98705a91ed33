"""Deterministic synthetic scenes for BASELINE.json's five configs.

All tensors are generated on the CPU in fp32 from torch.Generator().manual_seed(20261104 + k)
and copied to the GPU by the caller.  Camera matrices restate the reference's conventions:
  getWorld2View2 / getProjectionMatrix   /root/reference/utils/graphics_utils.py:38-71
  transposition + full_proj + centre     /root/reference/scene/cameras.py:80-89
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional

import numpy as np
import torch

SEED0 = 20261104
ZNEAR, ZFAR = 0.01, 100.0


@dataclass
class View:
    image_width: int
    image_height: int
    FoVx: float
    FoVy: float
    world_view_transform: torch.Tensor   # [4,4] transposed world->view
    full_proj_transform: torch.Tensor    # [4,4] transposed
    camera_center: torch.Tensor          # [3]

    @property
    def tanfovx(self):
        return math.tan(self.FoVx * 0.5)

    @property
    def tanfovy(self):
        return math.tan(self.FoVy * 0.5)


@dataclass
class Scene:
    name: str
    means3D: torch.Tensor      # [P,3]
    scales: torch.Tensor       # [P,3] post-activation
    rotations: torch.Tensor    # [P,4] normalised
    opacities: torch.Tensor    # [P,1] post-sigmoid
    shs: torch.Tensor          # [P,16,3]
    bg: torch.Tensor           # [3]
    views: List[View]
    sh_degree: int = 3

    @property
    def P(self):
        return self.means3D.shape[0]


def projection_matrix(znear, zfar, fovX, fovY) -> torch.Tensor:
    """utils/graphics_utils.py:51-71 (z_sign = +1)."""
    tY, tX = math.tan(fovY / 2), math.tan(fovX / 2)
    top, right = tY * znear, tX * znear
    bottom, left = -top, -right
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def make_view(w2c: np.ndarray, W: int, H: int, tanx: float = 0.625) -> View:
    """w2c: 4x4 world->camera (the matrix getWorld2View2 returns)."""
    focal = W / (2.0 * tanx)
    FoVx = 2.0 * math.atan(W / (2.0 * focal))
    FoVy = 2.0 * math.atan(H / (2.0 * focal))
    wvt = torch.tensor(np.float32(w2c)).transpose(0, 1).contiguous()
    proj = projection_matrix(ZNEAR, ZFAR, FoVx, FoVy).transpose(0, 1)
    full = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    center = wvt.inverse()[3, :3].contiguous()
    return View(W, H, FoVx, FoVy, wvt, full, center)


def _common(gen, P):
    rot = torch.randn(P, 4, generator=gen)
    rot = rot / rot.norm(dim=1, keepdim=True)
    opac = torch.sigmoid(2.0 * torch.randn(P, 1, generator=gen))
    shs = torch.cat([torch.randn(P, 1, 3, generator=gen), 0.15 * torch.randn(P, 15, 3, generator=gen)], dim=1)
    return rot.contiguous(), opac.contiguous(), shs.contiguous()


def camera_relative_scene(name: str, k: int, P: int, W: int, H: int, mu_px: float, behind_frac: float = 0.05,
                          bg: Optional[torch.Tensor] = None) -> Scene:
    """C1/C2/C3/C5: camera at the origin looking down +z."""
    gen = torch.Generator().manual_seed(SEED0 + k)
    tanx = 0.625
    focal = W / (2.0 * tanx)
    tany = H / (2.0 * focal)
    z = torch.exp(torch.rand(P, generator=gen) * math.log(20.0))
    nb = int(P * behind_frac)
    if nb > 0:
        z[:nb] = -5.0 + 5.2 * torch.rand(nb, generator=gen)
    az = z.abs().clamp(min=0.2)
    x = (torch.rand(P, generator=gen) * 2 - 1) * 1.15 * tanx * az
    y = (torch.rand(P, generator=gen) * 2 - 1) * 1.15 * tany * az
    means = torch.stack([x, y, z], dim=1)
    perm = torch.randperm(P, generator=gen)
    means = means[perm].contiguous()
    az = az[perm]
    scales = (az / focal)[:, None] * torch.exp(math.log(mu_px) + 0.7 * torch.randn(P, 3, generator=gen))
    rot, opac, shs = _common(gen, P)
    view = make_view(np.eye(4), W, H, tanx)
    return Scene(name, means, scales.contiguous(), rot, opac, shs, torch.zeros(3) if bg is None else bg, [view])


def orbit_scene(name: str, k: int, P: int, W: int, H: int, mu_px: float, n_views: int = 64) -> Scene:
    """C4: means uniform in [-3,3]^3, cameras on a circle of radius 6 at height 1.5 looking at the origin, +y down."""
    gen = torch.Generator().manual_seed(SEED0 + k)
    tanx = 0.625
    focal = W / (2.0 * tanx)
    means = (torch.rand(P, 3, generator=gen) * 6.0 - 3.0).contiguous()
    scales = (6.0 / focal) * torch.exp(math.log(mu_px) + 0.7 * torch.randn(P, 3, generator=gen))
    rot, opac, shs = _common(gen, P)
    bg = torch.rand(3, generator=gen)
    views = []
    for i in range(n_views):
        th = 2.0 * math.pi * i / n_views
        C = np.array([6.0 * math.cos(th), -1.5, 6.0 * math.sin(th)])
        fwd = -C / np.linalg.norm(C)
        down = np.array([0.0, 1.0, 0.0])
        right = np.cross(down, fwd)
        right /= np.linalg.norm(right)
        down = np.cross(fwd, right)
        Rw2c = np.stack([right, down, fwd], axis=0)
        w2c = np.eye(4)
        w2c[:3, :3] = Rw2c
        w2c[:3, 3] = -Rw2c @ C
        views.append(make_view(w2c, W, H, tanx))
    return Scene(name, means, scales.contiguous(), rot, opac, shs, bg, views)


# mu_px calibrated with the fp32 bit-model to the duplication targets of SURVEY 8d
# (R/P: C1 ~ 8, C2 ~ 2.5, C3 ~ 3, C4 ~ 3, C5 ~ 16); tests/test_gpu_fullsize.py prints the achieved ratios.
CONFIGS = {
    "C1": dict(k=1, P=10_000, W=256, H=256, mu_px=2.978, kind="cam"),
    "C2": dict(k=2, P=5_800_000, W=1297, H=840, mu_px=1.065, kind="cam"),
    "C3": dict(k=3, P=1_800_000, W=979, H=546, mu_px=1.279, kind="cam"),
    "C4": dict(k=4, P=3_000_000, W=1600, H=1064, mu_px=1.237, kind="orbit"),
    "C5": dict(k=5, P=20_000_000, W=3840, H=2160, mu_px=4.657, kind="cam"),
}


def make_config(name: str, P: Optional[int] = None, n_views: int = 64, scale_image: float = 1.0) -> Scene:
    """Build config `name`; P / image size can be scaled down for tests (same distributions)."""
    c = dict(CONFIGS[name])
    if P is not None:
        c["P"] = P
    W, H = max(16, int(round(c["W"] * scale_image))), max(16, int(round(c["H"] * scale_image)))
    if c["kind"] == "cam":
        return camera_relative_scene(name, c["k"], c["P"], W, H, c["mu_px"])
    return orbit_scene(name, c["k"], c["P"], W, H, c["mu_px"], n_views)
