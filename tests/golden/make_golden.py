"""Generate tests/golden/ref_substeps.npz by IMPORTING the reference's own Python maths.

Run in the authoring container only (needs /root/reference; the GPU box does not have it):
    python tests/golden/make_golden.py

The rasterizer kernels themselves have no source in /root/reference (empty submodule), so the
golden vectors pin the sub-steps the reference DOES carry in-tree:
  * eval_sh (+0.5, clamp)           /root/reference/utils/sh_utils.py:57-112,
                                    /root/reference/gaussian_renderer/__init__.py:76-80
  * build_scaling_rotation,
    strip_symmetric (cov3D packing) /root/reference/utils/general_utils.py:64-110,
                                    /root/reference/scene/gaussian_model.py:34-38
  * getWorld2View2,
    getProjectionMatrix, transposes,
    full_proj, camera centre        /root/reference/utils/graphics_utils.py:38-71,
                                    /root/reference/scene/cameras.py:80-89
general_utils hard-codes device="cuda"; torch.zeros is patched to ignore `device` while the
reference functions run (there is no GPU in the authoring container).
"""
import math
import os
import sys
from unittest import mock

import numpy as np
import torch

REF = "/root/reference"
sys.path.insert(0, REF)

from utils.sh_utils import eval_sh  # noqa: E402
from utils import general_utils as gu  # noqa: E402
from utils.graphics_utils import getWorld2View2, getProjectionMatrix, focal2fov  # noqa: E402

_zeros = torch.zeros


def _zeros_cpu(*a, **k):
    k.pop("device", None)
    return _zeros(*a, **k)


def main():
    g = torch.Generator().manual_seed(20261104)
    out = {}
    # ---- SH -> RGB, every degree, as render() does with convert_SHs_python
    N = 64
    xyz = torch.randn(N, 3, generator=g, dtype=torch.float64) * 3
    campos = torch.tensor([0.3, -0.2, 0.5], dtype=torch.float64)
    feats = torch.cat([torch.randn(N, 1, 3, generator=g, dtype=torch.float64),
                       0.3 * torch.randn(N, 15, 3, generator=g, dtype=torch.float64)], dim=1)  # [N,16,3]
    out["sh_xyz"], out["sh_campos"], out["sh_feats"] = xyz.numpy(), campos.numpy(), feats.numpy()
    for deg in range(4):
        shs_view = feats.transpose(1, 2).view(-1, 3, 16)
        d = xyz - campos.repeat(N, 1)
        d = d / d.norm(dim=1, keepdim=True)
        sh2rgb = eval_sh(deg, shs_view, d)
        out[f"sh_rgb_deg{deg}"] = torch.clamp_min(sh2rgb + 0.5, 0.0).numpy()
        out[f"sh_raw_deg{deg}"] = (sh2rgb + 0.5).numpy()
    # ---- covariance from scaling / rotation (GaussianModel.build_covariance_from_scaling_rotation)
    s = torch.exp(torch.randn(N, 3, generator=g)).float()
    q = torch.randn(N, 4, generator=g).float()
    q = q / q.norm(dim=1, keepdim=True)  # get_rotation normalises before the rasterizer sees it
    with mock.patch.object(torch, "zeros", _zeros_cpu):
        for mod in (1.0, 0.7):
            Lm = gu.build_scaling_rotation(mod * s, q)
            cov = Lm @ Lm.transpose(1, 2)
            out[f"cov_mod{mod}"] = gu.strip_symmetric(cov).numpy()
    out["cov_scales"], out["cov_rots"] = s.numpy(), q.numpy()
    # ---- camera matrices (scene/cameras.py:80-89)
    Rs, Ts = [], []
    for i in range(4):
        A = torch.randn(3, 3, generator=g, dtype=torch.float64).numpy()
        Q, _ = np.linalg.qr(A)
        if np.linalg.det(Q) < 0:
            Q[:, 0] = -Q[:, 0]
        Rs.append(Q)
        Ts.append(torch.randn(3, generator=g, dtype=torch.float64).numpy() * 2)
    W, H, focal = 979, 546, 1100.0
    FoVx, FoVy = focal2fov(focal, W), focal2fov(focal, H)
    out["cam_R"], out["cam_T"] = np.stack(Rs), np.stack(Ts)
    out["cam_fov"] = np.array([FoVx, FoVy, 0.01, 100.0])
    wvts, fulls, centers = [], [], []
    for R, T in zip(Rs, Ts):
        wvt = torch.tensor(getWorld2View2(R, T, np.array([0.0, 0.0, 0.0]), 1.0)).transpose(0, 1)
        proj = getProjectionMatrix(znear=0.01, zfar=100.0, fovX=FoVx, fovY=FoVy).transpose(0, 1)
        full = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0)
        wvts.append(wvt.numpy())
        fulls.append(full.numpy())
        centers.append(wvt.inverse()[3, :3].numpy())
    out["cam_wvt"], out["cam_full"], out["cam_center"] = np.stack(wvts), np.stack(fulls), np.stack(centers)
    out["cam_tan"] = np.array([math.tan(FoVx * 0.5), math.tan(FoVy * 0.5)])
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_substeps.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
