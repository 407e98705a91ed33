"""Generate tests/golden/ref_ssim.npz by IMPORTING the reference's own SSIM / L1 (pure torch, CPU-runnable).

Run in the authoring container only (needs /root/reference; the GPU box does not have it):
    python tests/golden/make_golden_ssim.py

Pins the image-loss row to the reference's CPU-capable statement of it:
  * ssim, _ssim, create_window, gaussian   /root/reference/utils/loss_utils.py:46-89
  * l1_loss                                /root/reference/utils/loss_utils.py:40-41
  * the training combination               /root/reference/train.py:153-159
For every case the npz holds the inputs (fp32), the reference's SSIM value, the loss
(1-l)*L1 + l*(1-SSIM) and d(loss)/d(img1) computed by the reference code + autograd, once in the reference's
own fp32 and once with the same functions fed float64 tensors (tight pin for the float64 oracle).
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from utils.loss_utils import l1_loss, ssim  # noqa: E402

LAMBDA = 0.2  # /root/reference/arguments/__init__.py lambda_dssim default
CASES = [("c37x53", 3, 37, 53, 11), ("c16x24", 3, 16, 24, 12), ("c9x7", 1, 9, 7, 13), ("c64x48", 3, 64, 48, 14)]


def run(img1, img2):
    img1 = img1.clone().requires_grad_(True)
    s = ssim(img1, img2)
    loss = (1.0 - LAMBDA) * l1_loss(img1, img2) + LAMBDA * (1.0 - s)
    (g,) = torch.autograd.grad(loss, img1, retain_graph=True)
    (gs,) = torch.autograd.grad(s, img1)
    return s.detach(), loss.detach(), g, gs


def main():
    out = {"lambda_dssim": np.float64(LAMBDA)}
    for name, ch, h, w, seed in CASES:
        gen = torch.Generator().manual_seed(seed)
        a = torch.rand(ch, h, w, generator=gen)
        b = (a + 0.25 * torch.randn(ch, h, w, generator=gen)).clamp(0, 1)  # correlated pair, like render vs photo
        s32, l32, g32, gs32 = run(a, b)
        s64, l64, g64, gs64 = run(a.double(), b.double())
        out[f"{name}_img1"], out[f"{name}_img2"] = a.numpy(), b.numpy()
        out[f"{name}_ssim_f32"], out[f"{name}_loss_f32"] = s32.numpy(), l32.numpy()
        out[f"{name}_grad_f32"], out[f"{name}_gradssim_f32"] = g32.numpy(), gs32.numpy()
        out[f"{name}_ssim_f64"], out[f"{name}_loss_f64"] = s64.numpy(), l64.numpy()
        out[f"{name}_grad_f64"], out[f"{name}_gradssim_f64"] = g64.numpy(), gs64.numpy()
        print(name, float(s64), float(l64))
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_ssim.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
