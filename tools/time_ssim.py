#!/usr/bin/env python
"""Time the fused training loss (pdgs_image_loss_ssim) and the L1-only loss at an image size; CUDA events."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "priordepth-gaussiansplatting_b200")):
    sys.path.insert(0, p)
import torch
from pdgs_b200 import rasterizer as rz
W, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1600, 1064)
dev = torch.device("cuda:0")
a, b = torch.rand(3, H, W, device=dev), torch.rand(3, H, W, device=dev)
acc = torch.zeros(3, device=dev)
scratch = torch.empty(rz.L.lib().pdgs_ssim_scratch_bytes(W, H, 3), dtype=torch.uint8, device=dev)
def t(fn, n=50):
    for _ in range(5): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
ms_ssim = t(lambda: rz.image_loss_ssim(a, b, lambda_dssim=0.2, loss_acc=acc, scratch=scratch))
ms_l1 = t(lambda: rz.image_loss(a, b, loss_acc=acc[:2]))
n = 3 * H * W
print(f"{W}x{H}: L1+SSIM loss+grad {ms_ssim*1e3:.1f} us ({44*n/ms_ssim/1e6:.0f} GB/s of 44 B/pixel-channel), L1 only {ms_l1*1e3:.1f} us")
