"""Drop-in `fused_ssim` package over libpdgs.so.

Public surface of the reference's module (/root/reference/submodules/fused-ssim/fused_ssim/__init__.py, imported by
/root/reference/train.py:41 and called at :156): `fused_ssim(img1, img2, padding="same", train=True)` on [B,CH,H,W]
CUDA tensors, differentiable w.r.t. `img1` only, returning the mean of the SSIM map; `FusedSSIMMap` (map-level
autograd function) and `allowed_padding`.  The two device entry points the reference binds with pybind
(ext.cpp: `fusedssim`, `fusedssim_backward`) are `pdgs_b200.rasterizer.ssim_forward` / `ssim_backward`
(C ABI pdgs_ssim_forward / pdgs_ssim_backward).  No CPU path.

Difference by design: for padding="same" the mean is reduced inside the forward kernel and the backward kernel takes
the uniform cotangent as a scalar, so neither the SSIM map nor a dL/dmap tensor is written to HBM.
"""
import torch

from pdgs_b200 import rasterizer as _rz

allowed_padding = ["same", "valid"]
_C1, _C2 = 0.01 ** 2, 0.03 ** 2
_BORDER = 5  # window radius: "valid" drops this many pixels on every side


def fusedssim(C1, C2, img1, img2, train=True):
    """(ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12), as the reference's extension function."""
    return _rz.ssim_forward(C1, C2, img1, img2, train=train)


def fusedssim_backward(C1, C2, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12):
    return _rz.ssim_backward(C1, C2, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12)


def _interior(t):
    return t[:, :, _BORDER:-_BORDER, _BORDER:-_BORDER]


class FusedSSIMMap(torch.autograd.Function):
    """SSIM map [B,CH,H,W] ("same") or its interior ("valid"); gradient flows to img1 only."""

    @staticmethod
    def forward(ctx, C1, C2, img1, img2, padding="same", train=True):
        out = fusedssim(C1, C2, img1, img2, train)
        ctx.consts = (C1, C2, padding == "valid")
        ctx.save_for_backward(img1.detach(), img2, *out[1:])
        return _interior(out[0]) if padding == "valid" else out[0]

    @staticmethod
    def backward(ctx, grad_map):
        C1, C2, crop = ctx.consts
        img1, img2, *partials = ctx.saved_tensors
        if partials[0].numel() == 0:
            raise RuntimeError("fused_ssim(..., train=False) keeps no gradient information")
        if crop:  # the cropped border receives no gradient
            full = torch.zeros_like(img1)
            _interior(full).copy_(grad_map)
            grad_map = full
        return None, None, fusedssim_backward(C1, C2, img1, img2, grad_map, *partials), None, None, None


class _MeanSSIM(torch.autograd.Function):
    """mean(ssim_map) with the reduction inside the kernel and a scalar cotangent in backward."""

    @staticmethod
    def forward(ctx, img1, img2, train):
        total = torch.zeros(1, dtype=torch.float32, device=img1.device)
        _, *partials = _rz.ssim_forward(_C1, _C2, img1, img2, train=train, want_map=False, ssim_sum=total)
        ctx.save_for_backward(img1.detach(), img2, *partials)
        ctx.count = max(img1.numel(), 1)
        return (total / ctx.count).reshape(())

    @staticmethod
    def backward(ctx, grad_mean):
        img1, img2, *partials = ctx.saved_tensors
        if partials[0].numel() == 0:
            raise RuntimeError("fused_ssim(..., train=False) keeps no gradient information")
        g = fusedssim_backward(_C1, _C2, img1, img2, 1.0 / ctx.count, *partials)
        return g * grad_mean, None, None


def fused_ssim(img1, img2, padding="same", train=True):
    assert padding in allowed_padding
    if padding == "same":
        return _MeanSSIM.apply(img1, img2, train)
    return FusedSSIMMap.apply(_C1, _C2, img1, img2, padding, train).mean()
