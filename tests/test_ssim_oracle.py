"""CPU: the SSIM / training-loss oracle against golden vectors produced by the reference's own Python
(tests/golden/make_golden_ssim.py imports /root/reference/utils/loss_utils.py)."""
import os

import numpy as np
import pytest
import torch

from oracle import ssim_oracle as so

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_ssim.npz"))
CASES = ["c37x53", "c16x24", "c9x7", "c64x48"]


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_ssim_and_loss(name):
    a = torch.from_numpy(G[f"{name}_img1"]).double().requires_grad_(True)
    b = torch.from_numpy(G[f"{name}_img2"]).double()
    lam = float(G["lambda_dssim"])
    loss, parts = so.training_loss(a, b, lam)
    # The reference rounds the 2-D window (outer product) to fp32 (loss_utils.py:52 `.float()`); the separable form
    # used by the oracle, by fused-ssim (ssim.cu:10-20) and by the CUDA kernel filters with the fp32 1-D taps
    # twice.  The two windows differ by fp32 rounding of each tap: ~1e-8 relative on the result.
    assert abs(float(parts["ssim"]) - float(G[f"{name}_ssim_f64"])) < 2e-8
    assert abs(float(loss) - float(G[f"{name}_loss_f64"])) < 2e-8
    (g,) = torch.autograd.grad(loss, a)
    ref = G[f"{name}_grad_f64"]
    assert float(np.abs(g.numpy() - ref).max()) < 1e-6 * float(np.abs(ref).max())
    # and the reference's own fp32 run agrees with the float64 pin to fp32 accuracy
    assert abs(float(G[f"{name}_ssim_f32"]) - float(G[f"{name}_ssim_f64"])) < 2e-6


@pytest.mark.parametrize("name", CASES[:2])
def test_analytic_backward_equals_autograd(name):
    """The fused-ssim formulation (partial maps + adjoint filtering) is the gradient of mean(ssim)."""
    a = torch.from_numpy(G[f"{name}_img1"]).double()
    b = torch.from_numpy(G[f"{name}_img2"]).double()
    g = so.ssim_backward(a, b, torch.full_like(a, 1.0 / a.numel()))
    ref = G[f"{name}_gradssim_f64"]
    assert float(np.abs(g.numpy() - ref).max()) < 1e-6 * float(np.abs(ref).max())


def test_window_and_edge_shapes():
    w = so.window_1d()
    assert w.numel() == 11 and abs(float(w.sum()) - 1.0) < 1e-6 and float(w[5]) == pytest.approx(0.26601171493530273, abs=1e-9)
    # a 1x1 image: every filtered moment is the centre tap squared times the pixel statistics
    a, b = torch.tensor([[[0.3]]], dtype=torch.float64), torch.tensor([[[0.6]]], dtype=torch.float64)
    m = so.ssim_map(a, b)
    k = float(w[5]) ** 2
    mu1, mu2 = k * 0.3, k * 0.6
    s1, s2, s12 = k * 0.09 - mu1 ** 2, k * 0.36 - mu2 ** 2, k * 0.18 - mu1 * mu2
    ref = ((2 * mu1 * mu2 + so.C1) * (2 * s12 + so.C2)) / ((mu1 ** 2 + mu2 ** 2 + so.C1) * (s1 + s2 + so.C2))
    assert float(m) == pytest.approx(ref, rel=1e-12)
    # identical images -> SSIM 1 everywhere
    x = torch.rand(2, 13, 9, dtype=torch.float64)
    assert torch.allclose(so.ssim_map(x, x), torch.ones_like(x), atol=1e-12)
