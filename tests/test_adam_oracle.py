"""CPU tests of oracle/adam_oracle.py: the bias-corrected mode is pinned against torch.optim.Adam itself (the reference's
default optimiser, /root/reference/scene/gaussian_model.py:194-195: Adam(l, lr=0.0, eps=1e-15) with per-group lr),
the sparse rule against its defining properties."""
import numpy as np
import torch

from oracle.adam_oracle import sparse_adam_step


def _tensors(N, M, seed):
    rng = np.random.default_rng(seed)
    p = rng.standard_normal((N, M)).astype(np.float32)
    return p


def test_bias_corrected_all_visible_matches_torch_adam():
    N, M, lr = 257, 45, 1.6e-4
    p0 = _tensors(N, M, 0)
    tp = torch.nn.Parameter(torch.from_numpy(p0.copy()))
    opt = torch.optim.Adam([{"params": [tp], "lr": lr}], lr=0.0, eps=1e-15)
    p, m, v = p0.copy(), np.zeros_like(p0), np.zeros_like(p0)
    vis = np.ones(N, bool)
    for step in range(1, 6):
        g = (_tensors(N, M, 10 + step) * 1e-3).astype(np.float32)
        tp.grad = torch.from_numpy(g.copy())
        opt.step()
        sparse_adam_step(p, g, m, v, vis, lr, eps=1e-15, step=step)
        ref = tp.detach().numpy()
        # torch uses lerp / addcdiv (different rounding order): agree to a few ulp of the update
        assert np.abs(p - ref).max() <= 1e-6 * lr / 1e-4 + 4e-7 * np.abs(ref).max()
        st = opt.state[tp]
        rm, rv = st["exp_avg"].numpy(), st["exp_avg_sq"].numpy()
        np.testing.assert_allclose(m, rm, rtol=2e-6, atol=5e-7 * np.abs(rm).max())  # sums of opposite signs cancel
        np.testing.assert_allclose(v, rv, rtol=2e-6, atol=0)


def test_sparse_rule_properties():
    N, M, lr = 100, 3, 0.01
    p0 = _tensors(N, M, 1)
    g = _tensors(N, M, 2)
    vis = np.random.default_rng(3).random(N) < 0.4
    p, m, v = p0.copy(), np.zeros_like(p0), np.zeros_like(p0)
    sparse_adam_step(p, g, m, v, vis, lr)
    # invisible rows untouched (parameter and both moments)
    assert np.array_equal(p[~vis], p0[~vis]) and not m[~vis].any() and not v[~vis].any()
    # first step from zero moments, no bias correction: m = 0.1 g, v = 0.001 g^2 -> |update| = lr * 0.1 / sqrt(0.001)
    upd = p0[vis] - p[vis]
    np.testing.assert_allclose(upd, np.sign(g[vis]) * lr * 0.1 / np.sqrt(0.001), rtol=1e-4)
    # zero gradient with zero moments: 0 / (0 + eps) = 0, parameters stay finite and unchanged
    p2, m2, v2 = p0.copy(), np.zeros_like(p0), np.zeros_like(p0)
    sparse_adam_step(p2, np.zeros_like(p0), m2, v2, np.ones(N, bool), lr)
    assert np.array_equal(p2, p0)
