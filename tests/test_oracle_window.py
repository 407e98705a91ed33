"""CPU check of oracle/window.py: inside its tile window, the windowed oracle (only the Gaussians whose rect touches
the window) reproduces the whole-view oracle -- image, inverse depth, n_contrib -- and, for a cotangent supported on the
window's pixels, its gradients.  This is what lets the full-size GPU tests and bench.py's CPU baseline run the float64
oracle on a window of a view that is far too large to blend whole."""
import torch

import helpers
from oracle import window as ow
from pdgs_b200 import scenes


def test_windowed_oracle_equals_whole_view_oracle_inside_the_window():
    sc = scenes.camera_relative_scene("w", 5, 1500, 112, 80, 3.0)
    view = sc.views[0]
    W, H = view.image_width, view.image_height
    full, res_full, st = helpers.run_oracle(sc, view)
    win = ow.place_window(W, H, (3, 2), (0.6, 0.4))
    assert win == (2, 1, 5, 3)
    o, ids, info = ow.window_oracle(sc, view, win, st=st)
    res = o.forward()
    ys, xs = ow.window_pixels(win, W, H)
    assert 0 < ids.numel() < sc.P and 0 < info["pairs_window"] < info["pairs_view"]
    assert torch.equal(res.color[:, ys, xs], res_full.color[:, ys, xs])
    assert torch.equal(res.invdepth[:, ys, xs], res_full.invdepth[:, ys, xs])
    assert torch.equal(res.n_contrib[ys, xs], res_full.n_contrib[ys, xs])
    gen = torch.Generator().manual_seed(1)
    m = torch.zeros(1, H, W, dtype=torch.float64)
    m[:, ys, xs] = 1.0
    gc = torch.randn(3, H, W, generator=gen, dtype=torch.float64) * m
    gd = torch.randn(1, H, W, generator=gen, dtype=torch.float64) * m
    g_full, g_win = full.backward(gc, gd), o.backward(gc, gd)
    other = torch.ones(sc.P, dtype=torch.bool)
    other[ids] = False
    for k in ("means3D", "means2D", "shs", "opacities", "scales", "rotations"):
        assert float(g_full[k][other].abs().max()) == 0.0, k
        assert torch.allclose(g_full[k][ids], g_win[k], rtol=1e-12, atol=1e-18), k
