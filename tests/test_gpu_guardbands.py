"""Out-of-bounds WRITE check of the CUDA path (compute-sanitizer is not available on the GPU pool).

Every device buffer the host side allocates for the C ABI -- geometry / image / binning state, backward scratch, outputs and
gradients, SSIM scratch, sort buffers -- is replaced by a slice of a larger allocation whose 4 KiB on either side hold a
canary pattern.  After forward, loss, backward and the batch step the canaries must be intact and the results must still match
an un-guarded run.  Sizes are deliberately ragged (image not a multiple of the 16x16 tile, P not a multiple of any tile
size, a pair count that ends mid-tile) because that is where an off-by-one of a staged or vectorised kernel would land.
"""
import math

import pytest
import torch

import helpers
from pdgs_b200 import rasterizer as rz
from pdgs_b200 import scenes

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GUARD = 4096
CANARY = 0xA5


class GuardedTorch:
    """Stands in for the `torch` module inside pdgs_b200: CUDA allocations get canary bands, everything else passes through."""

    def __init__(self):
        self.allocs = []

    def __getattr__(self, name):
        return getattr(torch, name)

    def _guarded(self, shape, dtype, device, fill=None):
        n = math.prod(shape)
        nbytes = n * torch.empty(0, dtype=dtype).element_size()
        pad = (-nbytes) % 16  # keep the trailing band 16-byte aligned like the body
        raw = torch.full((nbytes + pad + 2 * GUARD,), CANARY, dtype=torch.uint8, device=device)
        body = raw[GUARD:GUARD + nbytes].view(dtype).view(shape)
        if fill is not None:
            body.fill_(fill)
        self.allocs.append((raw, nbytes))
        return body

    @staticmethod
    def _shape(args):
        if len(args) == 1 and isinstance(args[0], (tuple, list, torch.Size)):
            return tuple(int(x) for x in args[0])
        return tuple(int(x) for x in args)

    def _alloc(self, fn, fill, args, kw):
        device = kw.get("device")
        if device is None or torch.device(device).type != "cuda" or kw.get("pin_memory"):
            return fn(*args, **kw)
        return self._guarded(self._shape(args), kw.get("dtype", torch.float32), device, fill)

    def empty(self, *args, **kw):
        return self._alloc(torch.empty, None, args, kw)

    def zeros(self, *args, **kw):
        return self._alloc(torch.zeros, 0, args, kw)

    def empty_like(self, t, **kw):
        if not t.is_cuda or kw:
            return torch.empty_like(t, **kw)
        return self._guarded(tuple(t.shape), t.dtype, t.device)

    def zeros_like(self, t, **kw):
        if not t.is_cuda or kw:
            return torch.zeros_like(t, **kw)
        return self._guarded(tuple(t.shape), t.dtype, t.device, 0)

    def check(self):
        torch.cuda.synchronize()
        assert self.allocs, "no guarded allocation was made: the patch did not take"
        for raw, nbytes in self.allocs:
            lo, hi = raw[:GUARD], raw[GUARD + nbytes:]
            assert bool((lo == CANARY).all()), f"write below a {nbytes}-byte buffer"
            assert bool((hi == CANARY).all()), f"write past the end of a {nbytes}-byte buffer"
        return len(self.allocs)


@pytest.fixture
def guarded(monkeypatch):
    g = GuardedTorch()
    monkeypatch.setattr(rz, "torch", g)
    return g


def ragged_scene(P, W, H, seed=11, mu_px=3.0):
    return scenes.camera_relative_scene("ragged", seed, P, W, H, mu_px)


def fwd_bwd(sc, view, antialiasing=False, colors_precomp=None, cov3d=None):
    lv, color, radii, invd = helpers.run_cuda(sc, view, DEV, antialiasing=antialiasing, colors_precomp=colors_precomp, cov3d=cov3d)
    gen = torch.Generator().manual_seed(5)
    dcol = torch.randn(color.shape, generator=gen).to(DEV)
    dinv = torch.randn(invd.shape, generator=gen).to(DEV)
    (color * dcol).sum().add((invd * dinv).sum()).backward()
    grads = {k: v.grad.clone() for k, v in lv.items() if v is not None and v.grad is not None}
    return color.detach().clone(), radii.clone(), invd.detach().clone(), grads


@pytest.mark.parametrize("P,W,H", [(3001, 250, 131), (1, 17, 9), (777, 16, 16), (5003, 641, 359)])
def test_forward_backward_guarded(guarded, P, W, H):
    sc = ragged_scene(P, W, H)
    view = sc.views[0]
    color, radii, invd, grads = fwd_bwd(sc, view)
    n = guarded.check()
    assert n >= 8  # outputs, state buffers, scratch, gradients
    # same results as an un-guarded run (radii / images bit-exact: same kernels, same inputs)
    rz.torch = torch
    color0, radii0, invd0, grads0 = fwd_bwd(sc, view)
    assert torch.equal(radii, radii0)
    assert torch.equal(color, color0) and torch.equal(invd, invd0)
    for k in grads0:
        assert helpers.rel_err(grads[k], grads0[k]) < 1e-4, k


def test_precomputed_inputs_guarded(guarded):
    sc = ragged_scene(2050, 203, 77, seed=12)
    view = sc.views[0]
    gen = torch.Generator().manual_seed(2)
    cols = torch.rand(sc.P, 3, generator=gen)
    fwd_bwd(sc, view, antialiasing=True, colors_precomp=cols)
    guarded.check()


def test_split_raw_guarded(guarded):
    from diff_gaussian_rasterization import GaussianRasterizer
    sc = ragged_scene(4099, 333, 211, seed=13)
    view = sc.views[0]
    rs = helpers.cuda_settings(sc, view, DEV)
    leaf = lambda t: t.to(DEV).clone().requires_grad_(True)  # noqa: E731
    lv = dict(means3D=leaf(sc.means3D), means2D=torch.zeros(sc.P, 3, device=DEV, requires_grad=True),
              opacities=leaf(torch.logit(sc.opacities.clamp(1e-6, 1 - 1e-6))), scales=leaf(torch.log(sc.scales)),
              rotations=leaf(sc.rotations * 1.3), dc=leaf(sc.shs[:, :1, :].contiguous()), shs=leaf(sc.shs[:, 1:, :].contiguous()))
    color, radii, invd = GaussianRasterizer(rs)(raw_params=True, **lv)
    (color.sum() + invd.sum()).backward()
    guarded.check()
    assert all(torch.isfinite(v.grad).all() for v in lv.values())


def test_losses_guarded(guarded):
    W, H = 203, 77
    gen = torch.Generator().manual_seed(4)
    r = lambda *s: torch.rand(*s, generator=gen).to(DEV)  # noqa: E731
    color, target, invd, prior = r(3, H, W), r(3, H, W), r(1, H, W), r(1, H, W)
    mask = (r(1, H, W) > 0.3).float()
    rz.image_loss(color, target, invd, prior, mask, 1.0, 0.5)
    rz.image_loss_ssim(color, target, invd, prior, mask, 0.2, 0.5)
    guarded.check()


@pytest.mark.parametrize("n", [1, 4097, 70001])
def test_sort_pairs_guarded(guarded, n):
    gen = torch.Generator().manual_seed(n)
    keys = torch.randint(0, 2 ** 40, (n,), generator=gen, dtype=torch.int64).to(DEV)
    vals = torch.arange(n, dtype=torch.int32).to(DEV)
    ref, idx = torch.sort(keys.cpu(), stable=True)  # before the call: sort_pairs uses its inputs as ping-pong buffers
    ko, vo = rz.sort_pairs(keys, vals, 40)
    guarded.check()
    assert torch.equal(ko.cpu().to(torch.int64), ref)
    assert torch.equal(vo.cpu().to(torch.int64), idx)


@pytest.mark.parametrize("split_raw", [False, True])
def test_batch_step_guarded(guarded, monkeypatch, split_raw):
    """The view-parallel batch step (three views in flight, shared gradient buffers, deferred SH backward)."""
    from pdgs_b200 import viewparallel as vpm
    monkeypatch.setattr(vpm, "torch", guarded)
    sc = scenes.make_config("C4", P=7001, n_views=5, scale_image=0.11)
    H, W = sc.views[0].image_height, sc.views[0].image_width
    gen = torch.Generator().manual_seed(11)
    targets = {i: torch.rand(3, H, W, generator=gen).to(DEV) for i in range(5)}
    priors = {i: torch.rand(1, H, W, generator=gen).to(DEV) for i in range(5)}
    g = {k: getattr(sc, k).to(DEV) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    if split_raw:
        g = dict(means3D=g["means3D"], opacities=torch.logit(g["opacities"].clamp(1e-6, 1 - 1e-6)), scales=torch.log(g["scales"]),
                 rotations=g["rotations"] * 1.3, shs_dc=g["shs"][:, :1].contiguous(), shs_rest=g["shs"][:, 1:].contiguous(),
                 raw_params=True)
    views = [helpers.cuda_settings(sc, v, DEV) for v in sc.views]
    vp = vpm.ViewParallelRasterizer(sc.P, 16, DEV, lanes=3)
    for lam in (0.0, 0.2):
        vp.step(g, views, targets, priors, None, 0.5, lambda_dssim=lam)
    guarded.check()
    assert torch.isfinite(vp.buf.flat).all()


def test_capacity_overflow_resume_guarded(guarded):
    """A binning buffer sized from an earlier, smaller view: the fused scan + duplicate kernel must stop at the capacity,
    report the true pair count, and the resumed call (stand-alone duplicate over the stored offsets) must give the same result."""
    sc = ragged_scene(5003, 641, 359)
    view = sc.views[0]
    ref = fwd_bwd(sc, view)
    key = (torch.device(DEV), view.image_width, view.image_height)
    hint = rz._capacity_hint[key]
    assert hint > 6000
    rz._capacity_hint[key] = 1200.0    # capacity 1200 * 1.25 + 4096 pairs: far below this view's pair count
    misses = rz.binning_stats()["misses"]
    out = fwd_bwd(sc, view)
    guarded.check()
    assert rz.binning_stats()["misses"] == misses + 1
    assert rz._capacity_hint[key] == hint
    assert torch.equal(out[1], ref[1]) and torch.equal(out[0], ref[0]) and torch.equal(out[2], ref[2])
    for k in ref[3]:
        assert helpers.rel_err(out[3][k], ref[3][k]) < 1e-4, k
