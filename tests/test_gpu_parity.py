"""GPU parity tests (run with -m gpu on a B200): the CUDA path, driven through the drop-in Python API
and therefore through the C ABI of libpdgs.so, against the float64 oracle and the fp32 bit-model on
identical seeded scenes.

Tolerances (BASELINE.json north_star): radii / tile ranges / sorted lists bit-exact; image and
inverse depth <= 1e-4 max-abs; gradients <= 1e-3 relative.  Pixels the oracle marks `fragile`
(an accept/terminate threshold within rounding distance in float64) are compared separately:
there the fp32 kernel may legitimately take the other branch.
"""
import numpy as np
import pytest
import torch

import helpers
from oracle import bitmodel
from oracle import raster_oracle as ro
from pdgs_b200 import scenes

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
IMG_TOL = 1e-4
GRAD_REL = 1e-3


def _view_state(lv, radii, color, view):
    from pdgs_b200 import rasterizer as rz
    fn = color.grad_fn
    geom, binning, img = fn.saved_tensors[8], fn.saved_tensors[9], fn.saved_tensors[10]
    binning._pdgs_capacity = fn.binning_capacity
    return rz.state_view(geom, binning, img, radii.shape[0], view.image_width, view.image_height, fn.num_rendered), fn


def _check_ints(sc, view, lv, color, radii, st, scale_modifier=1.0, cov3d=None):
    sv, fn = _view_state(lv, radii, color, view)
    b = bitmodel.bin_and_sort(st, view.image_width, view.image_height)
    np.testing.assert_array_equal(radii.cpu().numpy(), st["radii"])
    vis = st["radii"] > 0
    assert fn.num_rendered == st["num_rendered"]
    np.testing.assert_array_equal(sv["tiles_touched"].numpy().view(np.uint32)[vis], st["tiles"][vis])
    np.testing.assert_array_equal(sv["xy"].numpy()[vis, :2].view(np.uint32), st["xy"][vis].view(np.uint32))
    np.testing.assert_array_equal(sv["depth_bits"].numpy()[vis].view(np.uint32), st["depth_bits"][vis])
    np.testing.assert_array_equal(sv["point_list"].numpy().view(np.uint32), b["point_list"])
    np.testing.assert_array_equal(sv["sorted_keys"].numpy().view(np.uint64), b["keys"])
    np.testing.assert_array_equal(sv["ranges"].numpy().view(np.uint32), b["ranges"])
    return sv


def _check_images(res, color, invd, sv, view):
    frag = res.fragile
    dc = (color.double().cpu() - res.color).abs().max(dim=0).values
    dd = (invd.double().cpu() - res.invdepth).abs()[0]
    assert float(frag.float().mean()) < 0.01
    assert float(dc[~frag].max()) <= IMG_TOL, float(dc[~frag].max())
    assert float(dd[~frag].max()) <= IMG_TOL * max(1.0, float(res.invdepth.abs().max())), float(dd[~frag].max())
    # a flipped 1/255 or 1e-4 threshold changes a pixel by at most ~alpha_min*max(rgb) resp. T_eps*max(rgb)
    assert float(dc.max()) <= 0.05
    H, W = view.image_height, view.image_width
    nc = sv["n_contrib"].view(H, W)
    assert torch.equal(nc[~frag], res.n_contrib[~frag])
    fT = sv["final_T"].view(H, W).double()
    assert float((fT - res.final_T)[~frag].abs().max()) <= 1e-5


def _check_grads(o, lv, color, invd, seed=0, tol=GRAD_REL):
    H, W = color.shape[1:]
    gen = torch.Generator().manual_seed(100 + seed)
    gc = torch.randn(3, H, W, generator=gen) / (3 * H * W)
    gd = torch.randn(1, H, W, generator=gen) / (H * W)
    ref = o.backward(gc, gd)
    torch.autograd.backward([color, invd], [gc.to(DEV), gd.to(DEV)])
    worst = {}
    for k, t in lv.items():
        if t is None:
            continue
        r = ref[k]
        g = t.grad
        assert g is not None, k
        assert g.shape == t.shape
        if float(r.abs().max()) == 0.0:
            assert float(g.abs().max()) == 0.0, k
            continue
        worst[k] = helpers.rel_err(g, r.view_as(g))
        assert worst[k] <= tol, (k, worst[k])
        assert helpers.elem_err(g, r.view_as(g), floor_frac=2e-3) <= 0.05, (k, helpers.elem_err(g, r.view_as(g), 2e-3))
    return worst


def test_c1_forward_backward_parity():
    """BASELINE config 1: 10k Gaussians, SH3, 256x256, fwd+bwd."""
    sc = scenes.make_config("C1")
    view = sc.views[0]
    o, res, st = helpers.run_oracle(sc, view)
    lv, color, radii, invd = helpers.run_cuda(sc, view, DEV)
    sv = _check_ints(sc, view, lv, color, radii, st)
    _check_images(res, color, invd, sv, view)
    worst = _check_grads(o, lv, color, invd)
    print("C1 grad rel errs", worst)


@pytest.mark.parametrize("variant", ["aa", "mod0.7", "deg0", "deg1", "deg2", "precomp", "ragged", "bg"])
def test_variants(variant):
    W, H = (200, 120)
    if variant == "ragged":
        W, H = 203, 117
    sc = scenes.camera_relative_scene("v", 11, 4000, W, H, 3.5)
    view = sc.views[0]
    kw = {}
    ckw = {}
    if variant == "aa":
        kw["antialiasing"] = True
    if variant == "mod0.7":
        kw["scale_modifier"] = 0.7
    if variant.startswith("deg"):
        kw["sh_degree"] = int(variant[3])
    if variant == "bg":
        sc.bg = torch.tensor([0.9, 0.2, 0.4])
    if variant == "precomp":
        gen = torch.Generator().manual_seed(5)
        colors = torch.rand(sc.P, 3, generator=gen)
        cov = ro.cov3d_from_scale_rot(sc.scales.double(), 1.0, sc.rotations.double()).float().contiguous()
        ckw = dict(colors_precomp=colors, cov3d=cov)
    o, res, st = helpers.run_oracle(sc, view, **kw, **ckw)
    lv, color, radii, invd = helpers.run_cuda(sc, view, DEV, **kw, **ckw)
    sv = _check_ints(sc, view, lv, color, radii, st, cov3d=ckw.get("cov3d"))
    _check_images(res, color, invd, sv, view)
    _check_grads(o, lv, color, invd, seed=hash(variant) % 97)


def test_orbit_view_world_relative_camera():
    """C4-shaped: world-space Gaussians, rotated/translated cameras, random background."""
    sc = scenes.make_config("C4", P=6000, n_views=5, scale_image=0.15)
    for vi in (0, 3):
        view = sc.views[vi]
        o, res, st = helpers.run_oracle(sc, view)
        lv, color, radii, invd = helpers.run_cuda(sc, view, DEV)
        sv = _check_ints(sc, view, lv, color, radii, st)
        _check_images(res, color, invd, sv, view)
        _check_grads(o, lv, color, invd, seed=vi)


def test_depth_prior_loss_step_c3_shape():
    """BASELINE config 3's loss: L1 image + |invdepth - prior| * mask (train.py:174), reduced P."""
    sc = scenes.make_config("C3", P=20000, scale_image=0.25)
    view = sc.views[0]
    H, W = view.image_height, view.image_width
    gen = torch.Generator().manual_seed(9)
    gt, prior = torch.rand(3, H, W, generator=gen), torch.rand(1, H, W, generator=gen)
    mask = (torch.rand(1, H, W, generator=gen) > 0.3).float()
    o, res, st = helpers.run_oracle(sc, view)
    lv, color, radii, invd = helpers.run_cuda(sc, view, DEV)
    loss = (color - gt.to(DEV)).abs().mean() + ((invd - prior.to(DEV)) * mask.to(DEV)).abs().mean()
    loss.backward()
    gc = torch.sign(res.color - gt.double()) / gt.numel()
    gd = torch.sign((res.invdepth - prior.double()) * mask.double()) * mask.double() / prior.numel()
    ref = o.backward(gc, gd)
    for k in ("means3D", "shs", "opacities", "scales", "rotations", "means2D"):
        # sign() of a residual can differ between fp32 and fp64 where the residual is ~0: allow a little more
        assert helpers.rel_err(lv[k].grad, ref[k].view_as(lv[k].grad)) <= 5e-3, k


def test_edge_cases():
    from diff_gaussian_rasterization import GaussianRasterizer
    sc = scenes.camera_relative_scene("e", 3, 64, 48, 32, 3.0)
    view = sc.views[0]
    sc.bg = torch.tensor([0.25, 0.5, 0.75])
    # (a) everything behind the camera
    behind = torch.stack([sc.means3D[:, 0], sc.means3D[:, 1], -sc.means3D[:, 2].abs() - 1.0], dim=1).contiguous()
    sc_b = scenes.Scene(**{**sc.__dict__, "means3D": behind})
    lv, color, radii, invd = helpers.run_cuda(sc_b, view, DEV)
    assert int(radii.max()) == 0
    assert torch.allclose(color.cpu(), sc.bg[:, None, None].expand(3, 32, 48))
    assert float(invd.abs().max()) == 0.0
    (color.sum() + invd.sum()).backward()
    for k, t in lv.items():
        if t is not None:
            assert t.grad is not None and float(t.grad.abs().max()) == 0.0, k
    # (b) P = 0
    rs = helpers.cuda_settings(sc, view, DEV)
    z = lambda *s: torch.zeros(*s, device=DEV, requires_grad=True)
    color, radii, invd = GaussianRasterizer(rs)(means3D=z(0, 3), means2D=z(0, 3), opacities=z(0, 1), shs=z(0, 16, 3),
                                                scales=z(0, 3), rotations=z(0, 4))
    assert radii.numel() == 0 and torch.allclose(color.cpu(), sc.bg[:, None, None].expand(3, 32, 48))
    # (c) one huge Gaussian covering every tile + one tiny one
    sc1 = scenes.camera_relative_scene("e1", 4, 2, 80, 48, 3.0, behind_frac=0.0)
    sc1.means3D[:] = torch.tensor([[0.0, 0.0, 2.0], [0.1, 0.1, 3.0]])
    sc1.scales[:] = torch.tensor([[5.0, 5.0, 5.0], [0.01, 0.01, 0.01]])
    v1 = sc1.views[0]
    o, res, st = helpers.run_oracle(sc1, v1)
    lv, color, radii, invd = helpers.run_cuda(sc1, v1, DEV)
    sv = _check_ints(sc1, v1, lv, color, radii, st)
    assert int(st["tiles"][0]) == 5 * 3
    _check_images(res, color, invd, sv, v1)
    _check_grads(o, lv, color, invd)
    # (d) wrong shape raises
    with pytest.raises(Exception, match="num_points, 3"):
        GaussianRasterizer(rs)(means3D=z(4, 2), means2D=z(4, 3), opacities=z(4, 1), shs=z(4, 16, 3), scales=z(4, 3),
                               rotations=z(4, 4))


def test_non_finite_covariance_is_culled():
    """NaN / inf scales or rotations (or exp() overflow of a raw scale) must cull the Gaussian, as upstream's
    (int)NaN == 0 radius does, instead of reaching the binning with an inverted tile rect."""
    from diff_gaussian_rasterization import GaussianRasterizer
    sc = scenes.camera_relative_scene("nan", 17, 3000, 160, 96, 3.0, behind_frac=0.0)
    view = sc.views[0]
    bad = torch.arange(0, 3000, 7)
    sc.scales[bad[0::4], 1] = float("nan")
    sc.scales[bad[1::4], 0] = float("inf")
    sc.rotations[bad[2::4], 2] = float("nan")
    sc.means3D[bad[3::4], 0] = float("nan")
    st = helpers.bit_structure(sc, view)
    assert int(st["radii"][bad.numpy()].max()) == 0
    lv, color, radii, invd = helpers.run_cuda(sc, view, DEV)
    np.testing.assert_array_equal(radii.cpu().numpy(), st["radii"])
    assert int(radii[bad.to(DEV)].max()) == 0 and int((radii > 0).sum()) > 2000
    assert torch.isfinite(color).all() and torch.isfinite(invd).all()
    sv, fn = _view_state(lv, radii, color, view)
    assert fn.num_rendered == st["num_rendered"]
    T = ((view.image_width + 15) // 16) * ((view.image_height + 15) // 16)
    assert int(sv["sorted_tiles"].max()) < T
    # identical to the scene without the bad Gaussians
    keep = torch.ones(sc.P, dtype=torch.bool)
    keep[bad] = False
    sub = scenes.Scene("sub", sc.means3D[keep].contiguous(), sc.scales[keep].contiguous(), sc.rotations[keep].contiguous(),
                       sc.opacities[keep].contiguous(), sc.shs[keep].contiguous(), sc.bg, sc.views)
    with torch.no_grad():
        _, c2, r2, d2 = helpers.run_cuda(sub, view, DEV)
    assert torch.equal(color.detach(), c2) and torch.equal(invd.detach(), d2)
    (color.sum() + invd.sum()).backward()
    for k in ("means3D", "shs", "opacities", "scales", "rotations"):
        assert float(lv[k].grad[bad.to(DEV)].abs().max()) == 0.0, k
    # raw-parameter mode: exp() of a huge raw scale overflows to inf inside the kernel
    rs = helpers.cuda_settings(sc, view, DEV)
    raw_scales = torch.log(sub.scales).to(DEV)
    raw_scales[::5, 0] = 200.0
    with torch.no_grad():
        c3, r3, d3 = GaussianRasterizer(rs)(means3D=sub.means3D.to(DEV), means2D=torch.zeros(sub.P, 3, device=DEV),
                                            opacities=torch.logit(sub.opacities.clamp(1e-6, 1 - 1e-6)).to(DEV),
                                            shs=sub.shs.to(DEV), scales=raw_scales, rotations=sub.rotations.to(DEV),
                                            raw_params=True)
    assert int(r3[::5].max()) == 0 and torch.isfinite(c3).all() and torch.isfinite(d3).all()


def test_prefiltered_is_rejected():
    """prefiltered=True traps upstream when a point is culled; the reference always passes False
    (/root/reference/gaussian_renderer/__init__.py:47).  Here it raises instead of being silently ignored."""
    from diff_gaussian_rasterization import GaussianRasterizer
    sc = scenes.camera_relative_scene("pf", 5, 100, 64, 64, 3.0)
    rs = helpers.cuda_settings(sc, sc.views[0], DEV)._replace(prefiltered=True)
    lv = helpers.cuda_leaves(sc, DEV)
    with pytest.raises(Exception, match="prefiltered"):
        GaussianRasterizer(rs)(**lv)


def test_forward_only_and_none_cotangents():
    sc = scenes.camera_relative_scene("n", 13, 3000, 128, 96, 3.0)
    view = sc.views[0]
    with torch.no_grad():
        lv, c0, r0, d0 = helpers.run_cuda(sc, view, DEV)
    lv, c1, r1, d1 = helpers.run_cuda(sc, view, DEV)
    assert torch.equal(c0, c1) and torch.equal(d0, d1) and torch.equal(r0, r1)  # forward is deterministic
    c1.sum().backward()  # invdepth cotangent is None (is_depth_feedback = False in the reference's train.py:68)
    g_sh = lv["shs"].grad.clone()
    lv2, c2, r2, d2 = helpers.run_cuda(sc, view, DEV)
    (c2.sum() + 0.0 * d2.sum()).backward()
    assert helpers.rel_err(lv2["shs"].grad, g_sh) < 1e-5


def test_binning_capacity_regrow_and_speculative_paths_agree():
    from pdgs_b200 import rasterizer as rz
    small = scenes.camera_relative_scene("s", 21, 500, 96, 64, 2.0)
    big = scenes.camera_relative_scene("b", 22, 20000, 320, 240, 6.0)
    rz._capacity_hint.clear()
    rz.binning_stats(reset=True)
    key_b = (torch.device(DEV), 320, 240)
    with torch.no_grad():
        _, cs, _, _ = helpers.run_cuda(small, small.views[0], DEV)   # no hint: NEED_BINNING -> resume
        _, cb0, rb0, db0 = helpers.run_cuda(big, big.views[0], DEV)  # no hint for this image size either
        rz._capacity_hint[key_b] = 100.0
        _, cb1, rb1, db1 = helpers.run_cuda(big, big.views[0], DEV)  # hint too small: speculative miss -> resume
        _, cb2, rb2, db2 = helpers.run_cuda(big, big.views[0], DEV)  # hint large enough: speculative hit
        _, cs2, _, _ = helpers.run_cuda(small, small.views[0], DEV)  # hit
    assert torch.equal(cb1, cb2) and torch.equal(db1, db2) and torch.equal(rb1, rb2) and torch.equal(cb0, cb2)
    assert torch.equal(cs, cs2)
    assert rz.binning_stats() == dict(calls=5, misses=1, sizing_calls=2)
    # the hint is a high-water mark that decays by 1 % per call: a smaller view afterwards keeps the capacity
    hwm = rz._capacity_hint[key_b]
    sub = scenes.Scene("sub", big.means3D[:2000].contiguous(), big.scales[:2000].contiguous(), big.rotations[:2000].contiguous(),
                       big.opacities[:2000].contiguous(), big.shs[:2000].contiguous(), big.bg, big.views)
    with torch.no_grad():
        helpers.run_cuda(sub, sub.views[0], DEV)
        helpers.run_cuda(big, big.views[0], DEV)
    assert abs(rz._capacity_hint[key_b] - hwm) <= 1e-6 * hwm and rz.binning_stats()["misses"] == 1


@pytest.mark.parametrize("n,end_bit", [(1, 41), (255, 41), (3071, 45), (3072, 45), (3073, 47), (100_003, 44),
                                       (2_000_000, 47), (50_000, 64), (50_000, 8), (50_000, 13)])
def test_onesweep_sort_matches_stable_sort(n, end_bit):
    from pdgs_b200 import rasterizer as rz
    gen = torch.Generator().manual_seed(n + end_bit)
    hi = torch.randint(0, 1 << 15, (n,), generator=gen, dtype=torch.int64)
    lo = torch.randint(0, 1 << 32, (n,), generator=gen, dtype=torch.int64)
    if n > 1000:
        lo[: n // 3] = lo[0]  # heavy duplicates exercise stability
    keys = (hi << 32) | lo
    if end_bit == 64:
        keys = keys | (torch.randint(0, 2, (n,), generator=gen, dtype=torch.int64) << 63)
    mask = -1 if end_bit == 64 else (1 << end_bit) - 1
    vals = torch.arange(n, dtype=torch.int32)
    ko, vo = rz.sort_pairs(keys.to(DEV), vals.to(DEV), end_bit)
    # expected: stable sort by the low end_bit bits as an unsigned number
    sub = keys & mask
    as_u = sub.numpy().view(np.uint64)
    order = np.argsort(as_u, kind="stable")
    np.testing.assert_array_equal(vo.cpu().numpy(), vals.numpy()[order])
    np.testing.assert_array_equal(ko.cpu().numpy(), keys.numpy()[order])


def test_work_counters_match_the_oracle_counts():
    """pdgs_set_work_counters: N_test / N_acc of SURVEY 8(d)'s FLOP count, as bench.py's roofline uses them, against the
    oracle's exact counts (off by default; the forward runs its counter-free variant otherwise)."""
    from pdgs_b200 import rasterizer as rz
    sc = scenes.make_config("C1")
    view = sc.views[0]
    o, res, st = helpers.run_oracle(sc, view, need_grad=False)
    rz.set_work_counters(DEV, True)
    try:
        with torch.no_grad():
            rs = helpers.cuda_settings(sc, view, DEV)
            empty = torch.empty(0, device=DEV)
            g = {k: getattr(sc, k).to(DEV) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
            R, color, radii, geom, binning, img, invd = rz.rasterize_gaussians(
                rs.bg, g["means3D"], empty, g["opacities"], g["scales"], g["rotations"], 1.0, empty, rs.viewmatrix,
                rs.projmatrix, rs.tanfovx, rs.tanfovy, view.image_height, view.image_width, g["shs"], 3, rs.campos, False,
                False, False)
            c = rz.image_counters(geom, binning, img, sc.P, view.image_width, view.image_height)
    finally:
        rz.set_work_counters(DEV, False)
    assert abs(c["n_accepted"] - res.n_acc) <= 2e-3 * res.n_acc
    assert abs(c["n_tested"] - res.n_test) <= 2e-3 * res.n_test
    assert abs(c["n_contrib"] - res.n_test_bwd) <= 2e-3 * res.n_test_bwd
    # with the counters off the two planes are never written: asking for them raises / yields None, not garbage
    with pytest.raises(Exception, match="work counters are off"):
        rz.image_counters(geom, binning, img, sc.P, view.image_width, view.image_height)
    sv = rz.state_view(geom, binning, img, sc.P, view.image_width, view.image_height, R)
    assert sv["n_tested"] is None and sv["n_accepted"] is None and sv["n_contrib"].numel() == view.image_width * view.image_height


def test_mark_visible():
    from diff_gaussian_rasterization import GaussianRasterizer
    sc = scenes.make_config("C4", P=5000, n_views=3, scale_image=0.1)
    view = sc.views[1]
    rs = helpers.cuda_settings(sc, view, DEV)
    got = GaussianRasterizer(rs).markVisible(sc.means3D.to(DEV)).cpu()
    want = ro.mark_visible(sc.means3D, view.world_view_transform)
    z = sc.means3D.double() @ view.world_view_transform.double()[:3, 2] + view.world_view_transform.double()[3, 2]
    safe = (z - ro.NEAR_Z).abs() > 1e-5
    assert torch.equal(got[safe], want[safe])


def test_debug_mode_runs():
    sc = scenes.camera_relative_scene("d", 5, 1000, 64, 64, 3.0)
    lv, color, radii, invd = helpers.run_cuda(sc, sc.views[0], DEV, debug=True)
    color.sum().backward()
    assert torch.isfinite(color).all()


def test_full_size_properties_c2():
    """BASELINE config 2 at full size (5.8M Gaussians, 1297x840): size-independent properties."""
    sc = scenes.make_config("C2")
    view = sc.views[0]
    st = helpers.bit_structure(sc, view)
    lv, color, radii, invd = helpers.run_cuda(sc, view, DEV)
    np.testing.assert_array_equal(radii.cpu().numpy(), st["radii"])
    sv, fn = _view_state(lv, radii, color, view)
    R = fn.num_rendered
    assert R == st["num_rendered"]
    keys = sv["sorted_keys"].numpy().view(np.uint64)
    assert bool((keys[1:] >= keys[:-1]).all())                       # sortedness
    pl = sv["point_list"].numpy().view(np.uint32).astype(np.int64)
    assert np.array_equal(np.bincount(pl, minlength=sc.P), st["tiles"].astype(np.int64))  # a permutation of the pairs
    same = keys[1:] == keys[:-1]
    assert bool((pl[1:][same] > pl[:-1][same]).all())                # stability (ids ascend within equal keys)
    rg = sv["ranges"].numpy().view(np.uint32).astype(np.int64)
    ne = rg[:, 1] > rg[:, 0]
    assert int((rg[ne, 1] - rg[ne, 0]).sum()) == R                   # ranges partition the list
    tiles = (keys >> np.uint64(32)).astype(np.int64)
    assert np.array_equal(tiles[rg[ne, 0]], np.nonzero(ne)[0]) and np.array_equal(tiles[rg[ne, 1] - 1], np.nonzero(ne)[0])
    assert torch.isfinite(color).all() and torch.isfinite(invd).all()
    assert float(color.min()) >= 0.0
    fT = sv["final_T"]
    assert float(fT.min()) >= 9.0e-5 and float(fT.max()) <= 1.0
    # backward linearity: grads(2g) == 2 grads(g) up to atomic reordering
    gen = torch.Generator().manual_seed(1)
    gc = (torch.randn(color.shape, generator=gen) / color.numel()).to(DEV)
    color.backward(gc, retain_graph=True)
    g1 = {k: t.grad.clone() for k, t in lv.items() if t is not None}
    for t in lv.values():
        if t is not None:
            t.grad = None
    color.backward(2 * gc)
    for k, t in lv.items():
        if t is not None:
            assert torch.isfinite(t.grad).all(), k
            assert helpers.rel_err(t.grad, 2 * g1[k]) < 1e-4, k
    # culled Gaussians get exactly zero gradient rows
    culled = torch.from_numpy(st["radii"] == 0).to(DEV)
    assert float(lv["shs"].grad[culled].abs().max()) == 0.0 and float(lv["means3D"].grad[culled].abs().max()) == 0.0


def test_first_slot_of_a_batch_is_the_only_contributor():
    """The forward blend remembers the last blended list entry by the shared-memory address of its staged record; slot 0
    of a 256-entry batch is the lowest such address, so "none" must not be encoded as a value that slot can take.
    (a) one Gaussian: every covered pixel's only contributor is slot 0 of the first batch;
    (b) 257 Gaussians in one tile, the farthest alone at its own spot: the pixels there see only list entry 256 = slot 0
        of the SECOND batch.  n_contrib, image and gradients against the oracle."""
    # (a)
    sc = scenes.camera_relative_scene("s0", 5, 1, 32, 32, 3.0, behind_frac=0.0)
    sc.means3D[:] = torch.tensor([[0.05, -0.05, 2.0]])
    sc.scales[:] = torch.tensor([[0.05, 0.04, 0.05]])
    sc.opacities[:] = 0.9
    v = sc.views[0]
    o, res, st = helpers.run_oracle(sc, v)
    lv, color, radii, invd = helpers.run_cuda(sc, v, DEV)
    sv = _check_ints(sc, v, lv, color, radii, st)
    _check_images(res, color, invd, sv, v)
    nc = sv["n_contrib"].view(v.image_height, v.image_width)
    assert int(nc.max()) == 1 and int((nc == 1).sum()) >= 10
    _check_grads(o, lv, color, invd)
    assert float(lv["means3D"].grad.abs().max()) > 0
    # (b)
    P = 257
    sc = scenes.camera_relative_scene("s1", 6, P, 16, 16, 1.0, behind_frac=0.0)
    focal = 16 / (2.0 * 0.625)
    z = 2.0 + 0.01 * torch.arange(P, dtype=torch.float32)
    spot = torch.tensor([[-4.0, -4.0]]).repeat(P, 1)   # pixel offsets from the image centre
    spot[P - 1] = torch.tensor([4.0, 4.0])             # the farthest Gaussian sits alone
    sc.means3D[:] = torch.cat([spot * z[:, None] / focal, z[:, None]], dim=1)
    sc.scales[:] = (z / focal)[:, None] * 0.6
    sc.opacities[:] = 0.02                             # the stack stays far from saturating, so all 257 are listed
    sc.opacities[P - 1] = 0.9
    v = sc.views[0]
    o, res, st = helpers.run_oracle(sc, v)
    lv, color, radii, invd = helpers.run_cuda(sc, v, DEV)
    sv = _check_ints(sc, v, lv, color, radii, st)
    assert int(sv["ranges"].view(-1, 2)[0, 1]) == P and int(sv["point_list"][P - 1]) == P - 1
    _check_images(res, color, invd, sv, v)
    nc = sv["n_contrib"].view(16, 16)
    assert int(nc[12, 12]) == P and int(res.n_contrib[12, 12]) == P     # the lone Gaussian's centre pixel
    assert float(res.final_T[12, 12]) < 0.5
    _check_grads(o, lv, color, invd)
    assert float(lv["opacities"].grad[P - 1].abs()) > 0
