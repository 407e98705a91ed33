"""CPU tests: the oracle's sub-steps against golden vectors produced by the reference's own Python
maths (tests/golden/make_golden.py), its internal consistency, and the fp32 bit-model."""
import math
import os

import numpy as np
import pytest
import torch

from oracle import bitmodel
from oracle import raster_oracle as ro
from pdgs_b200 import scenes
import helpers

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_substeps.npz"))


@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_sh_basis_matches_reference_eval_sh(deg):
    xyz, campos, feats = (torch.from_numpy(G[k]) for k in ("sh_xyz", "sh_campos", "sh_feats"))
    d = xyz - campos[None]
    d = d / d.norm(dim=1, keepdim=True)
    basis = ro.eval_sh_basis(deg, d)
    n = basis.shape[1]
    raw = torch.einsum("pk,pkc->pc", basis, feats[:, :n]) + 0.5
    np.testing.assert_allclose(raw.numpy(), G[f"sh_raw_deg{deg}"], rtol=0, atol=1e-13)
    np.testing.assert_allclose(raw.clamp_min(0).numpy(), G[f"sh_rgb_deg{deg}"], rtol=0, atol=1e-13)


@pytest.mark.parametrize("mod", [1.0, 0.7])
def test_cov3d_matches_reference_build_scaling_rotation(mod):
    s, q = torch.from_numpy(G["cov_scales"]).double(), torch.from_numpy(G["cov_rots"]).double()
    cov = ro.cov3d_from_scale_rot(s, mod, q)
    np.testing.assert_allclose(cov.numpy(), G[f"cov_mod{mod}"], rtol=2e-5, atol=2e-6)  # golden is fp32


def test_camera_matrices_match_reference():
    FoVx, FoVy, zn, zf = G["cam_fov"]
    for i in range(G["cam_R"].shape[0]):
        R, T = G["cam_R"][i], G["cam_T"][i]
        w2c = np.zeros((4, 4))
        w2c[:3, :3] = R.T
        w2c[:3, 3] = T
        w2c[3, 3] = 1.0
        wvt = torch.tensor(np.float32(w2c)).transpose(0, 1)
        proj = scenes.projection_matrix(zn, zf, FoVx, FoVy).transpose(0, 1)
        full = wvt @ proj
        np.testing.assert_allclose(wvt.numpy(), G["cam_wvt"][i], rtol=0, atol=1e-6)
        np.testing.assert_allclose(full.numpy(), G["cam_full"][i], rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose(wvt.inverse()[3, :3].numpy(), G["cam_center"][i], rtol=1e-5, atol=1e-5)


def test_oracle_projection_uses_reference_matrix_convention():
    """p_view = (x,y,z,1) @ world_view_transform must equal R^T-based COLMAP transform of the point."""
    R, T = G["cam_R"][0], G["cam_T"][0]
    wvt = torch.from_numpy(G["cam_wvt"][0]).double()
    p = torch.tensor([[0.3, -1.2, 2.5]], dtype=torch.float64)
    pv = (torch.cat([p, torch.ones(1, 1, dtype=torch.float64)], 1) @ wvt)[0, :3].numpy()
    np.testing.assert_allclose(pv, R.T @ p[0].numpy() + T, atol=1e-5)


def _tiny(P=48, W=40, H=24, seed=7, mu_px=3.0):
    sc = scenes.camera_relative_scene("tiny", seed, P, W, H, mu_px)
    return sc, sc.views[0]


def test_oracle_gradcheck_against_finite_differences():
    """autograd of the float64 forward vs central differences on a tiny scene (K7-K9 conventions)."""
    sc, view = _tiny()
    sc.opacities.clamp_(max=0.95)  # keep alpha below the 0.99 clamp, whose backward is straight-through by convention
    gen = torch.Generator().manual_seed(3)
    gc = torch.randn(3, view.image_height, view.image_width, generator=gen, dtype=torch.float64)
    gd = torch.randn(1, view.image_height, view.image_width, generator=gen, dtype=torch.float64)

    def loss_of(scene):
        o, res, _ = helpers.run_oracle(scene, view, need_grad=False, use_structure=False, round_pixel_centre=False)
        return float((res.color * gc).sum() + (res.invdepth * gd).sum()), res

    o, res, _ = helpers.run_oracle(sc, view, need_grad=True, use_structure=False, round_pixel_centre=False)
    grads = o.backward(gc, gd)
    base_nc = res.n_contrib.clone()
    eps = 1e-6
    checked = 0
    for name in ("means3D", "opacities", "scales", "rotations", "shs"):
        t = getattr(sc, name)
        flat = t.view(-1)
        idxs = torch.randperm(flat.numel(), generator=gen)[:12]
        for i in idxs.tolist():
            orig = flat[i].item()
            t64 = t.double().clone()
            vals = []
            ok = True
            for sgn in (+1, -1):
                t2 = t64.clone()
                t2.view(-1)[i] = orig + sgn * eps
                sc2 = scenes.Scene(**{**sc.__dict__, name: t2})
                l, r2 = loss_of(sc2)
                ok &= bool((r2.n_contrib == base_nc).all()) and bool((r2.radii == res.radii).all())
                vals.append(l)
            if not ok:
                continue  # a discrete decision flipped inside the stencil
            fd = (vals[0] - vals[1]) / (2 * eps)
            an = float(grads[name].reshape(-1)[i])
            assert abs(fd - an) <= 2e-5 * max(1.0, abs(an), abs(fd)) + 1e-7, (name, i, fd, an)
            checked += 1
    assert checked >= 30


def test_oracle_float32_leaves_vs_float64_scene_is_consistent():
    sc, view = _tiny(P=200, W=64, H=48)
    _, a, _ = helpers.run_oracle(sc, view, need_grad=False, use_structure=False)
    _, b, st = helpers.run_oracle(sc, view, need_grad=False, use_structure=True)
    # float64 integer side agrees with the fp32 bit-model except at rounding boundaries (rare)
    same = (a.radii == b.radii)
    assert same.float().mean() > 0.98
    if bool(same.all()) and torch.equal(a.rect, b.rect):
        assert torch.equal(a.point_list, b.point_list)


def test_bitmodel_matches_oracle_binning_bitwise():
    """Given the same integer structure, the C bit-model's sort + ranges equal the oracle's stable argsort."""
    sc, view = _tiny(P=3000, W=160, H=96, mu_px=4.0)
    o, res, st = helpers.run_oracle(sc, view, need_grad=False, use_structure=True)
    b = bitmodel.bin_and_sort(st, view.image_width, view.image_height)
    assert res.aux["R"] == st["num_rendered"] == b["point_list"].shape[0] > 0
    np.testing.assert_array_equal(b["point_list"].astype(np.int64), res.point_list.numpy())
    np.testing.assert_array_equal(b["keys"].astype(np.int64), res.aux["sorted_keys"].numpy())
    np.testing.assert_array_equal(b["ranges"].astype(np.int64), res.ranges.numpy())
    np.testing.assert_array_equal(st["radii"], res.radii.numpy())
    np.testing.assert_array_equal(st["tiles"].astype(np.int64), res.tiles_touched.numpy())


def test_oracle_edge_cases():
    # all Gaussians behind the camera -> background only, R = 0
    sc, view = _tiny(P=32)
    sc.means3D[:, 2] = -1.0
    sc.bg = torch.tensor([0.25, 0.5, 0.75])
    o, res, st = helpers.run_oracle(sc, view, need_grad=True)
    assert res.aux["R"] == 0 and int(res.radii.max()) == 0
    assert torch.allclose(res.color, sc.bg.double()[:, None, None].expand_as(res.color))
    assert float(res.invdepth.abs().max()) == 0.0
    g = o.backward(torch.ones_like(res.color), torch.ones_like(res.invdepth))
    assert all(float(v.abs().max()) == 0.0 for v in g.values() if v is not None)
    # invalid argument combinations raise like the reference's GaussianRasterizer.forward
    with pytest.raises(Exception):
        ro.Oracle(helpers.oracle_settings(sc, view), sc.means3D, sc.opacities, shs=sc.shs,
                  colors_precomp=torch.zeros(sc.P, 3), scales=sc.scales, rotations=sc.rotations)
    with pytest.raises(Exception):
        ro.Oracle(helpers.oracle_settings(sc, view), sc.means3D, sc.opacities, shs=sc.shs)


def test_oracle_mark_visible():
    sc, view = _tiny(P=500)
    vis = ro.mark_visible(sc.means3D, view.world_view_transform)
    assert torch.equal(vis, sc.means3D[:, 2].double() > ro.NEAR_Z)


def test_scene_generator_is_deterministic_and_hits_duplication_target():
    a = scenes.make_config("C1")
    b = scenes.make_config("C1")
    for k in ("means3D", "scales", "rotations", "opacities", "shs"):
        assert torch.equal(getattr(a, k), getattr(b, k))
    st = helpers.bit_structure(a, a.views[0])
    ratio = st["num_rendered"] / a.P
    assert 6.0 < ratio < 10.0, ratio
    c4 = scenes.make_config("C4", P=2000, n_views=4, scale_image=0.1)
    assert len(c4.views) == 4 and c4.views[0].image_width == 160
