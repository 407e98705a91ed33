"""GPU tests of SURVEY 8(f) rank 2: fused activations (raw _opacity / _scaling / _rotation, activations inside the
kernels) and split SH (features_dc / features_rest, no torch.cat), through the drop-in API and the C ABI.

Two references:
  * the float64 oracle fed with float64 torch activations of the same raw parameters
    (/root/reference/scene/gaussian_model.py:40-48, 103-131), gradients chained back through those activations;
  * the un-fused CUDA path (torch activations + torch.cat in fp32 -> GaussianRasterizer), which must agree much more
    tightly because both run the same fp32 kernels downstream.
Tolerances: images <= 1e-4 max-abs, gradients <= 1e-3 relative (BASELINE.json); fused vs un-fused: 2e-6 on images, 1e-4 on gradients (the blend backward adds with atomics in no fixed order).
"""
import pytest
import torch

import helpers
from pdgs_b200 import scenes

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def raw_params(sc, seed=3):
    """Raw optimiser parameters whose activations are the scene's tensors (rotations left un-normalised)."""
    gen = torch.Generator().manual_seed(seed)
    raw = dict(
        means3D=sc.means3D.clone(),
        _opacity=torch.logit(sc.opacities.double().clamp(1e-6, 1 - 1e-6)).float(),
        _scaling=torch.log(sc.scales),
        _rotation=sc.rotations * (0.5 + 1.5 * torch.rand(sc.P, 1, generator=gen)),
        dc=sc.shs[:, :1, :].contiguous(),
        rest=sc.shs[:, 1:, :].contiguous())
    return raw


def activate(raw, dtype):
    op = torch.sigmoid(raw["_opacity"].to(dtype))
    s = torch.exp(raw["_scaling"].to(dtype))
    q = torch.nn.functional.normalize(raw["_rotation"].to(dtype))
    shs = torch.cat((raw["dc"].to(dtype), raw["rest"].to(dtype)), dim=1)
    return op, s, q, shs


def leaves(raw, device):
    lv = {k: v.detach().to(device).clone().requires_grad_(True) for k, v in raw.items()}
    lv["means2D"] = torch.zeros(raw["means3D"].shape[0], 3, device=device, requires_grad=True)
    return lv


def run_fused(sc, view, lv, split=True, raw=True, sh_degree=None):
    from diff_gaussian_rasterization import GaussianRasterizer
    rs = helpers.cuda_settings(sc, view, DEV, sh_degree=sh_degree)
    if raw:
        op, s, q = lv["_opacity"], lv["_scaling"], lv["_rotation"]
    else:
        op, s, q = torch.sigmoid(lv["_opacity"]), torch.exp(lv["_scaling"]), torch.nn.functional.normalize(lv["_rotation"])
    kw = dict(means3D=lv["means3D"], means2D=lv["means2D"], opacities=op, scales=s, rotations=q, raw_params=raw)
    if split:
        kw.update(dc=lv["dc"], shs=lv["rest"])
    else:
        kw.update(shs=torch.cat((lv["dc"], lv["rest"]), dim=1))
    return GaussianRasterizer(rs)(**kw)


def cotangents(color, invd, seed=0):
    gen = torch.Generator().manual_seed(200 + seed)
    gc = torch.randn(color.shape, generator=gen) / color.numel()
    gd = torch.randn(invd.shape, generator=gen) / invd.numel()
    return gc, gd


@pytest.mark.parametrize("split,raw", [(True, True), (True, False), (False, True)])
def test_fused_matches_unfused(split, raw):
    sc = scenes.camera_relative_scene("fz", 21, 6000, 208, 144, 3.0)
    view = sc.views[0]
    rp = raw_params(sc)
    la, lb = leaves(rp, DEV), leaves(rp, DEV)
    ca, ra, da = run_fused(sc, view, la, split=split, raw=raw)
    cb, rb, db = run_fused(sc, view, lb, split=False, raw=False)
    # exp / sigmoid are the same libdevice functions; normalize may differ by an ulp in q -> radii may flip rarely
    assert float((ra != rb).float().mean()) <= 1e-3
    same = bool((ra == rb).all())
    assert float((ca - cb).abs().max()) <= (2e-6 if same else 1e-3)
    assert float((da - db).abs().max()) <= (2e-6 if same else 1e-3)
    gc, gd = cotangents(ca, da)
    torch.autograd.backward([ca, da], [gc.to(DEV), gd.to(DEV)])
    torch.autograd.backward([cb, db], [gc.to(DEV), gd.to(DEV)])
    for k in la:
        assert la[k].grad is not None, k
        e = helpers.rel_err(la[k].grad, lb[k].grad)
        assert e <= 1e-4, (k, e)


@pytest.mark.parametrize("sh_degree", [3, 1])
def test_fused_against_oracle(sh_degree):
    sc = scenes.camera_relative_scene("fo", 22, 3000, 160, 112, 3.0)
    view = sc.views[0]
    rp = raw_params(sc)
    # float64 reference: torch activations -> oracle -> chain rule back through the activations
    r64 = {k: v.double().clone().requires_grad_(True) for k, v in rp.items()}
    op, s, q, shs = activate(r64, torch.float64)
    sc64 = scenes.Scene(name="fo64", means3D=rp["means3D"], opacities=op.detach().float(), scales=s.detach().float(),
                        rotations=q.detach().float(), shs=shs.detach().float(), sh_degree=sc.sh_degree, bg=sc.bg,
                        views=sc.views)
    o, res, st = helpers.run_oracle(sc64, view, sh_degree=sh_degree)
    lv = leaves(rp, DEV)
    color, radii, invd = run_fused(sc, view, lv, split=True, raw=True, sh_degree=sh_degree)
    # the bit-model saw float32-rounded float64 activations; the kernel's fp32 activations can differ by an ulp
    assert float((radii.cpu() != torch.from_numpy(st["radii"])).float().mean()) <= 2e-3
    ok = ~res.fragile
    assert float((color.double().cpu() - res.color).abs().max(dim=0).values[ok].max()) <= 1e-4
    assert float((invd.double().cpu() - res.invdepth).abs()[0][ok].max()) <= 1e-4 * max(1.0, float(res.invdepth.max()))
    gc, gd = cotangents(color, invd, 1)
    ref = o.backward(gc, gd)
    torch.autograd.backward([color, invd], [gc.to(DEV), gd.to(DEV)])
    g_act = [ref["opacities"].view_as(op).double(), ref["scales"].double(), ref["rotations"].double(),
             ref["shs"].view_as(shs).double()]
    g_raw = torch.autograd.grad([op, s, q, shs], [r64["_opacity"], r64["_scaling"], r64["_rotation"], r64["dc"], r64["rest"]],
                                g_act)
    expect = dict(zip(["_opacity", "_scaling", "_rotation", "dc", "rest"], g_raw))
    expect["means3D"] = ref["means3D"]
    expect["means2D"] = ref["means2D"]
    for k, r in expect.items():
        g = lv[k].grad
        assert g is not None, k
        if float(r.abs().max()) == 0.0:
            assert float(g.abs().max()) == 0.0, k
            continue
        e = helpers.rel_err(g, r.view_as(g))
        assert e <= 1e-3, (k, e)


def test_split_degree0_model_and_errors():
    """M == 1 (dc only) and the argument checks of the split path."""
    from diff_gaussian_rasterization import GaussianRasterizer
    sc = scenes.camera_relative_scene("f0", 23, 1500, 96, 80, 3.0)
    view = sc.views[0]
    rs = helpers.cuda_settings(sc, view, DEV, sh_degree=0)
    lv = helpers.cuda_leaves(sc, DEV)
    dc = sc.shs[:, :1, :].contiguous().to(DEV).requires_grad_(True)
    c1, r1, d1 = GaussianRasterizer(rs)(means3D=lv["means3D"], means2D=lv["means2D"], opacities=lv["opacities"],
                                        scales=lv["scales"], rotations=lv["rotations"], dc=dc)
    c2, r2, d2 = GaussianRasterizer(rs)(means3D=lv["means3D"], means2D=lv["means2D"], opacities=lv["opacities"],
                                        scales=lv["scales"], rotations=lv["rotations"], shs=lv["shs"][:, :1, :])
    assert torch.equal(r1, r2) and torch.equal(c1, c2) and torch.equal(d1, d2)
    c1.sum().backward()
    assert dc.grad is not None and float(dc.grad.abs().max()) > 0
    with pytest.raises(Exception):
        GaussianRasterizer(rs)(means3D=lv["means3D"], means2D=lv["means2D"], opacities=lv["opacities"],
                               scales=lv["scales"], rotations=lv["rotations"], dc=dc,
                               colors_precomp=torch.rand(sc.P, 3, device=DEV), shs=lv["shs"][:, 1:, :])


def test_batch_step_split_raw_equals_sum_of_single_view_autograd():
    """The view-parallel batch step fed with features_dc / features_rest and raw parameters accumulates the same
    gradients as per-view autograd through the fused drop-in call."""
    from pdgs_b200.viewparallel import ViewParallelRasterizer
    sc = scenes.make_config("C4", P=6000, n_views=3, scale_image=0.1)
    H, W = sc.views[0].image_height, sc.views[0].image_width
    rp = raw_params(sc)
    gen = torch.Generator().manual_seed(4)
    targets = {i: torch.rand(3, H, W, generator=gen).to(DEV) for i in range(3)}
    sums = None
    for i, view in enumerate(sc.views):
        lv = leaves(rp, DEV)
        color, radii, invd = run_fused(sc, view, lv, split=True, raw=True)
        (color.clamp(0, 1) - targets[i]).abs().mean().backward()   # the loss sees render().clamp(0, 1), as in the reference
        g = {k: lv[k].grad for k in ("means3D", "_opacity", "_scaling", "_rotation", "dc", "rest")}
        sums = g if sums is None else {k: sums[k] + g[k] for k in g}
    vp = ViewParallelRasterizer(sc.P, 16, DEV)
    gaussians = dict(means3D=rp["means3D"].to(DEV), opacities=rp["_opacity"].to(DEV), scales=rp["_scaling"].to(DEV),
                     rotations=rp["_rotation"].to(DEV), shs_dc=rp["dc"].to(DEV), shs_rest=rp["rest"].to(DEV),
                     raw_params=True)
    views = [helpers.cuda_settings(sc, v, DEV) for v in sc.views]
    vp.step(gaussians, views, targets)
    names = dict(means3D="means3D", _opacity="opacities", _scaling="scales", _rotation="rotations", dc="shs_dc", rest="shs_rest")
    for k, seg in names.items():
        e = helpers.rel_err(vp.buf.view(seg), sums[k].view_as(vp.buf.view(seg)))
        assert e < 1e-4, (k, e)


def test_overwrite_mode_with_densification_statistics():
    """accumulate_into={..., "overwrite": True}: every gradient row is written (garbage in the buffers does not matter,
    culled rows become zero) while the densification statistics are still accumulated."""
    from pdgs_b200 import rasterizer as rz
    sc = scenes.camera_relative_scene("ow", 31, 4000, 128, 96, 3.0)
    view = sc.views[0]
    rs = helpers.cuda_settings(sc, view, DEV)
    lv, color, radii, invd = helpers.run_cuda(sc, view, DEV)
    gc, gd = cotangents(color, invd, 2)
    fn = color.grad_fn
    geom, binning, img = fn.saved_tensors[8], fn.saved_tensors[9], fn.saved_tensors[10]  # before backward frees them
    num_rendered = fn.num_rendered
    binning._pdgs_capacity = fn.binning_capacity
    torch.autograd.backward([color, invd], [gc.to(DEV), gd.to(DEV)])
    P = sc.P
    f32 = dict(dtype=torch.float32, device=DEV)
    acc = dict(means3D=torch.full((P, 3), 7.0, **f32), means2D=torch.full((P, 3), 7.0, **f32), shs=torch.full((P, 16, 3), 7.0, **f32),
               opacities=torch.full((P, 1), 7.0, **f32), scales=torch.full((P, 3), 7.0, **f32), rotations=torch.full((P, 4), 7.0, **f32),
               densify_grad_accum=torch.ones(P, **f32), densify_denom=torch.ones(P, **f32),
               max_radii2D=torch.zeros(P, dtype=torch.int32, device=DEV), overwrite=True)
    empty = torch.empty(0, device=DEV)
    d = {k: v.detach() for k, v in lv.items() if v is not None}
    rz.rasterize_gaussians_backward(rs.bg, d["means3D"], radii, empty, d["opacities"], d["scales"], d["rotations"], 1.0, empty,
                                    rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, gc.to(DEV), gd.to(DEV), d["shs"],
                                    rs.sh_degree, rs.campos, geom, num_rendered, binning, img, False, False, accumulate_into=acc)
    for k in ("means3D", "shs", "opacities", "scales", "rotations"):
        assert helpers.rel_err(acc[k], lv[k].grad.view_as(acc[k])) <= 1e-4, k
    vis = radii > 0
    assert float(acc["means3D"][~vis].abs().max()) == 0.0
    expect = torch.ones(P, device=DEV)
    expect[vis] += lv["means2D"].grad[vis, :2].norm(dim=-1)
    assert helpers.rel_err(acc["densify_grad_accum"], expect) <= 1e-5
    assert torch.equal(acc["densify_denom"], 1.0 + vis.float())
    assert torch.equal(acc["max_radii2D"], radii * vis)
