"""GPU tests of the batch (view-parallel) entry point and the fused image loss, through the C ABI."""
import pytest
import torch

import helpers
from pdgs_b200 import scenes

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("W,H,depth", [(64, 48, False), (203, 117, True), (1, 1, True)])
def test_image_loss_matches_torch(W, H, depth):
    from pdgs_b200 import rasterizer as rz
    gen = torch.Generator().manual_seed(W * 7 + H)
    color = torch.rand(3, H, W, generator=gen).to(DEV).requires_grad_(True)
    target = torch.rand(3, H, W, generator=gen).to(DEV)
    target[0, 0, 0] = color[0, 0, 0].detach()  # sign(0) = 0
    invd = torch.rand(1, H, W, generator=gen).to(DEV).requires_grad_(True)
    prior = torch.rand(1, H, W, generator=gen).to(DEV)
    mask = (torch.rand(1, H, W, generator=gen) > 0.4).float().to(DEV)
    w = 0.37
    l1 = (color - target).abs().mean()
    ld = ((invd - prior) * mask).abs().mean()
    (l1 + w * ld).backward() if depth else l1.backward()
    dc, dd, acc = rz.image_loss(color.detach(), target, invd.detach() if depth else None, prior if depth else None,
                                mask if depth else None, 1.0, w)
    assert torch.allclose(dc, color.grad, rtol=0, atol=1e-12)
    assert abs(float(acc[0]) - float(l1)) <= 1e-6
    if depth:
        assert torch.allclose(dd, invd.grad, rtol=1e-6, atol=1e-12)
        assert abs(float(acc[1]) - float(ld)) <= 1e-6
    else:
        assert dd is None and float(acc[1]) == 0.0


def test_batch_step_equals_sum_of_single_view_autograd():
    """Gradients accumulated in-kernel over a batch == sum of per-view autograd gradients; densification
    statistics == the reference's torch ops (scene/gaussian_model.py:502-506, train.py:313-314)."""
    from pdgs_b200.viewparallel import ViewParallelRasterizer
    sc = scenes.make_config("C4", P=8000, n_views=4, scale_image=0.12)
    H, W = sc.views[0].image_height, sc.views[0].image_width
    gen = torch.Generator().manual_seed(3)
    targets = {i: torch.rand(3, H, W, generator=gen).to(DEV) for i in range(4)}
    priors = {i: torch.rand(1, H, W, generator=gen).to(DEV) for i in range(4)}
    masks = {i: (torch.rand(1, H, W, generator=gen) > 0.5).float().to(DEV) for i in range(4)}
    w_depth = 0.8
    # reference: four autograd passes
    sums, accum, denom, maxr, loss = None, torch.zeros(sc.P, device=DEV), torch.zeros(sc.P, device=DEV), \
        torch.zeros(sc.P, dtype=torch.int32, device=DEV), 0.0
    for i, view in enumerate(sc.views):
        lv, color, radii, invd = helpers.run_cuda(sc, view, DEV)
        image = color.clamp(0, 1)  # what render() returns (/root/reference/gaussian_renderer/__init__.py:121)
        l = (image - targets[i]).abs().mean() + w_depth * ((invd - priors[i]) * masks[i]).abs().mean()
        l.backward()
        loss += float((image - targets[i]).abs().mean())
        g = {k: lv[k].grad for k in ("means3D", "shs", "opacities", "scales", "rotations")}
        sums = g if sums is None else {k: sums[k] + g[k] for k in g}
        vis = radii > 0
        accum[vis] += lv["means2D"].grad[vis, :2].norm(dim=-1)
        denom[vis] += 1
        maxr[vis] = torch.max(maxr[vis], radii[vis])
    vp = ViewParallelRasterizer(sc.P, 16, DEV)
    gaussians = {k: getattr(sc, k).to(DEV) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    views = [helpers.cuda_settings(sc, v, DEV) for v in sc.views]
    out = vp.step(gaussians, views, targets, priors, masks, w_depth)
    assert out["local_views"] == [0, 1, 2, 3] and vp.last_launches > 0
    for k in sums:
        assert helpers.rel_err(vp.buf.view(k), sums[k].view_as(vp.buf.view(k))) < 1e-4, k
    assert helpers.rel_err(vp.buf.view("densify_grad_accum"), accum) < 1e-4
    assert torch.equal(vp.buf.view("densify_denom"), denom)
    assert torch.equal(vp.buf.max_radii2D, maxr)
    assert abs(float(vp.loss_acc[0]) - loss) < 1e-5
    # a second step starts from zero again
    vp.step(gaussians, views, targets, priors, masks, w_depth)
    assert helpers.rel_err(vp.buf.view("shs"), sums["shs"]) < 1e-4


def test_stage_timing_accumulates_without_sync():
    from pdgs_b200 import rasterizer as rz
    sc = scenes.camera_relative_scene("t", 8, 5000, 160, 120, 3.0)
    rz.set_stage_timing(DEV, True)
    for _ in range(3):
        lv, color, radii, invd = helpers.run_cuda(sc, sc.views[0], DEV)
        color.sum().backward()
    st = rz.stage_times(DEV)
    rz.set_stage_timing(DEV, False)
    assert st["preprocess"][1] == 3 and st["blend_bwd"][1] == 3 and st["sort"][1] >= 3
    assert all(ms >= 0.0 for ms, _ in st.values()) and st["blend_fwd"][0] > 0.0
    assert rz.stage_times(DEV)["preprocess"][1] == 0


def test_batch_step_with_ssim_training_loss_equals_autograd_through_the_dropin_packages():
    """The reference's training step (train.py:143-184): render -> (1-l)*L1 + l*(1-fused_ssim) + w*depth L1 ->
    backward, written with the two drop-in packages and torch autograd, equals the fused batch step."""
    from fused_ssim import fused_ssim
    from pdgs_b200.viewparallel import ViewParallelRasterizer
    sc = scenes.make_config("C4", P=6000, n_views=3, scale_image=0.1)
    H, W = sc.views[0].image_height, sc.views[0].image_width
    gen = torch.Generator().manual_seed(5)
    targets = {i: torch.rand(3, H, W, generator=gen).to(DEV) for i in range(3)}
    priors = {i: torch.rand(1, H, W, generator=gen).to(DEV) for i in range(3)}
    alphas = {i: (torch.rand(1, H, W, generator=gen) > 0.2).float().to(DEV) for i in range(3)}
    lam, w_depth = 0.2, 0.5
    sums, ssim_sum = None, 0.0
    over = 0
    for i, view in enumerate(sc.views):
        lv, color, radii, invd = helpers.run_cuda(sc, view, DEV)
        over += int((color > 1).sum())
        image = color.clamp(0, 1)    # /root/reference/gaussian_renderer/__init__.py:121
        image = image * alphas[i]    # /root/reference/train.py:148-150
        s = fused_ssim(image.unsqueeze(0), targets[i].unsqueeze(0))
        l = (1.0 - lam) * (image - targets[i]).abs().mean() + lam * (1.0 - s) + w_depth * (invd - priors[i]).abs().mean()
        l.backward()
        ssim_sum += float(s)
        g = {k: lv[k].grad for k in ("means3D", "shs", "opacities", "scales", "rotations")}
        sums = g if sums is None else {k: sums[k] + g[k] for k in g}
    vp = ViewParallelRasterizer(sc.P, 16, DEV)
    gaussians = {k: getattr(sc, k).to(DEV) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    views = [helpers.cuda_settings(sc, v, DEV) for v in sc.views]
    assert over > 0, "the scene must saturate some pixels, or the clamp is not exercised"
    vp.step(gaussians, views, targets, priors, None, w_depth, lambda_dssim=lam, alpha_masks=alphas)
    for k in sums:
        assert helpers.rel_err(vp.buf.view(k), sums[k].view_as(vp.buf.view(k))) < 1e-4, k
    assert abs(float(vp.loss_acc[2]) - ssim_sum) < 1e-4


def test_step_from_staged_host_gaussians():
    """step(gaussians=stage_gaussians(host)) -- parameters uploaded on the side stream, images from pinned host memory --
    gives the same gradients as the all-resident step."""
    from pdgs_b200.viewparallel import ViewParallelRasterizer
    sc = scenes.make_config("C4", P=5000, n_views=3, scale_image=0.1)
    H, W = sc.views[0].image_height, sc.views[0].image_width
    gen = torch.Generator().manual_seed(9)
    tg_host = {i: torch.rand(3, H, W, generator=gen).pin_memory() for i in range(3)}
    host = {k: getattr(sc, k).contiguous().pin_memory() for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    views = [helpers.cuda_settings(sc, v, DEV) for v in sc.views]
    a = ViewParallelRasterizer(sc.P, 16, DEV)
    a.step({k: v.to(DEV) for k, v in host.items()}, views, {i: t.to(DEV) for i, t in tg_host.items()})
    b = ViewParallelRasterizer(sc.P, 16, DEV)
    handle = b.stage_gaussians(host)
    nxt = b.stage_gaussians(host)  # a second upload in flight, as the pipelined loop does
    # next step's parameters uploaded slice by slice, and the first image(s) of the next step behind the last view
    out = b.step(handle, views, tg_host, prefetch_gaussians=host, prefetch_images=dict(targets=tg_host))
    assert len(b._pre_images) >= 1
    b.step(nxt, views, tg_host)
    assert len(b._pre_images) == 0   # consumed by the step that followed
    b.step(out["prefetched"], views, tg_host)
    torch.cuda.synchronize()
    assert helpers.rel_err(b.buf.flat, a.buf.flat) < 1e-4
    assert torch.equal(b.buf.max_radii2D, a.buf.max_radii2D)


@pytest.mark.parametrize("split_raw", [False, True])
def test_lanes_match_one_view_at_a_time(split_raw):
    """Three views in flight on separate streams / host threads, all adding into the same buffers with reductions,
    give the gradients, statistics and losses of the one-view-at-a-time step (up to the order of the fp32 sums)."""
    from pdgs_b200.viewparallel import ViewParallelRasterizer
    sc = scenes.make_config("C4", P=9000, n_views=7, scale_image=0.1)
    H, W = sc.views[0].image_height, sc.views[0].image_width
    gen = torch.Generator().manual_seed(11)
    targets = {i: torch.rand(3, H, W, generator=gen).pin_memory() for i in range(7)}   # host images: upload path too
    priors = {i: torch.rand(1, H, W, generator=gen).to(DEV) for i in range(7)}
    g = {k: getattr(sc, k).to(DEV) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    if split_raw:
        g = dict(means3D=g["means3D"], opacities=torch.logit(g["opacities"].clamp(1e-6, 1 - 1e-6)), scales=torch.log(g["scales"]),
                 rotations=g["rotations"] * 1.3, shs_dc=g["shs"][:, :1].contiguous(), shs_rest=g["shs"][:, 1:].contiguous(),
                 raw_params=True)
    views = [helpers.cuda_settings(sc, v, DEV) for v in sc.views]
    a = ViewParallelRasterizer(sc.P, 16, DEV, lanes=1)
    b = ViewParallelRasterizer(sc.P, 16, DEV, lanes=3)
    for lam in (0.0, 0.2):
        a.step(g, views, targets, priors, None, 0.5, lambda_dssim=lam)
        out = b.step(g, views, targets, priors, None, 0.5, lambda_dssim=lam)
        torch.cuda.synchronize()
        assert out["local_views"] == list(range(7)) and len(b.last_num_rendered) == 7
        assert b.last_num_rendered == a.last_num_rendered
        assert helpers.rel_err(b.buf.flat, a.buf.flat) < 1e-4
        assert torch.equal(b.buf.max_radii2D, a.buf.max_radii2D)
        assert torch.allclose(b.loss_acc, a.loss_acc, rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("split", [False, True])
def test_deferred_sh_backward_and_accumulating_k9_against_oracle(split):
    """Direct oracle comparison of the batch path's two special kernels (no CUDA-vs-CUDA step in between): K9 in
    accumulate mode (adds every view's gradients into the same buffers, leaves only the clamp-masked dL/drgb per view)
    and pdgs_sh_backward_views(_split) (SH gradient + view-direction term of all views in one pass), against the SUM
    over views of the float64 oracle's autograd gradients for random cotangents."""
    from pdgs_b200 import rasterizer as rz
    from pdgs_b200.viewparallel import GradBuffer
    n_views = 4
    sc = scenes.make_config("C4", P=6000, n_views=n_views, scale_image=0.12)
    H, W = sc.views[0].image_height, sc.views[0].image_width
    g = {k: getattr(sc, k).to(DEV) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    dc, rest = g["shs"][:, :1].contiguous(), g["shs"][:, 1:].contiguous()
    buf = GradBuffer(sc.P, 16, DEV)
    acc = buf.accumulate_dict(split)
    color_grads = torch.empty(n_views, sc.P, 3, device=DEV)
    empty = torch.empty(0, device=DEV)
    want = None
    accum_ref, denom_ref = torch.zeros(sc.P, dtype=torch.float64), torch.zeros(sc.P)
    for i, view in enumerate(sc.views):
        o, res, st = helpers.run_oracle(sc, view)
        gen = torch.Generator().manual_seed(40 + i)
        gc = torch.randn(3, H, W, generator=gen) / (3 * H * W)
        gd = torch.randn(1, H, W, generator=gen) / (H * W)
        ref = o.backward(gc, gd)
        want = ref if want is None else {k: (want[k] + ref[k] if ref[k] is not None else None) for k in ref}
        vis = torch.from_numpy(st["radii"] > 0)
        accum_ref += ref["means2D"][:, :2].norm(dim=-1) * vis
        denom_ref += vis.float()
        rs = helpers.cuda_settings(sc, view, DEV)
        R, color, radii, geom, binning, img, invd = rz.rasterize_gaussians(
            rs.bg, g["means3D"], empty, g["opacities"], g["scales"], g["rotations"], 1.0, empty, rs.viewmatrix, rs.projmatrix,
            rs.tanfovx, rs.tanfovy, H, W, rest if split else g["shs"], 3, rs.campos, False, False, False,
            dc=dc if split else None)
        np_radii = radii.cpu().numpy()
        assert (np_radii == st["radii"]).all()
        a = dict(acc)
        a["color_grad_view"] = color_grads[i]
        rz.rasterize_gaussians_backward(
            rs.bg, g["means3D"], radii, empty, g["opacities"], g["scales"], g["rotations"], 1.0, empty, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, gc.to(DEV), gd.to(DEV), rest if split else g["shs"], 3, rs.campos, geom, R,
            binning, img, False, False, accumulate_into=a, dc=dc if split else None)
    campos = torch.stack([v.camera_center for v in sc.views]).to(DEV)
    if split:
        rz.sh_backward_views_split(g["means3D"], dc, rest, 3, campos, color_grads, acc["shs_dc"], acc["shs_rest"],
                                   acc["means3D"], accumulate=False)
    else:
        rz.sh_backward_views(g["means3D"], g["shs"], 3, campos, color_grads, acc["shs"], acc["means3D"], accumulate=False)
    torch.cuda.synchronize()
    got_shs = torch.cat([buf.view("shs_dc"), buf.view("shs_rest")], dim=1) if split else buf.view("shs")
    for k, got in (("means3D", buf.view("means3D")), ("shs", got_shs), ("opacities", buf.view("opacities")),
                   ("scales", buf.view("scales")), ("rotations", buf.view("rotations"))):
        e = helpers.rel_err(got, want[k].reshape(got.shape))
        assert e <= 1e-3, (k, e)
        # element-wise: no entry is off by more than 0.2 % of the largest one (the sum over views cancels in some rows,
        # so an error relative to the entry itself is not meaningful there)
        w = want[k].reshape(got.shape)
        assert float((got.double().cpu() - w).abs().max()) <= 2e-3 * float(w.abs().max()), k
    assert helpers.rel_err(buf.view("densify_grad_accum"), accum_ref) <= 1e-3
    assert torch.equal(buf.view("densify_denom").cpu(), denom_ref)


@pytest.mark.parametrize("split", [False, True])
def test_sh_backward_views_in_chunks_along_p_equals_one_pass(split):
    """pdgs_sh_backward_views(_split) over [first, first + count) ranges -- what the multi-GPU step runs so that the
    all-reduce of a chunk's rows overlaps the next chunk -- writes bit-identical rows to the single whole-P pass."""
    from pdgs_b200 import rasterizer as rz
    P, V = 2600, 5
    gen = torch.Generator().manual_seed(12)
    means = torch.randn(P, 3, generator=gen).to(DEV)
    shs = torch.randn(P, 16, 3, generator=gen).to(DEV)
    dc, rest = shs[:, :1].contiguous(), shs[:, 1:].contiguous()
    campos = (3.0 * torch.randn(V, 3, generator=gen)).to(DEV)
    cg = torch.randn(V, P, 3, generator=gen).to(DEV)
    cg[:, ::7] = 0.0   # culled rows

    def run(chunks):
        g_sh = torch.full((P, 16, 3), 7.0, device=DEV)
        g_dc, g_rest = torch.full((P, 1, 3), 7.0, device=DEV), torch.full((P, 15, 3), 7.0, device=DEV)
        g_m = torch.ones(P, 3, device=DEV)
        for lo, hi in chunks:
            if split:
                rz.sh_backward_views_split(means, dc, rest, 3, campos, cg, g_dc, g_rest, g_m, accumulate=False, first=lo,
                                           count=hi - lo)
            else:
                rz.sh_backward_views(means, shs, 3, campos, cg, g_sh, g_m, accumulate=False, first=lo, count=hi - lo)
        return (torch.cat([g_dc, g_rest], dim=1) if split else g_sh), g_m

    whole = run([(0, 0)])
    parts = run([(0, 1024), (1024, 2304), (2304, P)])
    assert torch.equal(whole[0], parts[0]) and torch.equal(whole[1], parts[1])
    assert float((whole[0] - 7.0).abs().max()) > 0.0
    only = run([(1024, 2304)])   # rows outside the range are untouched
    assert bool((only[0][:1024] == 7.0).all()) and bool((only[0][2304:] == 7.0).all()) and bool((only[1][:1024] == 1.0).all())
    assert torch.equal(only[0][1024:2304], whole[0][1024:2304])
    with pytest.raises(Exception, match="multiple"):
        run([(10, 300)])
