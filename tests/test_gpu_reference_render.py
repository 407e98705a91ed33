"""The reference's OWN caller code on top of the drop-in package (north_star: "so
gaussian_renderer/render(), train.py and the depth-prior loss run unchanged").

What runs here is the reference's text, not a restatement: oracle/build_ref.py stages byte copies of
  gaussian_renderer/__init__.py   render()                      [REF gaussian_renderer/__init__.py:18-130]
  scene/gaussian_model.py         GaussianModel                 [REF scene/gaussian_model.py:30-131, 180-214, 502-506]
  scene/cameras.py                MiniCam                       [REF scene/cameras.py:91-102]
  utils/loss_utils.py             l1_loss, ssim                 [REF utils/loss_utils.py:40-89]
  arguments/__init__.py           PipelineParams, OptimizationParams
under oracle/_ref/ref_py (git-ignored; /root/reference does not exist on the GPU box), and this test imports them with
`diff_gaussian_rasterization` / `fused_ssim` resolving to THIS repo's packages.  The training step below follows
[REF train.py:143-184, 308-338] line by line (render -> clamp'ed image -> l1 + fused_ssim -> depth-prior L1 ->
backward -> max_radii2D / add_densification_stats -> optimizer.step).

Checked against the float64 oracle + fp32 bit-model: image, inverse depth, radii, visibility filter, the gradients of
the six raw parameter tensors (through the reference's activations), viewspace_points.grad, the densification
statistics, and -- with SparseGaussianAdam -- the parameters after `optimizer.step(visible, N)` bit-exactly against
oracle/adam_oracle.py.  Both import behaviours are covered: SparseGaussianAdam exported (separate_sh=True, the `dc=`
call [REF gaussian_renderer/__init__.py:92-101]) and hidden (separate_sh=False, [REF :103-112]).
"""
from argparse import ArgumentParser

import numpy as np
import pytest
import torch
from torch import nn

import helpers
import ref_loader
from oracle import ssim_oracle as so
from oracle.adam_oracle import sparse_adam_step
from pdgs_b200 import scenes

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_loader.have_ref_py(),
                                 reason="oracle/_ref/ref_py missing: run `python oracle/build_ref.py` where /root/reference exists")]
DEV = "cuda:0"


def _build_model(GaussianModel, sc, optimizer_type):
    """A GaussianModel holding a synthetic scene: the raw tensors create_from_pcd / load_ply would have produced
    [REF scene/gaussian_model.py:150-178]."""
    pc = GaussianModel(3, optimizer_type)
    pc._xyz = nn.Parameter(sc.means3D.to(DEV).requires_grad_(True))
    pc._features_dc = nn.Parameter(sc.shs[:, :1].contiguous().to(DEV).requires_grad_(True))
    pc._features_rest = nn.Parameter(sc.shs[:, 1:].contiguous().to(DEV).requires_grad_(True))
    pc._scaling = nn.Parameter(torch.log(sc.scales).to(DEV).requires_grad_(True))
    pc._rotation = nn.Parameter((sc.rotations * 1.3).to(DEV).requires_grad_(True))   # not unit: normalize() matters
    pc._opacity = nn.Parameter(torch.logit(sc.opacities.clamp(1e-6, 1 - 1e-6)).to(DEV).requires_grad_(True))
    pc.max_radii2D = torch.zeros(sc.P, device=DEV)
    pc.active_sh_degree = 3
    pc.spatial_lr_scale = 1.0
    pc._exposure = nn.Parameter(torch.eye(3, 4, device=DEV)[None])
    pc.pretrained_exposures = None
    return pc


def _oracle_scene(pc, sc):
    """The post-activation tensors render() hands to the rasterizer, as computed by the reference's getters on the
    GPU (so the fp32 bit-model sees the very same fp32 inputs)."""
    with torch.no_grad():
        return scenes.Scene("ref", pc.get_xyz.cpu(), pc.get_scaling.cpu(), pc.get_rotation.cpu(), pc.get_opacity.cpu(),
                            pc.get_features.cpu().contiguous(), sc.bg, sc.views)


def _chain_to_raw(pc, ref):
    """Oracle gradients w.r.t. the rasterizer inputs -> gradients w.r.t. the model's raw tensors through the
    reference's activations [REF scene/gaussian_model.py:40-48, 103-131], in float64."""
    raw_s = pc._scaling.detach().double().cpu().requires_grad_(True)
    raw_r = pc._rotation.detach().double().cpu().requires_grad_(True)
    raw_o = pc._opacity.detach().double().cpu().requires_grad_(True)
    outs = [torch.exp(raw_s), torch.nn.functional.normalize(raw_r), torch.sigmoid(raw_o)]
    cots = [ref["scales"], ref["rotations"], ref["opacities"].view_as(outs[2])]
    torch.autograd.backward(outs, cots)
    return dict(_xyz=ref["means3D"], _features_dc=ref["shs"][:, :1], _features_rest=ref["shs"][:, 1:], _scaling=raw_s.grad,
                _rotation=raw_r.grad, _opacity=raw_o.grad)


@pytest.mark.parametrize("sparse_adam,antialiasing", [(False, False), (True, False), (True, True)])
def test_reference_training_step_on_the_dropin(sparse_adam, antialiasing):
    with ref_loader.ref_py(sparse_adam):
        import diff_gaussian_rasterization as dgr
        from arguments import OptimizationParams, PipelineParams
        from fused_ssim import fused_ssim
        from gaussian_renderer import render
        from scene.cameras import MiniCam
        from scene.gaussian_model import GaussianModel
        from utils.loss_utils import l1_loss

        # what [REF train.py:47-52] computes
        SPARSE_ADAM_AVAILABLE = hasattr(dgr, "SparseGaussianAdam")
        assert SPARSE_ADAM_AVAILABLE == sparse_adam
        opt = OptimizationParams(ArgumentParser())
        pipe = PipelineParams(ArgumentParser())
        pipe.antialiasing = antialiasing
        opt.optimizer_type = "sparse_adam" if sparse_adam else "default"

        sc = scenes.make_config("C4", P=5000, n_views=3, scale_image=0.125)   # 200 x 133, world-space cameras
        v = sc.views[1]
        H, W = v.image_height, v.image_width
        cam = MiniCam(W, H, v.FoVy, v.FoVx, scenes.ZNEAR, scenes.ZFAR, v.world_view_transform.to(DEV),
                      v.full_proj_transform.to(DEV))
        gaussians = _build_model(GaussianModel, sc, opt.optimizer_type)
        gaussians.training_setup(opt)
        use_sparse_adam = opt.optimizer_type == "sparse_adam" and SPARSE_ADAM_AVAILABLE   # [REF train.py:82]
        assert type(gaussians.optimizer).__name__ == ("SparseGaussianAdam" if sparse_adam else "Adam")
        bg = sc.bg.to(DEV)
        gen = torch.Generator().manual_seed(31)
        gt_image = torch.rand(3, H, W, generator=gen).to(DEV)
        mono_invdepth = torch.rand(1, H, W, generator=gen).to(DEV)
        depth_mask = (torch.rand(1, H, W, generator=gen) > 0.3).float().to(DEV)
        depth_w = 0.7
        before = {k: getattr(gaussians, k).detach().clone() for k in
                  ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation")}

        # ---- [REF train.py:143-184] ----
        render_pkg = render(cam, gaussians, pipe, bg, use_trained_exp=False, separate_sh=SPARSE_ADAM_AVAILABLE)
        image, viewspace_point_tensor, visibility_filter, radii = (render_pkg["render"], render_pkg["viewspace_points"],
                                                                   render_pkg["visibility_filter"], render_pkg["radii"])
        Ll1 = l1_loss(image, gt_image)
        ssim_value = fused_ssim(image.unsqueeze(0), gt_image.unsqueeze(0))
        loss = (1.0 - opt.lambda_dssim) * Ll1 + opt.lambda_dssim * (1.0 - ssim_value)
        invDepth = render_pkg["depth"]
        Ll1depth_pure = torch.abs((invDepth - mono_invdepth) * depth_mask).mean()
        loss = loss + depth_w * Ll1depth_pure
        loss.backward()

        # ---- oracle on the same inputs ----
        osc = _oracle_scene(gaussians, sc)
        o, res, st = helpers.run_oracle(osc, v, antialiasing=antialiasing)
        np.testing.assert_array_equal(radii.cpu().numpy(), st["radii"])
        assert torch.equal(visibility_filter.cpu().view(-1), torch.from_numpy(np.nonzero(st["radii"] > 0)[0]))
        ok = ~res.fragile
        assert float(res.fragile.float().mean()) < 0.01
        assert float((image.double().cpu() - res.color.clamp(0, 1)).abs().max(dim=0).values[ok].max()) <= 1e-4
        assert float((invDepth.double().cpu() - res.invdepth).abs()[0][ok].max()) <= 1e-4 * max(1.0, float(res.invdepth.max()))
        assert int((res.color > 1).sum()) > 0, "no saturated pixel: the caller-side clamp is not exercised"
        xc = res.color.clone().requires_grad_(True)
        xd = res.invdepth.clone().requires_grad_(True)
        loss_ref, parts = so.training_loss(xc.clamp(0, 1), gt_image.double().cpu(), opt.lambda_dssim, xd,
                                           mono_invdepth.double().cpu(), depth_mask.double().cpu(), depth_w)
        assert abs(float(loss) - float(loss_ref)) <= 2e-5
        gc, gd = torch.autograd.grad(loss_ref, [xc, xd])
        ref = o.backward(gc, gd)
        want = _chain_to_raw(gaussians, ref)
        errs = {}
        for k, r in want.items():
            g = getattr(gaussians, k).grad
            assert g is not None and g.shape == getattr(gaussians, k).shape, k
            errs[k] = helpers.rel_err(g, r.reshape(g.shape))
            assert errs[k] <= 5e-3, (k, errs[k])   # sign() of an fp32 residual ~ 0 may differ from float64's
        g2d = viewspace_point_tensor.grad
        assert g2d is not None and g2d.shape == (sc.P, 3) and float(g2d[:, 2].abs().max()) == 0.0
        assert helpers.rel_err(g2d, ref["means2D"]) <= 5e-3
        print("reference training step on the drop-in: grad rel errs", errs)

        # ---- [REF train.py:308-338] ----
        with torch.no_grad():
            gaussians.max_radii2D[visibility_filter] = torch.max(gaussians.max_radii2D[visibility_filter],
                                                                 radii[visibility_filter])
            gaussians.add_densification_stats(viewspace_point_tensor, visibility_filter)
            vis = torch.from_numpy(st["radii"] > 0)
            assert torch.equal(gaussians.max_radii2D.cpu(), torch.from_numpy(st["radii"]).float())
            assert torch.equal(gaussians.denom.cpu().view(-1), vis.float())
            assert helpers.rel_err(gaussians.xyz_gradient_accum.view(-1), ref["means2D"][:, :2].norm(dim=-1) * vis) <= 5e-3
            grads = {k: getattr(gaussians, k).grad.detach().clone() for k in before}
            gaussians.exposure_optimizer.step()
            gaussians.exposure_optimizer.zero_grad(set_to_none=True)
            if use_sparse_adam:
                visible = radii > 0
                gaussians.optimizer.step(visible, radii.shape[0])
                gaussians.optimizer.zero_grad(set_to_none=True)
            else:
                gaussians.optimizer.step()
                gaussians.optimizer.zero_grad(set_to_none=True)
        lrs = {g["name"]: g["lr"] for g in gaussians.optimizer.param_groups}
        names = dict(_xyz="xyz", _features_dc="f_dc", _features_rest="f_rest", _opacity="opacity", _scaling="scaling",
                     _rotation="rotation")
        for k, p0 in before.items():
            p1 = getattr(gaussians, k).detach()
            assert torch.isfinite(p1).all(), k
            if use_sparse_adam:
                exp = p0.cpu().numpy().copy()
                sparse_adam_step(exp, grads[k].cpu().numpy(), np.zeros_like(exp), np.zeros_like(exp),
                                 (st["radii"] > 0), lrs[names[k]], eps=1e-15, step=0)
                np.testing.assert_array_equal(p1.cpu().numpy().view(np.uint32), exp.view(np.uint32), err_msg=k)
                assert torch.equal(p1[~visible], p0[~visible]), k   # invisible rows untouched
            else:
                assert not torch.equal(p1, p0), k


def test_reference_render_python_side_paths():
    """The other branches of render(): SH evaluated in Python (`convert_SHs_python`, [REF :72-80] -> colors_precomp),
    covariance built in Python (`compute_cov3D_python`, [REF :63-64] -> cov3D_precomp), `override_color`, and
    `scaling_modifier`; forward images agree with the in-kernel SH / covariance path."""
    with ref_loader.ref_py(False):
        from arguments import PipelineParams
        from gaussian_renderer import render
        from scene.cameras import MiniCam
        from scene.gaussian_model import GaussianModel
        sc = scenes.make_config("C4", P=4000, n_views=2, scale_image=0.1)
        v = sc.views[0]
        cam = MiniCam(v.image_width, v.image_height, v.FoVy, v.FoVx, scenes.ZNEAR, scenes.ZFAR,
                      v.world_view_transform.to(DEV), v.full_proj_transform.to(DEV))
        pc = _build_model(GaussianModel, sc, "default")
        bg = sc.bg.to(DEV)
        pipe = PipelineParams(ArgumentParser())
        with torch.no_grad():
            base = render(cam, pc, pipe, bg)
            pipe.convert_SHs_python = True
            py_sh = render(cam, pc, pipe, bg)
            pipe.convert_SHs_python, pipe.compute_cov3D_python = False, True
            py_cov = render(cam, pc, pipe, bg)
            pipe.compute_cov3D_python = False
            half = render(cam, pc, pipe, bg, scaling_modifier=0.5)
            over = render(cam, pc, pipe, bg, override_color=torch.full((sc.P, 3), 0.25, device=DEV))
        assert torch.equal(base["radii"], py_sh["radii"])
        assert float((base["render"] - py_sh["render"]).abs().max()) <= 1e-4
        assert float((base["depth"] - py_sh["depth"]).abs().max()) == 0.0
        assert float((base["radii"] != py_cov["radii"]).float().mean()) <= 1e-3
        assert float((base["render"] - py_cov["render"]).abs().max()) <= 2e-3
        assert int(half["radii"].max()) < int(base["radii"].max())
        # a constant colour blends to 0.25 * (1 - T) + T * bg
        osc = _oracle_scene(pc, sc)
        o, res, st = helpers.run_oracle(osc, v, need_grad=False)
        expect = 0.25 * (1 - res.final_T)[None] + res.final_T[None] * sc.bg.double()[:, None, None]
        ok = ~res.fragile
        assert float((over["render"].double().cpu() - expect.clamp(0, 1)).abs().max(dim=0).values[ok].max()) <= 1e-4
