"""Host side of the rasterizer above the C ABI: the role of the reference's rasterize_points.cu
glue (`RasterizeGaussiansCUDA`, `RasterizeGaussiansBackwardCUDA`, `markVisible`; source absent
from /root/reference, interface per SURVEY.md 8(a) rows a5/a6 and Appendix B, caller at
/root/reference/gaussian_renderer/__init__.py:104-112).

PyTorch is used for device memory and the current stream only; all arithmetic happens in
libpdgs.so.  Argument order and return tuples of `rasterize_gaussians` /
`rasterize_gaussians_backward` / `mark_visible` follow the reference module's `_C` entry points.
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import Optional

import torch

from . import _lib as L

# Speculative capacity of the binning buffer (pairs) per (device, image width, image height): a HIGH-WATER MARK of the
# pair counts seen, decaying by 1 % per call, so that a training loop that pops cameras in random order
# (/root/reference/train.py:130-135: R varies from view to view) does not miss every time R grows.  A miss costs a
# re-allocation and a second pass of duplicate + sort + blend; misses are counted (binning_stats()).
_capacity_hint = {}
_CAP_HEADROOM, _CAP_DECAY = 1.25, 0.99
_binning_stats = {"calls": 0, "misses": 0, "sizing_calls": 0}
_binning_lock = threading.Lock()   # the view-parallel lanes call in from several host threads


def _count(key: str) -> None:
    with _binning_lock:
        _binning_stats[key] += 1


def binning_stats(reset: bool = False) -> dict:
    """{"calls", "misses", "sizing_calls"}: forwards run, forwards whose speculative binning capacity was too small
    (duplicate + sort + blend ran twice), forwards that had no capacity hint yet (scan only, then the real pass)."""
    with _binning_lock:
        out = dict(_binning_stats)
        if reset:
            for k in _binning_stats:
                _binning_stats[k] = 0
    return out


class _Launches(threading.local):
    """Kernels of ours launched by the calling thread's last forward / backward / loss / SH-backward call (for
    bench.py's gpu_launches).  Thread-local: the view-parallel lanes call in from several host threads."""
    forward = 0
    backward = 0
    sh_bwd = 0
    loss = 0


LAUNCHES = _Launches()


def _ptr(t: Optional[torch.Tensor]):
    if t is None or t.numel() == 0:
        return None
    return C.c_void_p(t.data_ptr())


def _f32c(t: Optional[torch.Tensor], name: str, device) -> Optional[torch.Tensor]:
    if t is None or t.numel() == 0:
        return None
    if not t.is_cuda:
        raise L.PdgsError(f"{name} must be a CUDA tensor (this rasterizer has no CPU path)")
    if t.device != device:
        raise L.PdgsError(f"{name} is on {t.device}, expected {device}")
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _passes(tiles: int) -> int:
    bits = 0
    while (1 << bits) <= tiles and bits < 31:
        bits += 1
    return (32 + bits + 7) // 8


def _camera(bg, scale_modifier, viewmatrix, projmatrix, tanfovx, tanfovy, H, W, sh_degree, campos, prefiltered,
            antialiasing, debug) -> L.PdgsCamera:
    cam = L.PdgsCamera()
    cam.image_height, cam.image_width = int(H), int(W)
    cam.tanfovx, cam.tanfovy = float(tanfovx), float(tanfovy)
    cam.bg = _ptr(bg)
    cam.scale_modifier = float(scale_modifier)
    cam.viewmatrix, cam.projmatrix = _ptr(viewmatrix), _ptr(projmatrix)
    cam.sh_degree = int(sh_degree)
    cam.campos = _ptr(campos)
    cam.prefiltered, cam.debug, cam.antialiasing = int(bool(prefiltered)), int(bool(debug)), int(bool(antialiasing))
    return cam


def _scene(P, means3D, sh, colors_precomp, opacities, scales, rotations, cov3D_precomp, dc=None,
           raw_params=False) -> L.PdgsScene:
    """dc given: split SH -- `dc` is features_dc [P,1,3] and `sh` features_rest [P,M-1,3] (may be None/empty)."""
    sc = L.PdgsScene()
    sc.P = int(P)
    if dc is not None:
        sc.M = 1 + (int(sh.shape[1]) if sh is not None else 0)
        sc.shs_dc, sc.shs_rest, sc.shs = _ptr(dc), _ptr(sh), None
    else:
        sc.M = int(sh.shape[1]) if sh is not None else 0
        sc.shs = _ptr(sh)
    sc.raw_params = int(bool(raw_params))
    sc.means3D, sc.colors_precomp = _ptr(means3D), _ptr(colors_precomp)
    sc.opacities, sc.scales, sc.rotations = _ptr(opacities), _ptr(scales), _ptr(rotations)
    sc.cov3D_precomp = _ptr(cov3D_precomp)
    return sc


def _state(geom, image, binning, capacity) -> L.PdgsState:
    st = L.PdgsState()
    st.geom, st.geom_bytes = _ptr(geom), geom.numel()
    st.image, st.image_bytes = _ptr(image), image.numel()
    st.binning = _ptr(binning)
    st.binning_bytes = 0 if binning is None else binning.numel()
    st.capacity_pairs = int(capacity)
    return st


def _binning_capacity(binning: Optional[torch.Tensor]) -> int:
    """The capacity a binning blob was allocated for is stored in its python attribute."""
    return int(getattr(binning, "_pdgs_capacity", 0)) if binning is not None else 0


def rasterize_gaussians(bg, means3D, colors_precomp, opacities, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, tanfovx, tanfovy, image_height, image_width, sh, sh_degree, campos,
                        prefiltered, antialiasing, debug, dc=None, raw_params=False):
    """-> (num_rendered, color[3,H,W], radii[P] int32, geomBuffer, binningBuffer, imgBuffer, invdepths[1,H,W]).

    Extensions beyond the reference's 20 arguments: `dc` = features_dc [P,1,3], in which case
    `sh` is features_rest [P,M-1,3] (the pair render() passes when separate_sh is set,
    /root/reference/gaussian_renderer/__init__.py:82-101); `raw_params` = opacities / scales / rotations are the
    pre-activation parameters and sigmoid / exp / normalize run inside the preprocess kernel."""
    if means3D.dim() != 2 or means3D.shape[1] != 3:
        raise L.PdgsError("means3D must have dimensions (num_points, 3)")
    if prefiltered:
        # upstream traps the kernel when a point marked as prefiltered fails the frustum test; the reference always
        # passes False (/root/reference/gaussian_renderer/__init__.py:47).  Reject instead of silently ignoring it.
        raise L.PdgsError("prefiltered=True is not supported: this rasterizer always culls by itself (pass False)")
    lib = L.lib()
    dev = means3D.device
    if not means3D.is_cuda:
        raise L.PdgsError("means3D must be a CUDA tensor (this rasterizer has no CPU path)")
    P, H, W = means3D.shape[0], int(image_height), int(image_width)
    means3D = _f32c(means3D, "means3D", dev)
    sh = _f32c(sh, "sh", dev)
    dc = _f32c(dc, "dc", dev)
    colors_precomp = _f32c(colors_precomp, "colors_precomp", dev)
    opacities = _f32c(opacities, "opacities", dev)
    scales = _f32c(scales, "scales", dev)
    rotations = _f32c(rotations, "rotations", dev)
    cov3D_precomp = _f32c(cov3D_precomp, "cov3D_precomp", dev)
    bg = _f32c(bg, "bg", dev)
    viewmatrix = _f32c(viewmatrix, "viewmatrix", dev)
    projmatrix = _f32c(projmatrix, "projmatrix", dev)
    campos = _f32c(campos, "campos", dev)
    if sh is not None and sh.dim() != 3:
        raise L.PdgsError("sh must have dimensions (num_points, num_coeffs, 3)")
    if dc is not None and (dc.dim() != 3 or dc.shape[1] != 1 or dc.shape[0] != P):
        raise L.PdgsError("dc must have dimensions (num_points, 1, 3)")
    with torch.cuda.device(dev):
        h = L.handle(dev.index if dev.index is not None else torch.cuda.current_device())
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        u8 = dict(dtype=torch.uint8, device=dev)
        color = torch.empty(3, H, W, dtype=torch.float32, device=dev)
        invdepth = torch.empty(1, H, W, dtype=torch.float32, device=dev)
        radii = torch.empty(P, dtype=torch.int32, device=dev)
        geom = torch.empty(lib.pdgs_geom_bytes(P), **u8)
        image = torch.empty(lib.pdgs_image_bytes(W, H), **u8)
        hint_key = (dev, W, H)
        hwm = float(_capacity_hint.get(hint_key, 0.0))
        cap = (int(hwm * _CAP_HEADROOM) + 4096) if (P > 0 and hwm > 0) else 0
        binning = None
        if cap > 0:
            binning = torch.empty(lib.pdgs_binning_bytes(cap, P), **u8)
        cam = _camera(bg, scale_modifier, viewmatrix, projmatrix, tanfovx, tanfovy, H, W, sh_degree, campos,
                      prefiltered, antialiasing, debug)
        sc = _scene(P, means3D, sh, colors_precomp, opacities, scales, rotations, cov3D_precomp, dc, raw_params)
        st = _state(geom, image, binning, cap)
        out = L.PdgsForwardOut(_ptr(color), _ptr(invdepth), _ptr(radii))
        R = C.c_int64(0)
        rc = L.check(lib.pdgs_forward(h, C.byref(sc), C.byref(cam), C.byref(st), C.byref(out), 0, C.byref(R), stream),
                     "pdgs_forward")
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        launches = 1 if P > 0 else 0                       # preprocess
        if P > 0 and cap == 0:
            launches += 1                                  # stand-alone scan of the sizing call
        per_bin = 1 + 2 + _passes(tiles) + 1 + 1           # scan + duplicate, 2 histograms, radix passes (4 depth + tile id), ranges, blend
        if cap > 0:
            launches += per_bin
        _count("calls")
        if rc == L.PDGS_NEED_BINNING:
            _count("misses" if cap > 0 else "sizing_calls")
            cap = max(int(R.value * _CAP_HEADROOM) + 4096, 4096)
            binning = torch.empty(lib.pdgs_binning_bytes(cap, P), **u8)
            st = _state(geom, image, binning, cap)
            L.check(lib.pdgs_forward(h, C.byref(sc), C.byref(cam), C.byref(st), C.byref(out), 1, C.byref(R), stream),
                    "pdgs_forward(resume)")
            launches += per_bin
        if P > 0:
            _capacity_hint[hint_key] = max(float(R.value), hwm * _CAP_DECAY, 1.0)
        if binning is None:
            binning = torch.empty(0, **u8)
        binning._pdgs_capacity = cap
        LAUNCHES.forward = launches
    return int(R.value), color, radii, geom, binning, image, invdepth


def rasterize_gaussians_backward(bg, means3D, radii, colors_precomp, opacities, scales, rotations, scale_modifier,
                                 cov3D_precomp, viewmatrix, projmatrix, tanfovx, tanfovy, grad_out_color,
                                 grad_out_invdepth, sh, sh_degree, campos, geomBuffer, num_rendered, binningBuffer,
                                 imgBuffer, antialiasing, debug, accumulate_into: Optional[dict] = None, dc=None,
                                 raw_params=False):
    """-> (grad_means2D[P,3], grad_colors_precomp[P,3], grad_opacities[P,1], grad_means3D[P,3],
           grad_cov3Ds_precomp[P,6], grad_sh[P,M,3], grad_scales[P,3], grad_rotations[P,4]).

    accumulate_into (extension used by the view-parallel path): a dict of preallocated gradient
    tensors keyed like PdgsGrads; the kernel adds into them instead of writing fresh tensors (or overwrites them when
    the dict has "overwrite": True); "densify_grad_accum" / "densify_denom" / "max_radii2D" receive the statistics.
    dc / raw_params: as in rasterize_gaussians; with `dc` the tuple gains a ninth element grad_dc[P,1,3] and grad_sh is
    the gradient of features_rest; with raw_params the opacity / scale / rotation gradients are w.r.t. the raw tensors."""
    lib = L.lib()
    dev = means3D.device
    P = means3D.shape[0]
    H, W = int(grad_out_color.shape[-2]), int(grad_out_color.shape[-1])
    means3D = _f32c(means3D, "means3D", dev)
    sh = _f32c(sh, "sh", dev)
    dc = _f32c(dc, "dc", dev)
    colors_precomp = _f32c(colors_precomp, "colors_precomp", dev)
    opacities = _f32c(opacities, "opacities", dev)
    scales = _f32c(scales, "scales", dev)
    rotations = _f32c(rotations, "rotations", dev)
    cov3D_precomp = _f32c(cov3D_precomp, "cov3D_precomp", dev)
    bg = _f32c(bg, "bg", dev)
    viewmatrix = _f32c(viewmatrix, "viewmatrix", dev)
    projmatrix = _f32c(projmatrix, "projmatrix", dev)
    campos = _f32c(campos, "campos", dev)
    grad_out_color = _f32c(grad_out_color, "grad_out_color", dev)
    grad_out_invdepth = _f32c(grad_out_invdepth, "grad_out_invdepth", dev)
    M = sh.shape[1] if sh is not None else 0
    f32 = dict(dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        h = L.handle(dev.index if dev.index is not None else torch.cuda.current_device())
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        g = L.PdgsGrads()
        if accumulate_into is None:
            # the kernel writes every row (zeros for culled Gaussians), so no memset is needed
            dmeans2D = torch.empty(P, 3, **f32)
            dmeans3D = torch.empty(P, 3, **f32)
            dopac = torch.empty(P, 1, **f32)
            dcolors = torch.empty(P, 3, **f32) if colors_precomp is not None else torch.zeros(P, 3, **f32)
            dsh = torch.empty(P, M, 3, **f32) if sh is not None else torch.zeros(P, 0, 3, **f32)
            ddc = torch.empty(P, 1, 3, **f32) if dc is not None else None
            dcov = torch.empty(P, 6, **f32) if cov3D_precomp is not None else torch.zeros(P, 6, **f32)
            dscales = torch.empty(P, 3, **f32) if scales is not None else torch.zeros(P, 3, **f32)
            drots = torch.empty(P, 4, **f32) if rotations is not None else torch.zeros(P, 4, **f32)
            g.accumulate = 0
        else:
            a = accumulate_into
            dmeans2D, dmeans3D, dopac = a.get("means2D"), a.get("means3D"), a.get("opacities")
            dcolors, dsh, dcov = a.get("colors_precomp"), a.get("shs"), a.get("cov3D_precomp")
            dscales, drots = a.get("scales"), a.get("rotations")
            ddc = a.get("shs_dc")
            if dc is not None:
                dsh = a.get("shs_rest")
            # "overwrite": True -> the kernel writes every row of the given tensors (zeros for culled Gaussians) instead of
            # adding to them; the densification statistics are still accumulated
            g.accumulate = 0 if a.get("overwrite") else 1
        g.means3D, g.means2D = _ptr(dmeans3D), _ptr(dmeans2D)
        if dc is not None:
            g.shs_dc, g.shs_rest = _ptr(ddc), _ptr(dsh)
        else:
            g.shs = _ptr(dsh) if sh is not None else None
        g.colors_precomp = _ptr(dcolors) if colors_precomp is not None else None
        g.opacities = _ptr(dopac)
        g.scales = _ptr(dscales) if scales is not None else None
        g.rotations = _ptr(drots) if rotations is not None else None
        g.cov3D_precomp = _ptr(dcov) if cov3D_precomp is not None else None
        if accumulate_into is not None:
            g.densify_grad_accum = _ptr(accumulate_into.get("densify_grad_accum"))
            g.densify_denom = _ptr(accumulate_into.get("densify_denom"))
            g.max_radii2D = _ptr(accumulate_into.get("max_radii2D"))
            # deferred SH backward (sh_backward_views): this view only leaves its masked dL/drgb [P,3]
            g.color_grad_view = (_ptr(accumulate_into.get("color_grad_view"))
                                 if (sh is not None or dc is not None) else None)
        if P > 0:
            scratch = torch.empty(lib.pdgs_backward_scratch_bytes(P), dtype=torch.uint8, device=dev)
            cam = _camera(bg, scale_modifier, viewmatrix, projmatrix, tanfovx, tanfovy, H, W, sh_degree, campos,
                          False, antialiasing, debug)
            sc = _scene(P, means3D, sh, colors_precomp, opacities, scales, rotations, cov3D_precomp, dc, raw_params)
            st = _state(geomBuffer, imgBuffer, binningBuffer, _binning_capacity(binningBuffer))
            L.check(lib.pdgs_backward(h, C.byref(sc), C.byref(cam), C.byref(st), _ptr(radii), int(num_rendered),
                                      _ptr(grad_out_color), _ptr(grad_out_invdepth), _ptr(scratch), scratch.numel(),
                                      C.byref(g), stream), "pdgs_backward")
            LAUNCHES.backward = 2 if num_rendered > 0 else 1
    if dc is not None:
        return dmeans2D, dcolors, dopac, dmeans3D, dcov, dsh, dscales, drots, ddc
    return dmeans2D, dcolors, dopac, dmeans3D, dcov, dsh, dscales, drots


def sh_backward_views_split(means3D, dc, rest, sh_degree, campos_views, color_grads, grad_dc, grad_rest, grad_means3D,
                            accumulate=True, first=0, count=0):
    """sh_backward_views for a scene given as features_dc [P,1,3] / features_rest [P,M-1,3]
    (C ABI pdgs_sh_backward_views_split).  first / count: Gaussians [first, first + count) only (count 0 = all;
    `first` a multiple of 256)."""
    lib = L.lib()
    dev = means3D.device
    if not means3D.is_cuda:
        raise L.PdgsError("means3D must be a CUDA tensor (this rasterizer has no CPU path)")
    means3D, dc, rest = _f32c(means3D, "means3D", dev), _f32c(dc, "dc", dev), _f32c(rest, "rest", dev)
    campos_views, color_grads = _f32c(campos_views, "campos_views", dev), _f32c(color_grads, "color_grads", dev)
    P, V = means3D.shape[0], int(campos_views.shape[0])
    if color_grads.numel() != V * P * 3:
        raise L.PdgsError("color_grads must be [V,P,3]")
    with torch.cuda.device(dev):
        h = L.handle(dev.index if dev.index is not None else torch.cuda.current_device())
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        sc = _scene(P, means3D, rest, None, None, None, None, None, dc)
        L.check(lib.pdgs_sh_backward_views_split(h, C.byref(sc), int(sh_degree), _ptr(campos_views), _ptr(color_grads), V,
                                                 _ptr(grad_dc), _ptr(grad_rest), _ptr(grad_means3D), int(bool(accumulate)),
                                                 int(first), int(count), stream), "pdgs_sh_backward_views_split")
        LAUNCHES.sh_bwd = (V + 255) // 256
    return grad_dc, grad_rest, grad_means3D


def sh_backward_views(means3D, sh, sh_degree, campos_views, color_grads, grad_shs, grad_means3D, accumulate=True,
                      first=0, count=0):
    """SH part of the backward for a batch of views at once (C ABI pdgs_sh_backward_views): reads `color_grads`
    [V,P,3] (what backward left in accumulate_into["color_grad_view"] per view) and `campos_views` [V,3]; writes
    or accumulates grad_shs [P,M,3] and adds the view-direction term to grad_means3D [P,3].  first / count: Gaussians
    [first, first + count) only (count 0 = all; `first` a multiple of 4)."""
    lib = L.lib()
    dev = means3D.device
    if not means3D.is_cuda:
        raise L.PdgsError("means3D must be a CUDA tensor (this rasterizer has no CPU path)")
    means3D, sh = _f32c(means3D, "means3D", dev), _f32c(sh, "sh", dev)
    campos_views, color_grads = _f32c(campos_views, "campos_views", dev), _f32c(color_grads, "color_grads", dev)
    P, M, V = means3D.shape[0], sh.shape[1], int(campos_views.shape[0])
    if color_grads.numel() != V * P * 3:
        raise L.PdgsError("color_grads must be [V,P,3]")
    with torch.cuda.device(dev):
        h = L.handle(dev.index if dev.index is not None else torch.cuda.current_device())
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        sc = _scene(P, means3D, sh, None, None, None, None, None)
        L.check(lib.pdgs_sh_backward_views(h, C.byref(sc), int(sh_degree), _ptr(campos_views), _ptr(color_grads), V,
                                           _ptr(grad_shs), _ptr(grad_means3D), int(bool(accumulate)), int(first), int(count),
                                           stream), "pdgs_sh_backward_views")
        LAUNCHES.sh_bwd = (V + 255) // 256
    return grad_shs, grad_means3D


def image_loss(color, target, invdepth=None, prior=None, mask=None, w_color=1.0, w_depth=0.0, loss_acc=None,
               clamp=False, alpha_mask=None):
    """One-pass image loss + cotangents (C ABI pdgs_image_loss; the torch ops at
    /root/reference/train.py:153,174).  -> (dL_dcolor[3,H,W], dL_dinvdepth[1,H,W] or None, loss_acc[2]);
    loss_acc (device float[2]: mean L1, mean masked depth L1) is accumulated into when given.
    clamp / alpha_mask: the loss is taken on clamp(color, 0, 1) * alpha_mask as the reference's caller does
    (gaussian_renderer/__init__.py:121, train.py:148-150) and dL_dcolor is w.r.t. the raw `color`."""
    lib = L.lib()
    dev = color.device
    if not color.is_cuda:
        raise L.PdgsError("color must be a CUDA tensor (this rasterizer has no CPU path)")
    H, W = int(color.shape[-2]), int(color.shape[-1])
    color = _f32c(color, "color", dev)
    target = _f32c(target, "target", dev)
    if target is None or target.numel() != color.numel():
        raise L.PdgsError("target must have the shape of color")
    invdepth, prior, mask = _f32c(invdepth, "invdepth", dev), _f32c(prior, "prior", dev), _f32c(mask, "mask", dev)
    alpha_mask = _alpha_plane(alpha_mask, H, W, dev)
    with torch.cuda.device(dev):
        h = L.handle(dev.index if dev.index is not None else torch.cuda.current_device())
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        dcolor = torch.empty_like(color)
        ddepth = torch.empty(1, H, W, dtype=torch.float32, device=dev) if prior is not None else None
        if loss_acc is None:
            loss_acc = torch.zeros(2, dtype=torch.float32, device=dev)
        L.check(lib.pdgs_image_loss(h, _ptr(color), _ptr(target), _ptr(invdepth), _ptr(prior), _ptr(mask),
                                    _ptr(alpha_mask), int(bool(clamp)), W, H, float(w_color), float(w_depth), _ptr(dcolor),
                                    _ptr(ddepth), _ptr(loss_acc), stream), "pdgs_image_loss")
        LAUNCHES.loss = 1
    return dcolor, ddepth, loss_acc


def _alpha_plane(alpha_mask, H, W, dev):
    """alpha_mask [1,H,W] / [H,W] (the camera's alpha_mask, /root/reference/train.py:148-150) as one fp32 plane."""
    alpha_mask = _f32c(alpha_mask, "alpha_mask", dev)
    if alpha_mask is not None and alpha_mask.numel() != H * W:
        raise L.PdgsError("alpha_mask must hold one value per pixel ([1,H,W] or [H,W])")
    return alpha_mask


def image_loss_ssim(color, target, invdepth=None, prior=None, mask=None, lambda_dssim=0.2, w_depth=0.0, loss_acc=None,
                    scratch=None, clamp=True, alpha_mask=None):
    """The reference's training loss and its cotangents (C ABI pdgs_image_loss_ssim; the torch ops at
    /root/reference/train.py:148-159 on the image render() returns, i.e. `image = render.clamp(0, 1)`
    (gaussian_renderer/__init__.py:121), `image *= alpha_mask`, `(1-l)*l1 + l*(1-ssim)`, and :174 the depth-prior
    term).  `color` is the RAW rasterizer output; clamp=True (default) applies the clamp inside the kernels and
    dL_dcolor is zero where color is outside [0, 1].
    -> (dL_dcolor[3,H,W], dL_dinvdepth[1,H,W] or None, loss_acc[3] = mean L1, mean masked depth L1, mean SSIM)."""
    lib = L.lib()
    dev = color.device
    if not color.is_cuda:
        raise L.PdgsError("color must be a CUDA tensor (this rasterizer has no CPU path)")
    H, W = int(color.shape[-2]), int(color.shape[-1])
    color = _f32c(color, "color", dev)
    target = _f32c(target, "target", dev)
    if target is None or target.numel() != color.numel() or color.numel() != 3 * H * W:
        raise L.PdgsError("color and target must both be [3,H,W]")
    invdepth, prior, mask = _f32c(invdepth, "invdepth", dev), _f32c(prior, "prior", dev), _f32c(mask, "mask", dev)
    alpha_mask = _alpha_plane(alpha_mask, H, W, dev)
    with torch.cuda.device(dev):
        h = L.handle(dev.index if dev.index is not None else torch.cuda.current_device())
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        need = lib.pdgs_ssim_scratch_bytes(W, H, 3)
        if scratch is None or scratch.numel() < need:
            scratch = torch.empty(need, dtype=torch.uint8, device=dev)
        dcolor = torch.empty_like(color)
        ddepth = torch.empty(1, H, W, dtype=torch.float32, device=dev) if prior is not None else None
        if loss_acc is None:
            loss_acc = torch.zeros(3, dtype=torch.float32, device=dev)
        L.check(lib.pdgs_image_loss_ssim(h, _ptr(color), _ptr(target), _ptr(invdepth), _ptr(prior), _ptr(mask),
                                         _ptr(alpha_mask), int(bool(clamp)), W, H, float(lambda_dssim), float(w_depth),
                                         _ptr(scratch), scratch.numel(), _ptr(dcolor), _ptr(ddepth), _ptr(loss_acc), stream),
                "pdgs_image_loss_ssim")
        LAUNCHES.loss = 3 if prior is not None else 2
    return dcolor, ddepth, loss_acc


def ssim_scratch_bytes(width: int, height: int, planes: int = 3) -> int:
    return int(L.lib().pdgs_ssim_scratch_bytes(int(width), int(height), int(planes)))


def ssim_forward(C1, C2, img1, img2, train=True, want_map=True, ssim_sum=None):
    """fusedssim(C1, C2, img1, img2, train) of the reference's fused-ssim module
    (/root/reference/submodules/fused-ssim/ssim.cu:368-404): img [B,CH,H,W] fp32 CUDA ->
    (ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12); the partial maps are empty tensors when not train."""
    lib = L.lib()
    if not img1.is_cuda:
        raise L.PdgsError("img1 must be a CUDA tensor (no CPU path)")
    dev = img1.device
    if img1.dim() != 4 or img1.shape != img2.shape:
        raise L.PdgsError("img1 and img2 must be [B,CH,H,W] tensors of the same shape")
    img1, img2 = _f32c(img1, "img1", dev), _f32c(img2, "img2", dev)
    B, CH, H, W = (int(v) for v in img1.shape)
    with torch.cuda.device(dev):
        h = L.handle(dev.index if dev.index is not None else torch.cuda.current_device())
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        smap = torch.empty_like(img1) if want_map else None
        if train:
            d1, d2, d3 = torch.empty_like(img1), torch.empty_like(img1), torch.empty_like(img1)
        else:
            d1 = d2 = d3 = None
        if B * CH > 0 and H * W > 0:
            L.check(lib.pdgs_ssim_forward(h, float(C1), float(C2), _ptr(img1), _ptr(img2), B * CH, W, H, _ptr(smap), _ptr(d1),
                                          _ptr(d2), _ptr(d3), _ptr(ssim_sum), stream), "pdgs_ssim_forward")
    e = torch.empty(0, device=dev)
    return smap, (d1 if train else e), (d2 if train else e), (d3 if train else e)


def ssim_backward(C1, C2, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12):
    """fusedssim_backward of the reference (/root/reference/submodules/fused-ssim/ssim.cu:406-444) -> dL_dimg1.
    dL_dmap may be a python float (uniform cotangent of a mean)."""
    lib = L.lib()
    dev = img1.device
    img1, img2 = _f32c(img1, "img1", dev), _f32c(img2, "img2", dev)
    B, CH, H, W = (int(v) for v in img1.shape)
    uniform = 0.0
    if isinstance(dL_dmap, (int, float)):
        uniform, dL_dmap = float(dL_dmap), None
    else:
        dL_dmap = _f32c(dL_dmap, "dL_dmap", dev)
    d1, d2, d3 = _f32c(dm_dmu1, "dm_dmu1", dev), _f32c(dm_dsigma1_sq, "dm_dsigma1_sq", dev), _f32c(dm_dsigma12, "dm_dsigma12", dev)
    with torch.cuda.device(dev):
        h = L.handle(dev.index if dev.index is not None else torch.cuda.current_device())
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        out = torch.empty_like(img1)
        if B * CH > 0 and H * W > 0:
            L.check(lib.pdgs_ssim_backward(h, _ptr(img1), _ptr(img2), _ptr(dL_dmap), uniform, _ptr(d1), _ptr(d2), _ptr(d3),
                                           B * CH, W, H, _ptr(out), stream), "pdgs_ssim_backward")
    return out


def mark_visible(means3D, viewmatrix, projmatrix):
    """-> bool[P]; the near-plane test of the reference's checkFrustum."""
    lib = L.lib()
    if not means3D.is_cuda:
        raise L.PdgsError("means3D must be a CUDA tensor (this rasterizer has no CPU path)")
    dev = means3D.device
    P = means3D.shape[0]
    means3D = _f32c(means3D, "means3D", dev)
    viewmatrix = _f32c(viewmatrix, "viewmatrix", dev)
    projmatrix = _f32c(projmatrix, "projmatrix", dev)
    present = torch.zeros(P, dtype=torch.bool, device=dev)
    if P > 0:
        with torch.cuda.device(dev):
            stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            L.check(lib.pdgs_mark_visible(_ptr(means3D), _ptr(viewmatrix), _ptr(projmatrix), P, _ptr(present), stream),
                    "pdgs_mark_visible")
    return present


def sort_pairs(keys: torch.Tensor, vals: torch.Tensor, end_bit: int):
    """Stable radix sort of (uint64 key as int64 tensor, uint32 value as int32 tensor) pairs over
    key bits [0, end_bit) with the rasterizer's onesweep kernels.  Inputs are clobbered."""
    lib = L.lib()
    dev = keys.device
    n = keys.numel()
    assert keys.dtype == torch.int64 and vals.dtype == torch.int32 and vals.numel() == n
    keys_out, vals_out = torch.empty_like(keys), torch.empty_like(vals)
    if n == 0:
        return keys_out, vals_out
    with torch.cuda.device(dev):
        nb = lib.pdgs_sort_scratch_bytes(n)
        scratch = torch.empty(nb, dtype=torch.uint8, device=dev)
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        L.check(lib.pdgs_sort_pairs(_ptr(keys), _ptr(vals), _ptr(keys_out), _ptr(vals_out), n, int(end_bit),
                                    _ptr(scratch), nb, stream), "pdgs_sort_pairs")
    return keys_out, vals_out


def state_view(geom, binning, image, P: int, W: int, H: int, num_rendered: int, to_cpu: bool = True) -> dict:
    """Copies of the saved state as torch tensors (parity tests); to_cpu=False keeps the copies on the device."""
    lib = L.lib()
    st = _state(geom, image, binning if binning.numel() else None, _binning_capacity(binning))
    v = L.PdgsStateView()
    L.check(lib.pdgs_state_view(C.byref(st), P, W, H, C.byref(v)), "pdgs_state_view")
    dev = geom.device
    tiles = ((W + 15) // 16) * ((H + 15) // 16)

    def grab(ptr, count, dtype):
        if not ptr or count == 0:
            return torch.empty(0, dtype=dtype)
        nbytes = count * torch.empty(0, dtype=dtype).element_size()
        # locate the view inside its blob and slice
        for blob in (geom, binning, image):
            if blob.numel() and blob.data_ptr() <= ptr < blob.data_ptr() + blob.numel():
                off = ptr - blob.data_ptr()
                t = blob[off:off + nbytes].view(dtype).clone()
                return t.cpu() if to_cpu else t
        raise L.PdgsError("state view pointer outside the state blobs")

    counters = _work_counters_on(dev)
    R = int(num_rendered)
    point_list = grab(v.point_list, R, torch.int32)
    sorted_tiles = grab(v.sorted_tiles, R, torch.int32)
    conic_opacity = grab(v.conic_opacity, P * 4, torch.float32).view(-1, 4)
    rgb = grab(v.rgb, P * 4, torch.float32).view(-1, 4)
    # the reference's 64-bit sort key of every pair: (tile << 32) | depth bits of its Gaussian
    depth32 = conic_opacity[:, 3].contiguous().view(torch.int32)
    depth_bits = depth32.to(torch.int64) & 0xFFFFFFFF
    sorted_keys = (sorted_tiles.to(torch.int64) << 32) | depth_bits[point_list.to(torch.int64)]
    return dict(
        point_list=point_list, sorted_keys=sorted_keys, sorted_tiles=sorted_tiles,
        ranges=grab(v.ranges, tiles * 2, torch.int32).view(-1, 2), final_T=grab(v.final_T, W * H, torch.float32),
        n_contrib=grab(v.n_contrib, W * H, torch.int32),
        # written by the forward blend only while the work counters are on; otherwise the planes are uninitialised memory
        n_tested=grab(v.n_tested, W * H, torch.int32) if counters else None,
        n_accepted=grab(v.n_accepted, W * H, torch.int32) if counters else None,
        tiles_touched=grab(v.tiles_touched, P, torch.int32),
        point_offsets=grab(v.point_offsets, P, torch.int32), depth_order=grab(v.depth_order, P, torch.int32),
        xy=grab(v.xy, P * 4, torch.float32).view(-1, 4), conic_opacity=conic_opacity, rgb=rgb, depth_bits=depth32)


def image_counters(geom, binning, image, P: int, W: int, H: int) -> dict:
    """Sums over pixels of the forward blend's per-pixel counters, reduced on the device:
    n_tested (list entries evaluated), n_accepted (entries blended), n_contrib (last contributing position).
    The first two exist only for a forward made while set_work_counters(device, True) was in effect on the calling
    thread; without it the planes were never written, so this raises instead of summing uninitialised memory."""
    lib = L.lib()
    if not _work_counters_on(image.device):
        raise L.PdgsError("image_counters: the work counters are off for this thread's handle; call "
                          "set_work_counters(device, True) before the forward whose counters you want")
    st = _state(geom, image, binning if binning.numel() else None, _binning_capacity(binning))
    v = L.PdgsStateView()
    L.check(lib.pdgs_state_view(C.byref(st), P, W, H, C.byref(v)), "pdgs_state_view")
    out = {}
    for name in ("n_tested", "n_accepted", "n_contrib"):
        off = getattr(v, name) - image.data_ptr()
        out[name] = int(image[off:off + 4 * W * H].view(torch.int32).sum(dtype=torch.int64).item())
    return out


def set_stage_timing(device, enable: bool) -> None:
    """Turn the library's per-stage CUDA-event timing on/off for every thread's handle on `device` (autograd runs
    backward on its own thread, hence with its own handle)."""
    dev = torch.device(device)
    with torch.cuda.device(dev):
        idx = dev.index if dev.index is not None else torch.cuda.current_device()
        L.handle(idx)
        L._timing[int(idx)] = bool(enable)
        for h in L.device_handles(idx):
            L.check(L.lib().pdgs_set_stage_timing(h, int(bool(enable))), "pdgs_set_stage_timing")


_work_counters = {}  # (thread id, device index) -> work counters enabled on that thread's handle


def _work_counters_on(device) -> bool:
    import threading
    dev = torch.device(device)
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    return bool(_work_counters.get((threading.get_ident(), int(idx))))


def set_work_counters(device, enable: bool) -> None:
    """Turn the forward blend's per-pixel work counters (n_tested / n_accepted, read by image_counters / state_view) on
    or off for the CALLING thread's handle on `device`.  They are off by default (C ABI pdgs_set_work_counters)."""
    dev = torch.device(device)
    with torch.cuda.device(dev):
        idx = dev.index if dev.index is not None else torch.cuda.current_device()
        L.check(L.lib().pdgs_set_work_counters(L.handle(idx), int(bool(enable))), "pdgs_set_work_counters")
        import threading
        _work_counters[(threading.get_ident(), int(idx))] = bool(enable)


def stage_times(device) -> dict:
    """{stage: (summed duration in ms, number of launches of the stage)} since the previous call on `device`,
    summed over the handles of all threads (waits for the recorded work)."""
    dev = torch.device(device)
    tot = [0.0] * len(L.STAGES)
    num = [0] * len(L.STAGES)
    with torch.cuda.device(dev):
        idx = dev.index if dev.index is not None else torch.cuda.current_device()
        L.handle(idx)
        for h in L.device_handles(idx):
            buf = (C.c_float * len(L.STAGES))()
            cnt = (C.c_int32 * len(L.STAGES))()
            L.check(L.lib().pdgs_stage_times(h, buf, cnt, len(L.STAGES)), "pdgs_stage_times")
            for i in range(len(L.STAGES)):
                tot[i] += float(buf[i])
                num[i] += int(cnt[i])
    return {name: (tot[i], num[i]) for i, name in enumerate(L.STAGES)}
