"""GPU denominators for bench.py and the tests (never imported by the product package).

* CUB: the reference rasterizer's own library calls for K2 / K4 (tools/cub_baseline.cu) on the arrays one forward of
  the new pipeline produced -- `cub::DeviceScan::InclusiveSum(tiles_touched[P])` and
  `cub::DeviceRadixSort::SortPairs((tile << 32 | depth bits), Gaussian id)` over bits [0, 32 + bits(T)).
* the reference's fused-ssim CUDA extension (oracle/_ref/fused_ssim_cuda.so, built from
  /root/reference/submodules/fused-ssim by oracle/build_ref.py) timed next to pdgs_image_loss_ssim.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
_cub = None


def cub():
    """ctypes handle of tools/libcub_baseline.so, or None when it has not been built."""
    global _cub
    if _cub is None:
        path = os.path.join(HERE, "libcub_baseline.so")
        if not os.path.exists(path):
            return None
        l = C.CDLL(path)
        vp = C.c_void_p
        l.cubb_scan_temp_bytes.restype = C.c_size_t
        l.cubb_scan_temp_bytes.argtypes = [C.c_int64]
        l.cubb_inclusive_sum_u32.argtypes = [vp, vp, C.c_int64, vp, C.c_size_t, vp]
        for name in ("cubb_sort_u64_temp_bytes", "cubb_sort_u32_temp_bytes"):
            getattr(l, name).restype = C.c_size_t
            getattr(l, name).argtypes = [C.c_int64, C.c_int]
        for name in ("cubb_sort_pairs_u64", "cubb_sort_pairs_u32"):
            getattr(l, name).argtypes = [vp, vp, vp, vp, C.c_int64, C.c_int, vp, C.c_size_t, vp]
        l.cubb_duplicate_with_keys.argtypes = [vp, vp, vp, vp, vp, C.c_int64, C.c_int, C.c_int, vp, vp, vp, vp]
        _cub = l
    return _cub


def _p(t):
    return C.c_void_p(t.data_ptr())


def _time(fn, warmup, iters):
    """median / p10 / p90 of CUDA-event durations (ms) of fn() on the current stream."""
    for _ in range(warmup):
        fn()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in evs:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    q = lambda f: ts[min(len(ts) - 1, int(round(f * (len(ts) - 1))))]
    return dict(median=q(0.5), p10=q(0.1), p90=q(0.9))


def key_bits(tiles: int) -> int:
    """32 + bits(T), bits(n) = index of the MSB + 1."""
    return 32 + max(1, int(tiles).bit_length())


def reference_binning_chain(sv: dict, radii: torch.Tensor, W: int, H: int, timing: dict = None):
    """(keys_unsorted int64 [R], values_unsorted int32 [R]) as the reference's K2 + K3 emit them, computed WITHOUT any output
    of the new pipeline's binning: cub::DeviceScan::InclusiveSum over tiles_touched, then a reference-shaped
    duplicateWithKeys (tools/cub_baseline.cu: one thread per Gaussian in index order, rectangle recomputed from the pixel
    centre and radius, y outer / x inner, key = tile << 32 | depth bits).  Inputs are K1 outputs only:
    sv["xy"], sv["depth_bits"], sv["tiles_touched"] (device-side state view) and the radii the forward returned.
    timing: optional dict; receives duplicate_ms (median / p10 / p90 of the K3 kernel alone)."""
    l = cub()
    assert l is not None, "tools/libcub_baseline.so missing (make -C tools)"
    tt, xy, db = sv["tiles_touched"].contiguous(), sv["xy"].contiguous(), sv["depth_bits"].contiguous()
    dev = tt.device
    assert xy.is_cuda and radii.device == dev and xy.shape[1] == 4
    P = tt.numel()
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    tb = l.cubb_scan_temp_bytes(P)
    temp = torch.empty(max(tb, 1), dtype=torch.uint8, device=dev)
    offs = torch.empty_like(tt)
    assert l.cubb_inclusive_sum_u32(_p(tt), _p(offs), P, _p(temp), tb, stream) == 0
    R = int(offs[-1].item()) if P else 0
    keys = torch.empty(R, dtype=torch.int64, device=dev)
    vals = torch.empty(R, dtype=torch.int32, device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    rad = radii.to(torch.int32).contiguous()
    dup = lambda: l.cubb_duplicate_with_keys(_p(xy), _p(rad), _p(db), _p(offs), _p(tt), P, (W + 15) // 16, (H + 15) // 16,
                                             _p(keys), _p(vals), _p(err), stream)
    assert dup() == 0
    if timing is not None:  # the reference-shaped K3 timed alone (it rewrites the same pairs every time)
        timing["duplicate_ms"] = _time(dup, timing.get("warmup", 3), timing.get("iters", 20))
    assert int(err.item()) == 0, "tiles_touched disagrees with the rectangle recomputed from (xy, radius)"
    return keys, vals


def cub_binning(tiles_touched: torch.Tensor, keys_unsorted: torch.Tensor, vals_unsorted: torch.Tensor, tiles: int,
                warmup: int = 3, iters: int = 20) -> dict:
    """Times the reference's K2 and K4 library calls; returns their durations and the sorted values (parity)."""
    l = cub()
    assert l is not None, "tools/libcub_baseline.so missing (make -C tools)"
    dev = keys_unsorted.device
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    P, R = tiles_touched.numel(), keys_unsorted.numel()
    end_bit = key_bits(tiles)
    out = dict(P=P, R=R, end_bit=end_bit)
    tb = l.cubb_scan_temp_bytes(P)
    temp = torch.empty(max(tb, 1), dtype=torch.uint8, device=dev)
    offs = torch.empty_like(tiles_touched)
    out["scan_ms"] = _time(lambda: l.cubb_inclusive_sum_u32(_p(tiles_touched), _p(offs), P, _p(temp), tb, stream), warmup, iters)
    out["scan_total"] = int(offs[-1].item()) if P else 0
    tb = l.cubb_sort_u64_temp_bytes(R, end_bit)
    temp = torch.empty(max(tb, 1), dtype=torch.uint8, device=dev)
    ko, vo = torch.empty_like(keys_unsorted), torch.empty_like(vals_unsorted)
    out["sort_ms"] = _time(lambda: l.cubb_sort_pairs_u64(_p(keys_unsorted), _p(ko), _p(vals_unsorted), _p(vo), R, end_bit,
                                                        _p(temp), tb, stream), warmup, iters)
    out["sorted_keys"], out["sorted_vals"] = ko, vo
    return out


def cub_sort_u32(keys: torch.Tensor, vals: torch.Tensor, end_bit: int, warmup: int = 3, iters: int = 20) -> dict:
    l = cub()
    dev = keys.device
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    n = keys.numel()
    tb = l.cubb_sort_u32_temp_bytes(n, end_bit)
    temp = torch.empty(max(tb, 1), dtype=torch.uint8, device=dev)
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    ms = _time(lambda: l.cubb_sort_pairs_u32(_p(keys), _p(ko), _p(vals), _p(vo), n, end_bit, _p(temp), tb, stream), warmup, iters)
    return dict(ms=ms, sorted_keys=ko, sorted_vals=vo)


def ref_fused_ssim_module():
    """The reference's fused-ssim extension (fusedssim, fusedssim_backward) from oracle/_ref, or None."""
    import importlib.util
    import sys
    path = os.path.join(ROOT, "oracle", "_ref", "fused_ssim_cuda.so")
    if not os.path.exists(path):
        return None
    if "fused_ssim_cuda" in sys.modules:
        return sys.modules["fused_ssim_cuda"]
    spec = importlib.util.spec_from_file_location("fused_ssim_cuda", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    sys.modules["fused_ssim_cuda"] = mod
    return mod


def ssim_loss_baseline(H: int, W: int, device, lambda_dssim: float = 0.2, warmup: int = 5, iters: int = 30) -> dict:
    """The image loss of one training iteration, value + cotangent, three ways on a [3,H,W] image:
      reference   [REF train.py:153-159]: l1_loss (torch) + the reference's fused_ssim extension + autograd backward
      ours        pdgs_image_loss_ssim (3 launches, clamp inside)
    Returns median ms of each (None when the reference extension is unavailable)."""
    from pdgs_b200 import rasterizer as rz
    gen = torch.Generator().manual_seed(5)
    img = torch.rand(3, H, W, generator=gen).to(device)
    gt = torch.rand(3, H, W, generator=gen).to(device)
    out = dict(image=[W, H])
    scratch = torch.empty(rz.ssim_scratch_bytes(W, H), dtype=torch.uint8, device=device)
    acc = torch.zeros(3, device=device)
    out["pdgs_image_loss_ssim_ms"] = _time(lambda: rz.image_loss_ssim(img, gt, lambda_dssim=lambda_dssim, loss_acc=acc,
                                                                      scratch=scratch), warmup, iters)
    mod = ref_fused_ssim_module()
    if mod is None:
        out["reference_ms"] = None
        return out
    C1, C2 = 0.01 ** 2, 0.03 ** 2

    def ref_step():
        # value and gradient exactly as autograd runs them for [REF train.py:153-159] with FUSED_SSIM_AVAILABLE:
        # forward: clamp, |a-b|.mean(), fusedssim(...).mean(); backward: fusedssim_backward + l1 backward + clamp backward
        a = img.clamp(0, 1).unsqueeze(0)
        b = gt.unsqueeze(0)
        l1 = (a - b).abs().mean()
        smap, d1, d2, d3 = mod.fusedssim(C1, C2, a, b, True)
        s = smap.mean()
        loss = (1.0 - lambda_dssim) * l1 + lambda_dssim * (1.0 - s)
        dmap = torch.full_like(smap, -lambda_dssim / smap.numel())
        g = mod.fusedssim_backward(C1, C2, a, b, dmap, d1, d2, d3)
        g = g + (1.0 - lambda_dssim) / a.numel() * torch.sign(a - b)
        g = g * ((img >= 0) & (img <= 1)).unsqueeze(0)
        return loss, g

    out["reference_ms"] = _time(ref_step, warmup, iters)
    out["reference_kernels_only_ms"] = _time(
        lambda: mod.fusedssim_backward(C1, C2, img.unsqueeze(0), gt.unsqueeze(0), torch.ones(1, 3, H, W, device=device),
                                       *mod.fusedssim(C1, C2, img.unsqueeze(0), gt.unsqueeze(0), True)[1:]), warmup, iters)
    return out
