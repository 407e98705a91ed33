"""GPU denominators and pins (run with -m gpu):

* the reference's fused-SSIM CUDA extension, compiled from /root/reference/submodules/fused-ssim by oracle/build_ref.py
  (oracle/_ref/fused_ssim_cuda.so), against pdgs_ssim_forward / pdgs_ssim_backward and the `fused_ssim` drop-in package
  -- the comparison [REF submodules/fused-ssim/tests/test.py:77-91] makes against its own baseline, here between the
  reference's kernel and this repo's.  This PINS SURVEY 8(f) rank 3 to the reference's own CUDA kernel.
* CUB (the reference rasterizer's K2 / K4 library calls, tools/cub_baseline.cu) against the hand-written onesweep sort
  on the reference-format (u64 tile|depth key, u32 id) pairs of a real view.
Timings are printed (pytest -s) and recorded by bench.py (`library_baselines`); the assertions here are parity only.
"""
import os
import sys

import numpy as np
import pytest
import torch

import helpers
import ref_loader
from pdgs_b200 import scenes

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import gpu_baselines as gb  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
C1, C2 = 0.01 ** 2, 0.03 ** 2

needs_ref_ssim = pytest.mark.skipif(not os.path.exists(os.path.join(ref_loader.REF_DIR, "fused_ssim_cuda.so")),
                                    reason="oracle/_ref/fused_ssim_cuda.so missing: run `python oracle/build_ref.py`")


@needs_ref_ssim
@pytest.mark.parametrize("shape", [(1, 3, 64, 96), (2, 3, 37, 53), (1, 1, 203, 117), (1, 3, 546, 979), (1, 3, 1064, 1600)])
def test_ssim_kernels_match_the_reference_cuda_extension(shape):
    from pdgs_b200 import rasterizer as rz
    ref = ref_loader.ref_fused_ssim()
    gen = torch.Generator().manual_seed(shape[-1])
    a = torch.rand(*shape, generator=gen).to(DEV)
    b = (a.cpu() + 0.25 * torch.randn(*shape, generator=gen)).clamp(0, 1).to(DEV)
    r_map, r1, r2, r3 = ref.fusedssim(C1, C2, a, b, True)
    m_map, m1, m2, m3 = rz.ssim_forward(C1, C2, a, b, train=True)
    # two fp32 kernels, each within ~2e-5 of float64 (sigma = E[x^2] - mu^2 cancels in fp32)
    assert float((m_map - r_map).abs().max()) <= 4e-5
    for got, want in ((m1, r1), (m2, r2), (m3, r3)):
        assert float((got - want).abs().max()) <= 4e-4 * float(want.abs().max())
    gm = torch.randn(*shape, generator=gen).to(DEV)
    r_g = ref.fusedssim_backward(C1, C2, a, b, gm, r1, r2, r3)
    m_g = rz.ssim_backward(C1, C2, a, b, gm, m1, m2, m3)
    assert float((m_g - r_g).abs().max()) <= 2e-4 * float(r_g.abs().max())


@needs_ref_ssim
@pytest.mark.parametrize("padding", ["same", "valid"])
def test_fused_ssim_package_matches_the_reference_package(padding):
    """The reference's python wrapper (staged copy of fused_ssim/__init__.py) over the reference's extension vs the
    drop-in package: value and gradient, like [REF submodules/fused-ssim/tests/test.py:77-91]."""
    if not ref_loader.have_ref_py():
        pytest.skip("oracle/_ref/ref_py missing")
    ref_loader.ref_fused_ssim()   # registers `fused_ssim_cuda` for the wrapper's import
    sys.path.insert(0, ref_loader.REF_PY)
    try:
        from ref_fused_ssim_py import fused_ssim as ref_fused_ssim
    finally:
        sys.path.remove(ref_loader.REF_PY)
    from fused_ssim import fused_ssim
    gen = torch.Generator().manual_seed(3)
    for _ in range(5):
        x = torch.rand(2, 3, 70, 90, generator=gen)
        y = torch.rand(2, 3, 70, 90, generator=gen).to(DEV)
        p_ref = torch.nn.Parameter(x.clone().to(DEV))
        p_mine = torch.nn.Parameter(x.clone().to(DEV))
        v_ref = ref_fused_ssim(p_ref, y, padding)
        v_mine = fused_ssim(p_mine, y, padding)
        assert torch.isclose(v_ref, v_mine, rtol=1e-5, atol=1e-6)
        v_ref.backward()
        v_mine.backward()
        assert float((p_ref.grad - p_mine.grad).abs().max()) <= 2e-4 * float(p_ref.grad.abs().max())


@needs_ref_ssim
def test_ssim_loss_timing_next_to_the_reference_kernel():
    for (W, H) in ((979, 546), (1600, 1064)):
        r = gb.ssim_loss_baseline(H, W, DEV)
        print(f"\n  image loss {W}x{H}: pdgs_image_loss_ssim {r['pdgs_image_loss_ssim_ms']['median']:.4f} ms | reference "
              f"clamp + l1 + fused_ssim fwd + bwd {r['reference_ms']['median']:.4f} ms | reference kernels only "
              f"{r['reference_kernels_only_ms']['median']:.4f} ms")
        assert r["reference_ms"]["median"] > 0 and r["pdgs_image_loss_ssim_ms"]["median"] > 0


@pytest.mark.skipif(gb.cub() is None, reason="tools/libcub_baseline.so missing: make -C tools")
@pytest.mark.parametrize("cfg", [dict(name="C1"), dict(name="C4", P=200_000, scale_image=0.5)])
def test_sorted_list_equals_cub_sort_of_reference_keys(cfg):
    """cub::DeviceRadixSort::SortPairs over bits [0, 32 + bits(T)) of the (tile << 32 | depth, id) pairs in the
    reference's emission order == this pipeline's (depth sort, duplicate, tile sort) list, bit for bit; the stand-alone
    64-bit onesweep sort (pdgs_sort_pairs) gives the same list too."""
    from pdgs_b200 import rasterizer as rz
    sc = scenes.make_config(cfg["name"], P=cfg.get("P"), n_views=2, scale_image=cfg.get("scale_image", 1.0))
    view = sc.views[0]
    W, H = view.image_width, view.image_height
    T = ((W + 15) // 16) * ((H + 15) // 16)
    lv, color, radii, invd = helpers.run_cuda(sc, view, DEV)
    fn = color.grad_fn
    geom, binning, img = fn.saved_tensors[8], fn.saved_tensors[9], fn.saved_tensors[10]
    binning._pdgs_capacity = fn.binning_capacity
    sv = rz.state_view(geom, binning, img, sc.P, W, H, fn.num_rendered, to_cpu=False)
    ku, vu = gb.reference_binning_chain(sv, radii, W, H)   # from K1 outputs only, not from the sorted list under test
    # the unsorted pairs really are in the reference's duplicateWithKeys order: ids ascending, tiles ascending within an id
    assert bool((vu[1:] >= vu[:-1]).all())
    same = vu[1:] == vu[:-1]
    assert bool(((ku[1:] >> 32)[same] > (ku[:-1] >> 32)[same]).all())
    cub = gb.cub_binning(sv["tiles_touched"], ku, vu, T, warmup=1, iters=5)
    assert cub["scan_total"] == fn.num_rendered
    assert torch.equal(cub["sorted_vals"], sv["point_list"]) and torch.equal(cub["sorted_keys"], sv["sorted_keys"])
    ko, vo = rz.sort_pairs(ku.clone(), vu.clone(), cub["end_bit"])
    assert torch.equal(vo, sv["point_list"]) and torch.equal(ko, sv["sorted_keys"])
    print(f"\n  {cfg}: R={fn.num_rendered}, cub scan {cub['scan_ms']['median']:.4f} ms, cub sort {cub['sort_ms']['median']:.4f} ms")
