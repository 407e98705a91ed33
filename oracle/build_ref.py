#!/usr/bin/env python
"""Recipe for oracle/_ref/ -- the REAL reference, built / staged from the sources where they lie under
/root/reference.  TEST INFRASTRUCTURE ONLY: nothing here is imported by the product path; only tests/ and bench.py's
baseline legs load what this script produces.  oracle/_ref/ is git-ignored (no reference source enters the history)
but travels to the GPU box with the snapshot, like the repo's own built .so files.

Two artefacts:

1. oracle/_ref/fused_ssim_cuda*.so -- the reference's own fused-SSIM CUDA extension
   (/root/reference/submodules/fused-ssim: ssim.cu, ext.cpp, ssim.h; its setup.py:8-12 names exactly these two
   sources), compiled for sm_100a with torch's extension builder straight from /root/reference.  It is the parity
   target and the timed GPU denominator of SURVEY 8(f) rank 3 (tests/test_gpu_ref_fused_ssim.py, bench.py
   `library_baselines`).

2. oracle/_ref/ref_py/ -- byte copies of the reference's own Python for the hot path's CALLER:
   gaussian_renderer/__init__.py (render()), scene/gaussian_model.py (GaussianModel: activations, training_setup,
   add_densification_stats), scene/cameras.py (MiniCam), utils/{sh_utils,general_utils,loss_utils,graphics_utils,
   system_utils}.py, arguments/__init__.py, and the fused-ssim package's fused_ssim/__init__.py (as
   ref_fused_ssim_py/__init__.py).  tests/test_gpu_reference_render.py runs the reference's render() and training-step
   text on top of the drop-in `diff_gaussian_rasterization` package with them.  /root/reference does not exist on the
   GPU box, hence the staged copy.

The rasterizer itself (submodules/diff-gaussian-rasterization) is an EMPTY directory in this checkout
(.SUBMODULES.json: status "missing"), so there is nothing to build for it: unbuildable, parity unpinned.

  python oracle/build_ref.py [--force]
"""
from __future__ import annotations

import glob
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("PDGS_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")

REF_PY = [
    "gaussian_renderer/__init__.py",
    "scene/gaussian_model.py",
    "scene/cameras.py",
    "utils/sh_utils.py",
    "utils/general_utils.py",
    "utils/loss_utils.py",
    "utils/graphics_utils.py",
    "utils/system_utils.py",
    "arguments/__init__.py",
]


def stage_ref_py(force: bool = False) -> str:
    dst_root = os.path.join(OUT, "ref_py")
    for rel in REF_PY:
        src, dst = os.path.join(REF, rel), os.path.join(dst_root, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if force or not os.path.exists(dst) or os.path.getmtime(dst) < os.path.getmtime(src):
            shutil.copyfile(src, dst)
    # package markers that are NOT the reference's: its scene/__init__.py and utils have import-time side effects
    # (dataset readers, COLMAP loaders) outside the hot path
    for pkg in ("scene", "utils"):
        marker = os.path.join(dst_root, pkg, "__init__.py")
        if not os.path.exists(marker):
            open(marker, "w").write("# package marker written by oracle/build_ref.py (not a reference file)\n")
    fs = os.path.join(dst_root, "ref_fused_ssim_py")
    os.makedirs(fs, exist_ok=True)
    shutil.copyfile(os.path.join(REF, "submodules/fused-ssim/fused_ssim/__init__.py"), os.path.join(fs, "__init__.py"))
    return dst_root


def built_fused_ssim():
    hits = sorted(glob.glob(os.path.join(OUT, "fused_ssim_cuda*.so")))
    return hits[0] if hits else None


def build_fused_ssim(force: bool = False, verbose: bool = False) -> str:
    """nvcc/g++ via torch.utils.cpp_extension.load (ninja), sources read in place from /root/reference."""
    src_dir = os.path.join(REF, "submodules", "fused-ssim")
    sources = [os.path.join(src_dir, "ssim.cu"), os.path.join(src_dir, "ext.cpp")]
    have = built_fused_ssim()
    if have and not force and all(os.path.getmtime(have) >= os.path.getmtime(s) for s in sources):
        return have
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")  # no GPU in the authoring container to probe
    os.environ.setdefault("MAX_JOBS", "4")
    from torch.utils import cpp_extension
    build_dir = os.path.join(OUT, "build_fused_ssim")
    os.makedirs(build_dir, exist_ok=True)
    cpp_extension.load(name="fused_ssim_cuda", sources=sources, extra_include_paths=[src_dir],
                       extra_cuda_cflags=["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3"],
                       build_directory=build_dir, verbose=verbose, is_python_module=False)
    so = os.path.join(build_dir, "fused_ssim_cuda.so")
    dst = os.path.join(OUT, "fused_ssim_cuda.so")
    shutil.copyfile(so, dst)
    return dst


def main(force: bool = False, verbose: bool = False) -> dict:
    if not os.path.isdir(REF):
        raise SystemExit(f"{REF} not present: oracle/_ref can only be (re)built in the authoring container")
    os.makedirs(OUT, exist_ok=True)
    out = dict(ref_py=stage_ref_py(force))
    try:
        out["fused_ssim_cuda"] = build_fused_ssim(force, verbose)
    except Exception as e:  # recorded, not fatal: the tests that need it skip with this reason
        out["fused_ssim_cuda"] = None
        out["fused_ssim_error"] = f"{type(e).__name__}: {e}"[:2000]
        open(os.path.join(OUT, "fused_ssim_build_error.txt"), "w").write(out["fused_ssim_error"])
    return out


if __name__ == "__main__":
    print(main(force="--force" in sys.argv, verbose="-v" in sys.argv))
