"""Loads the REAL reference artefacts staged under oracle/_ref/ by oracle/build_ref.py (tests only).

  ref_fused_ssim()  -> the reference's fused-SSIM CUDA extension module (fusedssim / fusedssim_backward), or None
  ref_py()          -> context manager that puts the staged copy of the reference's own Python (render(), GaussianModel,
                       MiniCam, loss_utils, arguments) on sys.path with the non-hot-path third-party imports stubbed
"""
from __future__ import annotations

import contextlib
import importlib
import importlib.util
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
REF_PY = os.path.join(REF_DIR, "ref_py")
_REF_PKGS = ("gaussian_renderer", "scene", "utils", "arguments", "ref_fused_ssim_py")


def ref_fused_ssim():
    path = os.path.join(REF_DIR, "fused_ssim_cuda.so")
    if not os.path.exists(path):
        return None
    import torch  # noqa: F401  (libtorch must be loaded first)
    if "fused_ssim_cuda" in sys.modules:
        return sys.modules["fused_ssim_cuda"]
    spec = importlib.util.spec_from_file_location("fused_ssim_cuda", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    sys.modules["fused_ssim_cuda"] = mod
    return mod


def have_ref_py() -> bool:
    return os.path.exists(os.path.join(REF_PY, "gaussian_renderer", "__init__.py"))


@contextlib.contextmanager
def ref_py(sparse_adam: bool):
    """Import context for the staged reference Python.  `sparse_adam` selects what the reference's optional-symbol
    probes see (/root/reference/train.py:47-52, scene/gaussian_model.py:25-28): True = the drop-in package exports
    SparseGaussianAdam (its default), False = it does not (PDGS_NO_SPARSE_ADAM=1, the pinned commit's behaviour).
    Third-party modules the reference imports at module level but never uses on the hot path are stubbed:
    plyfile (PLY I/O), simple_knn (initialisation from a point cloud)."""
    saved = {k: v for k, v in sys.modules.items()
             if k.split(".")[0] in _REF_PKGS + ("diff_gaussian_rasterization", "plyfile", "simple_knn")}
    for k in saved:
        del sys.modules[k]
    old_env = os.environ.get("PDGS_NO_SPARSE_ADAM")
    os.environ["PDGS_NO_SPARSE_ADAM"] = "0" if sparse_adam else "1"
    ply = types.ModuleType("plyfile")
    ply.PlyData = ply.PlyElement = object
    knn = types.ModuleType("simple_knn")
    knn_c = types.ModuleType("simple_knn._C")
    knn_c.distCUDA2 = lambda *a, **k: (_ for _ in ()).throw(RuntimeError("simple_knn is not on the hot path"))
    knn._C = knn_c
    sys.modules.update({"plyfile": ply, "simple_knn": knn, "simple_knn._C": knn_c})
    sys.path.insert(0, REF_PY)
    try:
        yield
    finally:
        sys.path.remove(REF_PY)
        for k in [k for k in sys.modules if k.split(".")[0] in _REF_PKGS + ("diff_gaussian_rasterization", "plyfile", "simple_knn")]:
            del sys.modules[k]
        sys.modules.update(saved)
        if old_env is None:
            os.environ.pop("PDGS_NO_SPARSE_ADAM", None)
        else:
            os.environ["PDGS_NO_SPARSE_ADAM"] = old_env
