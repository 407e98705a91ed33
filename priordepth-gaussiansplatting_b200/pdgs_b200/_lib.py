"""ctypes binding of libpdgs.so (C ABI declared in include/pdgs.h).

This is the only way the Python side reaches the rasterizer.  There is no CPU path and no
fallback: if the library is missing or no sm_100 device is present, every call raises.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
# PDGS_LIB: development override used to A/B kernel variants of the same library (never another backend)
LIB_PATH = os.environ.get("PDGS_LIB") or os.path.join(_HERE, "..", "lib", "libpdgs.so")

PDGS_OK = 0
PDGS_NEED_BINNING = 1
PDGS_E_INVALID = -1
PDGS_E_CUDA = -2
PDGS_E_OVERFLOW = -3
PDGS_E_NO_DEVICE = -4

vp = C.c_void_p


class PdgsCamera(C.Structure):
    _fields_ = [("image_height", C.c_int32), ("image_width", C.c_int32), ("tanfovx", C.c_float),
                ("tanfovy", C.c_float), ("bg", vp), ("scale_modifier", C.c_float), ("viewmatrix", vp),
                ("projmatrix", vp), ("sh_degree", C.c_int32), ("campos", vp), ("prefiltered", C.c_int32),
                ("debug", C.c_int32), ("antialiasing", C.c_int32)]


class PdgsScene(C.Structure):
    _fields_ = [("P", C.c_int64), ("M", C.c_int32), ("means3D", vp), ("shs", vp), ("colors_precomp", vp),
                ("opacities", vp), ("scales", vp), ("rotations", vp), ("cov3D_precomp", vp), ("shs_dc", vp),
                ("shs_rest", vp), ("raw_params", C.c_int32)]


class PdgsState(C.Structure):
    _fields_ = [("geom", vp), ("geom_bytes", C.c_size_t), ("image", vp), ("image_bytes", C.c_size_t),
                ("binning", vp), ("binning_bytes", C.c_size_t), ("capacity_pairs", C.c_int64)]


class PdgsForwardOut(C.Structure):
    _fields_ = [("color", vp), ("invdepth", vp), ("radii", vp)]


class PdgsGrads(C.Structure):
    _fields_ = [("means3D", vp), ("means2D", vp), ("shs", vp), ("colors_precomp", vp), ("opacities", vp),
                ("scales", vp), ("rotations", vp), ("cov3D_precomp", vp), ("accumulate", C.c_int32),
                ("densify_grad_accum", vp), ("densify_denom", vp), ("max_radii2D", vp), ("color_grad_view", vp),
                ("shs_dc", vp), ("shs_rest", vp)]


class PdgsAdamGroup(C.Structure):
    _fields_ = [("param", vp), ("grad", vp), ("exp_avg", vp), ("exp_avg_sq", vp), ("lr", C.c_double), ("M", C.c_int32)]


class PdgsStateView(C.Structure):
    _fields_ = [("point_list", vp), ("sorted_tiles", vp), ("ranges", vp), ("final_T", vp), ("n_contrib", vp), ("n_tested", vp), ("n_accepted", vp),
                ("tiles_touched", vp), ("point_offsets", vp), ("depth_order", vp), ("xy", vp), ("conic_opacity", vp), ("rgb", vp)]


# every symbol include/pdgs.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "pdgs_last_error": (C.c_char_p, []),
    "pdgs_version": (C.c_char_p, []),
    "pdgs_create": (C.c_int, [C.POINTER(vp)]),
    "pdgs_destroy": (C.c_int, [vp]),
    "pdgs_set_stage_timing": (C.c_int, [vp, C.c_int32]),
    "pdgs_set_work_counters": (C.c_int, [vp, C.c_int32]),
    "pdgs_stage_times": (C.c_int, [vp, C.POINTER(C.c_float), C.POINTER(C.c_int32), C.c_int32]),
    "pdgs_image_loss": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, C.c_int32, C.c_int32, C.c_int32, C.c_float, C.c_float, vp, vp,
                                  vp, vp]),
    "pdgs_ssim_forward": (C.c_int, [vp, C.c_float, C.c_float, vp, vp, C.c_int32, C.c_int32, C.c_int32, vp, vp, vp, vp, vp, vp]),
    "pdgs_ssim_backward": (C.c_int, [vp, vp, vp, vp, C.c_float, vp, vp, vp, C.c_int32, C.c_int32, C.c_int32, vp, vp]),
    "pdgs_ssim_scratch_bytes": (C.c_size_t, [C.c_int32, C.c_int32, C.c_int32]),
    "pdgs_image_loss_ssim": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, C.c_int32, C.c_int32, C.c_int32, C.c_float, C.c_float, vp,
                                       C.c_size_t, vp, vp, vp, vp]),
    "pdgs_geom_bytes": (C.c_size_t, [C.c_int64]),
    "pdgs_image_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "pdgs_binning_bytes": (C.c_size_t, [C.c_int64, C.c_int64]),
    "pdgs_backward_scratch_bytes": (C.c_size_t, [C.c_int64]),
    "pdgs_forward": (C.c_int, [vp, C.POINTER(PdgsScene), C.POINTER(PdgsCamera), C.POINTER(PdgsState),
                               C.POINTER(PdgsForwardOut), C.c_int32, C.POINTER(C.c_int64), vp]),
    "pdgs_backward": (C.c_int, [vp, C.POINTER(PdgsScene), C.POINTER(PdgsCamera), C.POINTER(PdgsState), vp,
                                C.c_int64, vp, vp, vp, C.c_size_t, C.POINTER(PdgsGrads), vp]),
    "pdgs_sh_backward_views": (C.c_int, [vp, C.POINTER(PdgsScene), C.c_int32, vp, vp, C.c_int32, vp, vp, C.c_int32, C.c_int64,
                                         C.c_int64, vp]),
    "pdgs_sh_backward_views_split": (C.c_int, [vp, C.POINTER(PdgsScene), C.c_int32, vp, vp, C.c_int32, vp, vp, vp, C.c_int32,
                                               C.c_int64, C.c_int64, vp]),
    "pdgs_sparse_adam": (C.c_int, [vp, C.POINTER(PdgsAdamGroup), C.c_int32, vp, vp, C.c_int64, C.c_double, C.c_double, C.c_double,
                                   C.c_int32, vp]),
    "pdgs_mark_visible": (C.c_int, [vp, vp, vp, C.c_int64, vp, vp]),
    "pdgs_sort_scratch_bytes": (C.c_size_t, [C.c_int64]),
    "pdgs_sort_pairs": (C.c_int, [vp, vp, vp, vp, C.c_int64, C.c_int32, vp, C.c_size_t, vp]),
    "pdgs_sort_pairs_static": (C.c_int, [vp, vp, vp, vp, C.c_int64, C.c_int32, vp]),
    "pdgs_state_view": (C.c_int, [C.POINTER(PdgsState), C.c_int64, C.c_int32, C.c_int32, C.POINTER(PdgsStateView)]),
}

STAGES = ("preprocess", "scan", "duplicate", "sort", "ranges", "blend_fwd", "bwd_zero", "blend_bwd", "preprocess_bwd", "loss",
          "depth_sort", "sh_bwd", "adam")

_lib = None
_lock = threading.Lock()
_handles = {}  # (thread id, device index) -> PdgsHandle*
_timing = {}   # device index -> stage timing enabled (applies to every thread's handle on that device)


class PdgsError(RuntimeError):
    pass


def lib():
    """Load libpdgs.so (once).  Raises if it has not been built: there is nothing to fall back to."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                path = os.path.abspath(LIB_PATH)
                if not os.path.exists(path):
                    raise PdgsError(
                        f"{path} is missing: build it with `python priordepth-gaussiansplatting_b200/build.py` "
                        "(nvcc, sm_100a).  This rasterizer has no CPU or PyTorch fallback.")
                l = C.CDLL(path)
                for name, (res, args) in SYMBOLS.items():
                    fn = getattr(l, name)
                    fn.restype = res
                    fn.argtypes = args
                _lib = l
    return _lib


def last_error() -> str:
    return lib().pdgs_last_error().decode("utf-8", "replace")


def check(rc: int, what: str) -> int:
    if rc < 0:
        raise PdgsError(f"{what} failed ({rc}): {last_error()}")
    return rc


def handle(device_index: int):
    """Per (thread, device) PdgsHandle; the caller must have made `device_index` current."""
    key = (threading.get_ident(), int(device_index))
    h = _handles.get(key)
    if h is None:
        out = vp()
        check(lib().pdgs_create(C.byref(out)), "pdgs_create")
        h = out
        with _lock:
            _handles[key] = h
        if _timing.get(int(device_index)):
            check(lib().pdgs_set_stage_timing(h, 1), "pdgs_set_stage_timing")
    return h


def device_handles(device_index: int):
    """Every thread's handle on `device_index` (autograd runs backward on its own thread)."""
    with _lock:
        return [h for (_, d), h in _handles.items() if d == int(device_index)]
