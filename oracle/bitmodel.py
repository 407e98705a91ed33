"""ctypes wrapper of oracle/bitmodel.c (fp32 bit-model of the integer side).  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "libbitmodel.so")


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "bitmodel.c")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _LIB


_lib = None


def _load():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.bm_preprocess.restype = ctypes.c_int64
        _lib.bm_bin.restype = ctypes.c_int
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _f32(t):
    if t is None:
        return None
    if hasattr(t, "detach"):
        t = t.detach().cpu().numpy()
    return np.ascontiguousarray(t, dtype=np.float32)


def preprocess(means, scales, mod, rots, cov3d, viewmatrix, projmatrix, W, H, tanfovx, tanfovy, antialiasing=False):
    lib = _load()
    means = _f32(means)
    P = means.shape[0]
    scales, rots, cov3d = _f32(scales), _f32(rots), _f32(cov3d)
    vm, pm = _f32(viewmatrix).reshape(-1), _f32(projmatrix).reshape(-1)
    radii = np.zeros(P, np.int32)
    radius_raw = np.zeros(P, np.int32)
    rect = np.zeros((P, 4), np.int32)
    depth_bits = np.zeros(P, np.uint32)
    xy = np.zeros((P, 2), np.float32)
    tiles = np.zeros(P, np.uint32)
    total = lib.bm_preprocess(ctypes.c_int64(P), _p(means), _p(scales), ctypes.c_float(mod), _p(rots), _p(cov3d),
                              _p(vm), _p(pm), ctypes.c_int(W), ctypes.c_int(H), ctypes.c_float(tanfovx),
                              ctypes.c_float(tanfovy), ctypes.c_int(int(antialiasing)), _p(radii), _p(radius_raw),
                              _p(rect), _p(depth_bits), _p(xy), _p(tiles))
    return dict(radii=radii, radius_raw=radius_raw, rect=rect, depth_bits=depth_bits, xy=xy, tiles=tiles,
                num_rendered=int(total))


def bin_and_sort(pre: dict, W: int, H: int):
    lib = _load()
    gx, gy = (W + 15) // 16, (H + 15) // 16
    R = pre["num_rendered"]
    keys = np.zeros(max(R, 1), np.uint64)
    vals = np.zeros(max(R, 1), np.uint32)
    ranges = np.zeros((gx * gy, 2), np.uint32)
    P = pre["radii"].shape[0]
    rc = lib.bm_bin(ctypes.c_int64(P), _p(pre["radii"]), _p(pre["rect"]), _p(pre["depth_bits"]), ctypes.c_int(gx),
                    ctypes.c_int(gy), ctypes.c_int64(R), _p(keys), _p(vals), _p(ranges))
    if rc != 0:
        raise RuntimeError(f"bm_bin failed: {rc}")
    return dict(keys=keys[:R], point_list=vals[:R], ranges=ranges)
