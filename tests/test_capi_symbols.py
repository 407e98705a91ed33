"""CPU tests: libpdgs.so builds/loads, exports every symbol include/pdgs.h declares, sizing functions
behave, and the host-side mirror of the reference API validates arguments like the reference does.
No compute call is made here (there is no GPU in the authoring container)."""
import ctypes
import os
import re
import subprocess

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pdgs.h")


@pytest.fixture(scope="module")
def lib():
    import sys
    sys.path.insert(0, os.path.join(ROOT, "priordepth-gaussiansplatting_b200"))
    import build as pdgs_build
    pdgs_build.build()
    from pdgs_b200 import _lib
    return _lib


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pdgs_[a-z_]+)\s*\(", src)))


def test_header_and_binding_declare_the_same_symbols(lib):
    assert _declared_symbols() == sorted(lib.SYMBOLS.keys())


def test_library_exports_every_declared_symbol(lib):
    l = lib.lib()
    for name in _declared_symbols():
        assert hasattr(l, name), name
    out = subprocess.run(["nm", "-D", "--defined-only", os.path.abspath(lib.LIB_PATH)], capture_output=True, text=True)
    exported = set(re.findall(r" T (pdgs_[a-z_]+)", out.stdout))
    assert set(_declared_symbols()) <= exported


def test_library_is_sm100a_only(lib):
    out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "--list-elf", os.path.abspath(lib.LIB_PATH)],
                         capture_output=True, text=True)
    archs = set(re.findall(r"sm_(\d+a?)", out.stdout))
    assert archs == {"100a"}, out.stdout


def test_version_and_sizes(lib):
    l = lib.lib()
    assert b"sm_100a" in l.pdgs_version()
    g1, g2 = l.pdgs_geom_bytes(1000), l.pdgs_geom_bytes(2000)
    assert 0 < g1 < g2 and g1 % 128 == 0
    assert l.pdgs_geom_bytes(0) > 0
    assert l.pdgs_image_bytes(256, 256) >= 256 * 256 * 8
    b = l.pdgs_binning_bytes(1 << 20, 1000)
    assert b >= (1 << 20) * 16  # two (u32 tile id, u32 index) pair buffers
    assert l.pdgs_backward_scratch_bytes(1000) >= 1000 * 48
    assert l.pdgs_sort_scratch_bytes(1 << 20) > 0


def test_no_device_fails_loudly(lib):
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    h = ctypes.c_void_p()
    rc = lib.lib().pdgs_create(ctypes.byref(h))
    assert rc == lib.PDGS_E_NO_DEVICE
    assert b"no CPU path" in lib.lib().pdgs_last_error()


def test_python_api_surface_and_argument_validation(lib):
    import diff_gaussian_rasterization as dgr
    assert dgr.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug", "antialiasing")
    for name in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):
        assert callable(getattr(dgr._C, name))
    # the accelerated-rasterizer surface the reference probes for (train.py:47-52); fusedssim lives in fused_ssim
    assert issubclass(dgr.SparseGaussianAdam, torch.optim.Adam) and not hasattr(dgr._C, "fusedssim")
    import inspect
    assert {"dc", "raw_params"} <= set(inspect.signature(dgr.GaussianRasterizer.forward).parameters)
    rs = dgr.GaussianRasterizationSettings(16, 16, 0.5, 0.5, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 3,
                                           torch.zeros(3), False, False, False)
    r = dgr.GaussianRasterizer(raster_settings=rs)
    assert isinstance(r, torch.nn.Module)
    P = 4
    m, m2, op = torch.zeros(P, 3), torch.zeros(P, 3), torch.zeros(P, 1)
    sh, col = torch.zeros(P, 16, 3), torch.zeros(P, 3)
    sc, rot, cov = torch.ones(P, 3), torch.zeros(P, 4), torch.zeros(P, 6)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=m, means2D=m2, opacities=op, scales=sc, rotations=rot)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=m, means2D=m2, opacities=op, shs=sh, colors_precomp=col, scales=sc, rotations=rot)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=m, means2D=m2, opacities=op, shs=sh)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=m, means2D=m2, opacities=op, shs=sh, scales=sc, rotations=rot, cov3D_precomp=cov)
    # CPU tensors never silently fall back: the product path refuses them
    with pytest.raises(Exception, match="no CPU path"):
        r(means3D=m, means2D=m2, opacities=op, shs=sh, scales=sc, rotations=rot)
    with pytest.raises(Exception, match="no CPU path"):
        r.markVisible(m)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "priordepth-gaussiansplatting_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f
                assert "raster_oracle" not in src and "bitmodel" not in src, f


def test_prefiltered_is_refused_at_the_c_abi(lib):
    """include/pdgs.h (PdgsCamera.prefiltered): pdgs_forward / pdgs_backward return PDGS_E_INVALID for 1.  The flag is
    checked on the host before any CUDA call, so dummy non-NULL pointers are enough and no GPU is needed."""
    l = lib.lib()
    d = ctypes.c_void_p(256)
    cam = lib.PdgsCamera(image_height=16, image_width=16, tanfovx=0.5, tanfovy=0.5, bg=d, scale_modifier=1.0,
                         viewmatrix=d, projmatrix=d, sh_degree=0, campos=d, prefiltered=1, debug=0, antialiasing=0)
    sc = lib.PdgsScene(P=0, M=0)
    st, out, grads = lib.PdgsState(), lib.PdgsForwardOut(), lib.PdgsGrads()
    R = ctypes.c_int64(0)
    rc = l.pdgs_forward(d, ctypes.byref(sc), ctypes.byref(cam), ctypes.byref(st), ctypes.byref(out), 0, ctypes.byref(R), None)
    assert rc == lib.PDGS_E_INVALID and "prefiltered" in lib.last_error()
    rc = l.pdgs_backward(d, ctypes.byref(sc), ctypes.byref(cam), ctypes.byref(st), d, 0, d, None, d, 0, ctypes.byref(grads), None)
    assert rc == lib.PDGS_E_INVALID and "prefiltered" in lib.last_error()
    cam.prefiltered = 0  # the same call gets past the flag (and then stops at the missing output buffers)
    rc = l.pdgs_forward(d, ctypes.byref(sc), ctypes.byref(cam), ctypes.byref(st), ctypes.byref(out), 0, ctypes.byref(R), None)
    assert rc == lib.PDGS_E_INVALID and "prefiltered" not in lib.last_error()


def test_ctypes_structs_match_the_header_layout(lib, tmp_path):
    """Sizes and field offsets of pdgs_b200/_lib.py's ctypes structures against what a C compiler makes of
    include/pdgs.h (a struct that is short or shifted makes the library read past its end)."""
    structs = {"PdgsCamera": lib.PdgsCamera, "PdgsScene": lib.PdgsScene, "PdgsState": lib.PdgsState,
               "PdgsForwardOut": lib.PdgsForwardOut, "PdgsGrads": lib.PdgsGrads, "PdgsAdamGroup": lib.PdgsAdamGroup,
               "PdgsStateView": lib.PdgsStateView}
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', 'int main(void) {']
    for name, cls in structs.items():
        lines.append(f'  printf("{name} size %zu\\n", sizeof({name}));')
        for field, _ in cls._fields_:
            lines.append(f'  printf("{name} {field} %zu\\n", offsetof({name}, {field}));')
    lines += ['  return 0;', '}']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c11", "-o", str(exe), str(src)], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout
    seen = 0
    for line in out.splitlines():
        name, field, value = line.split()
        cls = structs[name]
        got = ctypes.sizeof(cls) if field == "size" else getattr(cls, field).offset
        assert got == int(value), (name, field, got, value)
        seen += 1
    assert seen == sum(len(c._fields_) + 1 for c in structs.values())
    # the header's field lists and the binding's agree in count too (offsetof above only sees the binding's names)
    hdr = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    for name, cls in structs.items():
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (name, name), hdr, flags=re.S).group(1)
        n_fields = sum(len(decl.split(",")) for decl in body.split(";") if decl.strip())
        assert n_fields == len(cls._fields_), (name, n_fields, len(cls._fields_))
