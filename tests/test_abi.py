"""CPU tests: the C-ABI library builds, loads and exports every symbol that
include/splat_b200.h declares; size queries behave; the drop-in Python packages import
and validate arguments like the reference.  No compute calls (no GPU here)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "splat_b200.h")).read()
    return sorted(set(re.findall(r"SPLAT_API\s+[\w\s\*]+?\b(\w+)\s*\(", text)))


def test_header_symbols_exported(native_lib):
    syms = _declared_symbols()
    assert len(syms) >= 11
    for s in syms:
        assert hasattr(native_lib, s), f"{s} declared in the header but not exported"
    import splat_b200_native as n
    assert sorted(n.EXPORTS) == sorted(s for s in syms)
    assert native_lib.splat_abi_version() == 8


def test_no_cuda_runtime_static_state(native_lib):
    # the library must not need a GPU just to be loaded / queried
    assert native_lib.dssr_geometry_bytes(0) > 0
    assert native_lib.splat_last_error() is not None


def test_size_queries(native_lib):
    g1, g2 = native_lib.dssr_geometry_bytes(1000), native_lib.dssr_geometry_bytes(500000)
    assert g2 > g1 >= 1000 * (64 + 16 + 4 + 8 + 4 + 4)
    assert native_lib.dssr_image_bytes(1024, 64) >= 64 * 1024 * 20 + 256 * 8
    b0 = native_lib.dssr_binning_bytes(0, 1024, 64, 500000, 0)
    b1 = native_lib.dssr_binning_bytes(2000000, 1024, 64, 500000, 0)
    b2 = native_lib.dssr_binning_bytes(2000000, 1024, 64, 500000, 1)
    assert b0 > 0 and b1 >= 2000000 * 28 and b2 >= b1 + 2000000 * 8
    assert native_lib.sknn_workspace_bytes(4000000) >= 4000000 * 32


def test_null_arguments_are_errors_not_crashes(native_lib):
    n = ctypes.c_int(0)
    assert native_lib.dssr_forward(None, ctypes.byref(n)) != 0
    assert b"NULL" in native_lib.splat_last_error()
    assert native_lib.dssr_backward(None) != 0
    assert native_lib.sknn_dist2(-1, None, None, None, 0, None) != 0
    assert native_lib.sknn_dist2(0, None, None, None, 0, None) == 0   # empty cloud is legal


def test_python_surface(native_lib):
    import torch
    import diff_surfel_spherical_rasterization as d
    import simple_knn._C as k
    assert d.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "scale_modifier", "viewmatrix", "projmatrix",
        "prefiltered", "debug")
    for fn in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):
        assert callable(getattr(d._C, fn))
    assert callable(k.distCUDA2)
    rs = d.GaussianRasterizationSettings(16, 32, 1.0, torch.eye(4), torch.eye(4), False, False)
    r = d.GaussianRasterizer(rs)
    x = torch.zeros(4, 3)
    # exactly one of scale/rotation or the precomputed matrix (reference __init__.py:232-240)
    with pytest.raises(Exception, match="exactly one"):
        r(x, x, torch.zeros(4, 1))
    with pytest.raises(Exception, match="exactly one"):
        r(x, x, torch.zeros(4, 1), scales=torch.zeros(4, 2), rotations=torch.zeros(4, 4),
          cov3D_precomp=torch.zeros(4, 12))
    # CPU tensors are rejected, there is no CPU fallback (rasterize_points.cu:21-22)
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        d._C.rasterize_gaussians(x, x[:, :1], x[:, :2], torch.zeros(4, 4), 1.0, torch.zeros(0),
                                 torch.eye(4), torch.eye(4), 16, 32, False, False)
    with pytest.raises(RuntimeError, match="num_points, 3"):
        d._C.rasterize_gaussians(torch.zeros(4, 2), x[:, :1], x[:, :2], torch.zeros(4, 4), 1.0,
                                 torch.zeros(0), torch.eye(4), torch.eye(4), 16, 32, False, False)
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        k.distCUDA2(x)


def test_missing_library_fails_loudly(native_lib, tmp_path, monkeypatch):
    import splat_b200_native as n
    monkeypatch.setattr(n, "_lib", None)
    monkeypatch.setattr(n, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(ImportError, match="no CPU or PyTorch fallback"):
        n.load()


def test_header_is_plain_c_and_cpp():
    """The drop-in boundary must be bindable from C (cgo / ctypes / JNI style FFI) as well as
    C++: the header compiles as C99 and as C++11 with warnings as errors."""
    import shutil
    import subprocess
    hdr = os.path.join(ROOT, "include", "splat_b200.h")
    for cc, lang, std in (("gcc", "c", "-std=c99"), ("g++", "c++", "-std=c++11")):
        if shutil.which(cc) is None:
            pytest.skip(f"{cc} not installed")
        p = subprocess.run([cc, std, "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only",
                            "-x", lang, hdr], capture_output=True, text=True)
        assert p.returncode == 0, p.stderr


def test_widened_surfaces_reject_cpu_tensors(native_lib):
    """gaussian_renderer / mapping_loss / fused_adam have no CPU path either: CPU tensors are an
    error, not a silent fallback."""
    import torch
    import gaussian_renderer as gr
    import mapping_loss as ml
    from fused_adam import activation_adam_step
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        gr.postprocess(torch.zeros(7, 4, 4), torch.eye(4), torch.eye(4), 0.0)

    class Cam:
        world_view_transform = torch.eye(4)
        projection_matrix = torch.eye(4)
        image_depth = torch.zeros(1, 4, 4)
        image_valid = torch.ones(1, 4, 4, dtype=torch.uint8)
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        ml.mapping_image_losses(torch.zeros(7, 4, 4), Cam)
    p = torch.nn.Parameter(torch.zeros(4, 3))
    opt = torch.optim.Adam([{"params": [p], "name": "xyz"}], lr=1e-3)
    with pytest.raises(RuntimeError, match="CUDA"):
        activation_adam_step(opt, torch.zeros(4, 10))
