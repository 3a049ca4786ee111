import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "splat-loam_b200"), os.path.join(ROOT, "tests"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def rel_l2(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30))


def cosine(a, b):
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    return float(a @ b / (np.linalg.norm(a) * np.linalg.norm(b) + 1e-30))


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


@pytest.fixture(scope="session")
def native_lib():
    """The C-ABI library, built on demand (nvcc cross-compiles without a GPU)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "splat_build", os.path.join(ROOT, "splat-loam_b200", "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.build()
    import splat_b200_native
    return splat_b200_native.load()


def assert_n_contrib_equal(a, b, HW):
    """n_contrib = [last_contributor | median_contributor].  The median entry of a pixel that
    blended nothing is the reference's float -1 pushed through a float->uint32 conversion,
    whose result is not defined (the reference build leaves 0 or stale register contents);
    the backward never reads it there (backward.cu:132,212).  Compare it only where the pixel
    blended something."""
    a = np.asarray(a).reshape(2, HW).view(np.uint32)
    b = np.asarray(b).reshape(2, HW).view(np.uint32)
    assert np.array_equal(a[0], b[0])
    m = b[0] > 0
    assert np.array_equal(a[1][m], b[1][m])
