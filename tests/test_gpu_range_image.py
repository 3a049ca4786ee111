"""GPU: the range-image front end (C-ABI rimg_*, splat-loam_b200/range_image.py) against the
numpy restatement oracle/range_image.py and against the round trip through the reference's own
back-projected cloud (tests/golden/range_image_roundtrip.npz).  PARITY UNPINNED for the two
pyprojections calls - see csrc/range_image.cu."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import range_image as ri
from test_oracle_range_image import _spinning_scan

pytestmark = pytest.mark.gpu


def _product(cloud, H, W, dmin=0.5, dmax=80.0, K=None):
    from range_image import RangeImagePreprocessor
    out = RangeImagePreprocessor(H, W, dmin, dmax)(cloud, K=K)
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in out.items()}


def _compare(cloud, H, W, dmin=0.5, dmax=80.0):
    got = _product(cloud, H, W, dmin, dmax)
    K = ri.spherical_intrinsics(cloud, H, W)
    # atan2f differs from numpy's arctan2 in the last place: the extrema, hence K, to ~1e-6
    assert np.abs(got["K"] - K).max() <= 2e-6 * np.abs(K).max()
    # the projection itself is compared under the product's own K
    lut, rng, nrm, valid = ri.project(cloud, got["K"], H, W, dmin, dmax)
    same = got["lut"] == lut
    # (a point whose image coordinate is within float rounding of an integer may fall either side)
    assert same.mean() > 0.999, same.mean()
    assert np.array_equal(got["image_valid"][0][same], valid[0][same])
    assert np.array_equal(got["image_depth"][0][same], rng[0][same])          # bit-exact ranges
    assert np.abs(got["image_normal"][:, same] - nrm[:, same]).max() <= 1e-6
    return got


def test_spinning_scan_matches_the_oracle(native_lib):
    got = _compare(_spinning_scan(64, 1024), 64, 1024)
    v = got["image_valid"][0]
    assert v[1:, 1:].mean() > 0.95 and not v[0].any() and not v[:, 0].any()


def test_dense_cloud_with_collisions(native_lib):
    """Many points per pixel: the nearest wins, equal ranges go to the smaller index."""
    rng = np.random.default_rng(5)
    d = rng.standard_normal((400000, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:, 2] *= 0.3
    cloud = (d * rng.uniform(0.3, 90.0, (len(d), 1))).astype(np.float32)
    cloud[1000:2000] = cloud[:1000]                      # exact duplicates: ties on the range
    got = _compare(cloud, 32, 256)
    lut = got["lut"]
    assert not np.isin(lut, np.arange(1000, 2000)).any()


def test_caller_supplied_K_and_round_trip_of_the_reference_cloud(native_lib):
    g = np.load(os.path.join(GOLDEN, "range_image_roundtrip.npz"))
    got = _product(g["cloud"], 64, 1024, K=g["K"])
    row, col = g["row"].astype(int), g["col"].astype(int)
    assert np.array_equal(got["K"], g["K"])
    assert np.array_equal(got["lut"][row, col], np.arange(len(row)))
    assert got["image_valid"].sum() == len(row)
    assert np.abs(got["image_depth"][0][row, col] - g["depth"]).max() < 1e-5


def test_empty_and_out_of_range(native_lib):
    from range_image import RangeImagePreprocessor
    pre = RangeImagePreprocessor(8, 16, 0.5, 80.0)
    with pytest.raises(RuntimeError):
        pre(np.zeros((0, 3), np.float32))
    K = np.array([[-2.0, 0, 8], [0, -4.0, 4], [0, 0, 1]], np.float32)
    out = pre(np.zeros((0, 3), np.float32), K=K)
    assert not out["image_valid"].any() and bool((out["lut"] == -1).all())
    far = np.array([[100.0, 0, 0], [0.1, 0, 0], [np.nan, 1, 1]], np.float32)
    out = pre(far, K=K)
    assert not out["image_valid"].any() and not out["image_depth"].any()
    with pytest.raises(RuntimeError):
        pre(np.zeros((4, 2), np.float32))


def test_feeds_the_camera_and_the_rasterizer_shapes(native_lib):
    """The outputs have the shapes and dtypes scene/cameras.py:10-50 takes."""
    got = _product(_spinning_scan(16, 128, seed=2), 16, 128)
    assert got["K"].shape == (3, 3) and got["image_depth"].shape == (1, 16, 128)
    assert got["image_normal"].shape == (3, 16, 128) and got["image_valid"].shape == (1, 16, 128)
    assert got["image_valid"].dtype == np.bool_ and got["lut"].dtype == np.int64
