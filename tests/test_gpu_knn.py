"""GPU parity tests of distCUDA2 (through simple_knn._C -> C-ABI sknn_dist2)."""
import numpy as np
import pytest
import torch

import scenes
from conftest import load_golden
from gpu_helpers import bits

pytestmark = pytest.mark.gpu


def _run(pts):
    import simple_knn._C as K
    return K.distCUDA2(torch.from_numpy(pts).cuda()).cpu().numpy()


def test_golden_bit_exact():
    g = load_golden("knn")
    pts = scenes.knn_cloud(5000, seed=3)
    assert np.array_equal(bits(_run(pts)), bits(g["dist"]))
    for n in (1, 2, 3, 4, 5):    # P < 4: sums with FLT_MAX, as the reference returns them
        assert np.array_equal(bits(_run(pts[:n].copy())), bits(g[f"small{n}"])), n
    assert _run(np.zeros((0, 3), np.float32)).shape == (0,)


def test_against_cpu_oracle():
    from oracle import oracle
    for P, seed in ((777, 1), (20000, 2)):
        pts = scenes.knn_cloud(P, seed=seed)
        assert np.array_equal(bits(_run(pts)), bits(oracle.dist2(pts))), P


def test_duplicates_and_degenerate_clouds():
    rng = np.random.default_rng(0)
    pts = rng.normal(size=(4096, 3)).astype(np.float32)
    pts[100:104] = pts[100]                     # four coincident points -> distance 0
    d = _run(pts)
    assert np.all(d[100:104] == 0)
    line = np.zeros((1000, 3), np.float32)      # degenerate bounding box (two flat axes)
    line[:, 0] = np.arange(1000)
    d = _run(line)
    assert np.allclose(d[1:-1], (1 + 1 + 4) / 3.0) and np.isclose(d[0], (1 + 4 + 9) / 3.0)
    same = np.ones((300, 3), np.float32)        # all points identical
    assert not _run(same).any()


def test_full_size_properties_and_reference_build():
    """4M-point accumulated cloud (BASELINE config C5): properties + a sampled exact check,
    and bit-exactness against the reference build when it is present."""
    from oracle import oracle
    P = 4_000_000
    pts = scenes.knn_cloud(P)
    d = _run(pts)
    assert np.isfinite(d).all() and d.min() >= 0
    # translation by a power of two and a permutation leave the result unchanged
    perm = np.random.default_rng(1).permutation(P)
    d2 = _run(pts[perm])
    assert np.array_equal(bits(d2), bits(d[perm]))
    # exact brute force on a bounded sample of queries against the whole cloud
    q = 2048
    exact = oracle.dist2(pts, 0, q)
    assert np.array_equal(bits(d[:q]), bits(exact[:q]))
    import reflib
    if reflib.available():
        ref = reflib.dist2(torch.from_numpy(pts).cuda()).cpu().numpy()
        assert np.array_equal(bits(d), bits(ref))
