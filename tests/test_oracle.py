"""CPU tests: the plain-C oracle (oracle/oracle.c) against the golden vectors produced by
the unmodified reference CUDA build (tests/golden/*.npz, see make_golden.py).

Tolerances: the oracle evaluates sin/cos/exp/atan2 with libm where the reference uses
__sincosf and libdevice, so the agreement is ~1e-5 relative, not bit-exact.  Integer
results (radii, tile counts, sort order, ranges, n_contrib) are compared exactly where the
floating-point inputs to the discrete decision are not borderline."""
import numpy as np
import pytest

import scenes
from conftest import assert_n_contrib_equal, cosine, load_golden, rel_l2
from oracle import oracle


def _fwd(s, precomp=None):
    return oracle.rasterize_forward(s["means3D"], None if precomp is not None else s["scales"],
                                    None if precomp is not None else s["rotations"],
                                    s["opacities"], s["viewmatrix"], s["projmatrix"], s["H"],
                                    s["W"], precomp=precomp)


def _check_inputs(s, g):
    assert bytes(g["digest"]).decode() == scenes.input_digest(s), "scene generator drifted"


@pytest.mark.parametrize("name,scene", [("c0", lambda: scenes.make_scene("C0")),
                                        ("edge", scenes.edge_scene)])
def test_forward_backward_vs_reference(name, scene):
    s, g = scene(), load_golden(name)
    _check_inputs(s, g)
    f = _fwd(s)
    P, H, W = f["P"], s["H"], s["W"]
    assert f["num_rendered"] == int(g["num_rendered"])
    assert np.array_equal(f["radii"], g["radii"])
    assert np.array_equal(f["tiles_touched"], g["tiles_touched"].view(np.uint32))
    vis = g["radii"] > 0
    assert rel_l2(f["depths"][vis], g["depths"][vis]) < 1e-6
    assert rel_l2(f["transMat"][vis], g["transMat"].reshape(P, 12)[vis]) < 1e-5
    assert np.abs(f["means2D"][vis] - g["means2D"].reshape(P, 2)[vis]).max() < 1e-3
    # sort order and ranges: identical (depth ties at the ulp level would show up here)
    assert np.array_equal(f["point_list"], g["point_list"].view(np.uint32))
    assert np.array_equal(f["ranges"].ravel(), g["ranges"].view(np.uint32))
    assert np.array_equal(f["keys"] >> np.uint64(32), g["keys"].view(np.uint64) >> np.uint64(32))
    for c in range(7):
        scale = np.abs(g["out"][c]).max() + 1e-30
        assert np.abs(f["out"][c] - g["out"][c]).max() <= 1e-4 * scale, f"map channel {c}"
    assert_n_contrib_equal(f["n_contrib"], g["n_contrib"], H * W)
    gr = oracle.rasterize_backward(f, s["dL_dout"])
    for k in ("opacity", "means3D", "transMat", "scales", "rotations", "normal"):
        ref = g["grad_" + k].reshape(gr[k].shape)
        assert rel_l2(gr[k], ref) < 1e-4, k
        assert cosine(gr[k], ref) > 0.9999, k
    assert not gr["means2D"].any() and not g["grad_means2D"].any()
    # transMat entries 3, 7, 11 are never written (backward.cu:269-277)
    assert not gr["transMat"][:, [3, 7, 11]].any()


def test_precomputed_transmat_path():
    s, g = scenes.precomp_scene(), load_golden("c0_precomp")
    _check_inputs(s, g)
    f = _fwd(s, precomp=g["precomp"])
    assert f["num_rendered"] == int(g["num_rendered"])
    assert np.array_equal(f["radii"], g["radii"])
    for c in range(7):
        scale = np.abs(g["out"][c]).max() + 1e-30
        assert np.abs(f["out"][c] - g["out"][c]).max() <= 1e-4 * scale


def test_reference_test_scene():
    """DSSR/tests/generate_input.py: the one fixture the reference ships.  Opacity 1.0
    everywhere puts the T*(1-alpha) < 1e-4 test on its rounding edge, so n_contrib is not
    compared and the (cancellation-dominated) distortion channel gets an absolute bound."""
    s, g = scenes.reference_test_scene(), load_golden("g0")
    _check_inputs(s, g)
    f = _fwd(s)
    assert f["num_rendered"] == int(g["num_rendered"])
    assert np.array_equal(f["radii"], g["radii"])
    assert np.array_equal(f["point_list"], g["point_list"].view(np.uint32))
    assert np.array_equal(f["ranges"].ravel(), g["ranges"].view(np.uint32))
    sub = f["out"][:, ::8, ::8]
    for c in range(6):
        scale = np.abs(g["out_sub"][c]).max() + 1e-30
        assert np.abs(sub[c] - g["out_sub"][c]).max() <= 1e-4 * scale, c
    dmax = float(g["out_sub"][5].max())
    assert np.abs(sub[6] - g["out_sub"][6]).max() <= 1e-5 * dmax * dmax
    assert rel_l2(f["out"].astype(np.float64).sum(axis=(1, 2))[:6], g["out_sum"][:6]) < 1e-5
    gr = oracle.rasterize_backward(f, s["dL_dout"])
    for k in ("opacity", "means3D", "scales", "rotations"):
        assert cosine(gr[k], g["grad_" + k]) > 0.9999, k


def test_knn_vs_reference():
    g = load_golden("knn")
    pts = scenes.knn_cloud(5000, seed=3)
    d = oracle.dist2(pts)
    assert np.array_equal(d.view(np.uint32), g["dist"].view(np.uint32))   # bit-exact
    for n in (1, 2, 3, 4, 5):   # FLT_MAX sums for P < 4 (simple_knn.cu:155,183)
        a, b = oracle.dist2(pts[:n].copy()), g[f"small{n}"]
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), n


def test_oracle_invariants():
    s = scenes.make_scene("C0", seed=11)
    f = _fwd(s)
    N = f["num_rendered"]
    mask = (np.uint64(1) << np.uint64(32 + 5)) - np.uint64(1)   # 32 tiles -> 5+... bits
    k = f["keys"]
    assert np.all(k[1:] >= k[:-1])                      # sorted by (tile, depth bits)
    r = f["ranges"].astype(np.int64)
    nonempty = r[:, 1] > r[:, 0]
    assert (r[nonempty, 1] - r[nonempty, 0]).sum() == N  # ranges partition the list
    assert sorted(f["keys_unsorted"].tolist()) == k.tolist()
    assert np.array_equal(f["point_offsets"][-1:], [N])
    a = f["out"][1]
    assert a.min() >= 0 and a.max() <= 1.0
    assert np.allclose(f["final_T"][0], 1 - a, atol=1e-7)
    # empty input
    e = oracle.rasterize_forward(np.zeros((0, 3)), np.zeros((0, 2)), np.zeros((0, 4)),
                                 np.zeros((0, 1)), s["viewmatrix"], s["projmatrix"], 16, 32)
    assert e["num_rendered"] == 0 and not e["out"].any()
    vis = oracle.mark_visible(s["means3D"], s["viewmatrix"])
    d = np.linalg.norm(s["means3D"], axis=1)   # identity pose for keyframe 0
    assert np.array_equal(vis, (d > 0.1) & (d <= 120.0))
