"""CPU: the numpy restatement of the range-image front end (oracle/range_image.py) - a
hand-made case, its own invariants, and the structural facts read off the two artefacts of
this step that the reference ships (tests/golden/range_image_evidence.json, written by
tests/golden/make_range_image_evidence.py).  PARITY UNPINNED: the package that does the
projection in the reference (`pyprojections`) is not in its tree."""
import json
import os

import numpy as np

import scenes
from conftest import GOLDEN
from oracle import range_image as ri


def _spinning_scan(H=64, W=1024, vfov_deg=42.2, seed=0):
    """A spinning-LiDAR scan: H beams x W firings, full turn, ranges from a box room."""
    rng = np.random.default_rng(seed)
    el = np.radians(np.linspace(vfov_deg / 2 - 0.1, -vfov_deg / 2 - 0.1, H))
    az = -np.pi + 2 * np.pi * (np.arange(W) + 0.5) / W
    el, az = np.meshgrid(el, az, indexing="ij")
    d = np.stack([np.cos(el) * np.cos(az), np.cos(el) * np.sin(az), np.sin(el)], -1).reshape(-1, 3)
    half = np.array([12.0, 8.0, 2.5])
    t = np.min(half / np.maximum(np.abs(d), 1e-9), axis=1)
    t = t * (1 + 0.002 * rng.standard_normal(t.shape))
    return (d * t[:, None]).astype(np.float32)


def test_hand_made_case():
    # four points on the +x / +y / -y axes and one above; a 4 x 8 image
    cloud = np.array([[2, 0, 0], [0, 3, 0], [0, -3, 0], [4, 0, 4], [1, 0, 0], [0.1, 0, 0]], np.float32)
    K = ri.spherical_intrinsics(cloud, 4, 8)
    # azimuth spans [-pi/2, pi/2], elevation [0, pi/4]
    assert np.isclose(K[0, 0], -7 / np.pi) and np.isclose(K[1, 1], -3 / (np.pi / 4))
    assert np.isclose(K[0, 2], 4.0) and np.isclose(K[1, 2], 2.0 + 3 / (np.pi / 4) * (np.pi / 8))
    lut, rng, nrm, valid = ri.project(cloud, K, 4, 8, 0.5, 10.0)
    # +x axis -> u = 4, v = 3.5 -> row 4: outside; elevation pi/4 -> v = 0.5 -> row 1, column 5
    assert lut[1, 5] == 3 and np.isclose(rng[0, 1, 5], np.sqrt(32.0))
    # +y -> u = 0.5 -> column 1 (row 4: outside); -y -> u = 7.5 -> column 8: outside
    assert valid.sum() == 1 and (lut >= 0).sum() == 1
    cloud[:, 2] += np.float32(0.3)         # lift everything: rows move inside
    lut, rng, nrm, valid = ri.project(cloud, ri.spherical_intrinsics(cloud, 4, 8), 4, 8, 0.5, 10.0)
    got = {int(i) for i in lut[lut >= 0]}
    assert 4 in got and 5 not in got       # nearest of the three on one ray; 0.1 m is below depth_min
    r, c = np.argwhere(lut == 4)[0]
    assert np.isclose(rng[0, r, c], np.linalg.norm(cloud[4]))
    assert np.allclose(nrm[:, r, c], -cloud[4] / np.linalg.norm(cloud[4]))


def test_invariants_on_a_scan():
    H, W = 64, 1024
    cloud = _spinning_scan(H, W)
    K = ri.spherical_intrinsics(cloud, H, W)
    assert K[0, 0] < 0 and K[1, 1] < 0 and K[2, 2] == 1
    lut, rng, nrm, valid = ri.project(cloud, K, H, W, 0.5, 80.0)
    v = valid[0]
    assert np.array_equal(v, lut >= 0) and np.array_equal(v, rng[0] > 0)
    assert np.array_equal(rng[0][v], np.linalg.norm(cloud, axis=1)[lut[v]])
    assert np.all(nrm[:, ~v] == 0) and np.abs(np.linalg.norm(nrm, axis=0)[v] - 1).max() < 1e-6
    assert len(np.unique(lut[v])) == v.sum()            # a point wins at most one pixel
    # every point inside the image is at least as far as the winner of its pixel
    az, el = ri.angles(cloud)
    u, v = K[0, 0] * az + K[0, 2], K[1, 1] * el + K[1, 2]
    col, row = np.floor(u).astype(int) + 1, np.floor(v).astype(int) + 1
    inside = (col >= 0) & (col < W) & (row >= 0) & (row < H)
    edge = (np.abs(u - np.rint(u)) < 1e-3) | (np.abs(v - np.rint(v)) < 1e-3)
    m = inside & ~edge
    assert np.all(rng[0][row[m], col[m]] > 0)
    assert np.all(np.linalg.norm(cloud, axis=1)[m] >= rng[0][row[m], col[m]])


def test_depth_limits_and_empty():
    cloud = np.array([[0.5, 0, 0], [0.6, 0.1, 0.1], [80, 1, 1], [81, -1, -1]], np.float32)
    K = ri.spherical_intrinsics(cloud, 8, 16)
    lut, rng, nrm, valid = ri.project(cloud, K, 8, 16, 0.5, 80.0)
    kept = {int(i) for i in lut[lut >= 0]}
    assert 0 not in kept and 3 not in kept              # depth_min < r <= depth_max
    lut, rng, nrm, valid = ri.project(np.zeros((0, 3), np.float32), K, 8, 16, 0.5, 80.0)
    assert not valid.any() and np.all(lut == -1) and not rng.any() and not nrm.any()


def test_agrees_with_what_the_reference_ships():
    """The structure of the one camera matrix and the one range image of this step in the
    reference tree: negative focal lengths of the (W-1)/fov form, centre at W/2 for a full
    turn; row 0 and column 0 empty, the other rows filled, the last column filled."""
    ev = json.load(open(os.path.join(GOLDEN, "range_image_evidence.json")))
    H, W = ev["H"], ev["W"]
    cloud = _spinning_scan(H, W, vfov_deg=ev["vfov_from_fy_deg"])
    K = ri.spherical_intrinsics(cloud, H, W)
    Kr = np.array(ev["K_ref"])
    assert abs(K[0, 0] - Kr[0, 0]) < 0.2 and abs(K[1, 1] - Kr[1, 1]) < 0.2
    assert abs(K[0, 2] - Kr[0, 2]) < 0.6 and abs(K[1, 2] - Kr[1, 2]) < 0.6
    lut, rng, nrm, valid = ri.project(cloud, K, H, W, 0.5, 80.0)
    per_row, per_col = valid[0].sum(1), valid[0].sum(0)
    ref_rows = np.array(ev["query_valid_per_row"])
    assert ref_rows[0] == 0 and per_row[0] == 0
    assert ev["query_valid_first_cols"][0] == 0 and per_col[0] == 0
    assert (ref_rows[1:] > 0).all() and (per_row[1:] > 0).all()
    assert ev["query_valid_last_cols"][-1] > 0 and per_col[-1] > 0


def test_round_trip_through_the_reference_back_projection():
    """8192 points of the reference's cloud_query.ply - its own back-projection
    (utils/graphic_utils.py:27-66) of its own range image depth_query.npy - projected with the
    camera matrix its test pairs with them land on the pixels they came from, with the range
    that was there."""
    g = np.load(os.path.join(GOLDEN, "range_image_roundtrip.npz"))
    H, W = 64, 1024
    lut, rng, nrm, valid = ri.project(g["cloud"], g["K"], H, W, 0.5, 80.0)
    row, col = g["row"].astype(int), g["col"].astype(int)
    assert np.array_equal(lut[row, col], np.arange(len(row)))
    assert valid.sum() == len(row)
    assert np.abs(rng[0][row, col] - g["depth"]).max() < 1e-5
