"""GPU: the photometric registration (gsa_photo_* through gsaligner_photometric.py) against
the UNMODIFIED reference implementation (oracle/_ref/libref_photo.so: the reference's
photometric_cuda.cu, photometric.cpp, image_set*.c* compiled where they lie), the numpy
oracle, and the golden outputs of the reference.

Tolerances: the prepared reference image (mask, clamped derivatives) bit-exact; the sums of
one linearisation 1e-3 relative on H, b and chi with the good / inlier counts within 0.2 % (a
z-buffer tie or a pixel on a rounding boundary can go either way: the products are rounded in
a different order); the transform after 10 iterations 2e-4 absolute."""
import os

import numpy as np
import pytest
import torch

import scenes
from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _cases():
    pose = scenes.pose(1)
    pose[:3, 3] *= 0.3
    a, K = scenes.room_range_image(32, 256, 40.0)
    b, _ = scenes.room_range_image(32, 256, 40.0, pose=pose, seed=1)
    yield "room", a, b, K
    a, K = scenes.room_range_image(64, 1024, 33.2, seed=2)
    pose = scenes.pose(2)
    pose[:3, 3] *= 0.2
    b, _ = scenes.room_range_image(64, 1024, 33.2, pose=pose, seed=3)
    yield "room_c2", a, b, K
    g = np.load(os.path.join(GOLDEN, "photo_ref.npz"))
    yield "fixture", g["ref"], g["query"], g["K"]


def _ours(ref, query, K, iterations, T0=None):
    from gsaligner_photometric import PhotometricAligner
    H, W = ref.shape
    proj = torch.eye(4)
    proj[:3, :3] = torch.from_numpy(K).T
    al = PhotometricAligner(image_height=H, image_width=W, photo_iterations=iterations,
                            geom_iterations=10)
    al.set_reference(torch.from_numpy(ref)[None].cuda(), None, proj.cuda())
    al.set_query(torch.from_numpy(query)[None].cuda(), None, proj.cuda())
    T, fit, err = al.align_photometric(torch.eye(4) if T0 is None else torch.from_numpy(T0))
    return al, T.numpy(), fit, err


def _close_sums(a, b):
    a, b = np.asarray(a), np.asarray(b)
    scale_H, scale_b = np.abs(b[:21]).max(), np.abs(b[21:27]).max()
    assert np.abs(a[:21] - b[:21]).max() <= 1e-3 * scale_H
    assert np.abs(a[21:27] - b[21:27]).max() <= 1e-3 * scale_b + 1e-9
    assert abs(a[27] - b[27]) <= 1e-3 * b[27]
    assert abs(a[28] - b[28]) <= 2e-3 * b[28] + 2 and abs(a[29] - b[29]) <= 2e-3 * b[29] + 2


@pytest.mark.parametrize("case", ["room", "room_c2", "fixture"])
def test_against_the_reference_build(native_lib, case):
    import reflib
    if not reflib.photo_available():
        pytest.skip("oracle/_ref/libref_photo.so not built")
    name, ref, query, K = next(c for c in _cases() if c[0] == case)
    H, W = ref.shape
    r = reflib.RefPhotometric(H, W)
    r.set_reference(torch.from_numpy(ref).cuda(), K)
    r.set_query(torch.from_numpy(query).cuda())
    T0 = np.eye(4, dtype=np.float32)
    want_img = r.reference_image().numpy()
    want_sums = r.linearize(T0)
    al, T1, _, _ = _ours(ref, query, K, 1)
    got_img = al.debug_reference_image().cpu().numpy()
    assert np.array_equal(got_img[3], want_img[3])                          # mask
    assert np.array_equal(got_img.view(np.uint32), want_img.view(np.uint32))  # every bit
    _close_sums(al.last_system, want_sums)
    assert want_sums[28] > 0.1 * H * W
    for it in (1, 10):
        wT, wfit, werr = r.align(T0, it)
        _, T, fit, err = _ours(ref, query, K, it)
        assert np.abs(T - wT.numpy()).max() < 2e-4, it
        assert abs(fit - wfit) < 3e-3 and abs(err - werr) <= 2e-3 * abs(werr) + 1e-6, it
    # a second alignment on the same object from another initial guess (state is re-armed)
    T0b = scenes.pose(1).astype(np.float32)
    T0b[:3, 3] *= 0.1
    wT, _, _ = r.align(T0b, 5)
    _, T, _, _ = _ours(ref, query, K, 5, T0b)
    assert np.abs(T - wT.numpy()).max() < 2e-4


def test_against_the_numpy_oracle(native_lib):
    from oracle import photometric as ph
    name, ref, query, K = next(_cases())
    T0 = np.eye(4, dtype=np.float32)
    al, T, fit, err = _ours(ref, query, K, 1)
    sums = ph.linearize(ph.prepare_reference(ref), ph.query_cloud(query, K, 0.5, 80.0), T0, K,
                        0.5, 80.0, 5.0, 0.8, 1.0)
    _close_sums(al.last_system, sums)
    d, dc, dr, mask = ph.prepare_reference(ref)
    img = al.debug_reference_image().cpu().numpy()
    assert np.array_equal(img[3] > 0, mask) and np.array_equal(img[1], dc) and np.array_equal(img[2], dr)
    wT, stats, _ = ph.align(ref, query, K, T0, 10)
    _, T, fit, err = _ours(ref, query, K, 10)
    assert np.abs(T - wT).max() < 5e-4
    # and the alignment recovers the motion between the two scans
    pose = scenes.pose(1)
    pose[:3, 3] *= 0.3
    assert np.abs(T - pose).max() < 5e-3


@pytest.mark.parametrize("name", ["ref", "room"])
def test_golden_outputs_of_the_reference(native_lib, name):
    path = os.path.join(GOLDEN, f"photo_{name}.npz")
    g = np.load(path) if os.path.exists(path) else None
    if g is None or "sums0" not in g.files:
        pytest.skip("golden outputs not generated yet")
    al, T1, _, _ = _ours(g["ref"], g["query"], g["K"], 1, g["T0"])
    assert np.array_equal(al.debug_reference_image().cpu().numpy().view(np.uint32),
                          g["ref_image"].view(np.uint32))
    _close_sums(al.last_system, g["sums0"])
    assert np.abs(T1 - g["T1"]).max() < 2e-4
    _, T10, fit, err = _ours(g["ref"], g["query"], g["K"], 10, g["T0"])
    assert np.abs(T10 - g["T10"]).max() < 2e-4
    assert abs(err - g["stats10"][0]) <= 2e-3 * abs(g["stats10"][0]) + 1e-6
    assert abs(fit - g["stats10"][1]) < 3e-3


def test_argument_errors(native_lib):
    from gsaligner_photometric import PhotometricAligner
    al = PhotometricAligner(image_height=16, image_width=64)
    with pytest.raises(RuntimeError, match="CUDA"):
        al.set_reference(torch.zeros(1, 16, 64), None, torch.eye(4))
    with pytest.raises(RuntimeError, match="1xHxW"):
        al.set_reference(torch.zeros(1, 8, 64).cuda(), None, torch.eye(4))
    with pytest.raises(RuntimeError, match="before"):
        al.align_photometric(torch.eye(4))
    # nothing to align: no valid pixel -> pose unchanged, zero inliers
    al.set_reference(torch.zeros(1, 16, 64).cuda(), None, torch.eye(4))
    al.set_query(torch.zeros(1, 16, 64).cuda(), None, torch.eye(4))
    T, fit, err = al.align_photometric(torch.eye(4))
    assert torch.equal(T, torch.eye(4)) and fit == 0.0
