"""GPU: further parity cases of the rasterizer against the unmodified reference CUDA build
(oracle/_ref) and behavioural checks of the C-ABI (streams, host threads, prefiltered)."""
import threading

import numpy as np
import pytest
import torch

import scenes
from conftest import assert_n_contrib_equal, cosine, rel_l2
from gpu_helpers import bits, run_product, to_cuda

pytestmark = pytest.mark.gpu
GRADS = ("means3D", "opacity", "scales", "rotations", "transMat")


def _reference(s, scale_modifier=1.0, backward=True):
    import reflib
    if not reflib.available():
        pytest.skip("oracle/_ref/libref_splat.so not built")
    c = to_cuda(s)
    ref = reflib.RefRasterizer()
    R, out, radii = ref.forward(c["means3D"], c["opacities"].reshape(-1), c["scales"], c["rotations"],
                                c["viewmatrix"], c["projmatrix"], s["H"], s["W"],
                                scale_modifier=scale_modifier)
    g = {"num_rendered": R, "out": out.cpu().numpy(), "radii": radii.cpu().numpy()}
    for nm in ("keys", "point_list", "ranges", "n_contrib"):
        g[nm] = ref.state(nm).numpy()
    if backward:
        for k, v in ref.backward(c["dL_dout"], scale_modifier=scale_modifier).items():
            g["grad_" + k] = v.cpu().numpy()
    return g


def _product(s, scale_modifier=1.0):
    from diff_surfel_spherical_rasterization import _C
    c = to_cuda(s)
    empty = torch.empty(0, device="cuda")
    R, out, radii, geom, binning, img = _C.rasterize_gaussians(
        c["means3D"], c["opacities"], c["scales"], c["rotations"], scale_modifier, empty,
        c["viewmatrix"], c["projmatrix"], s["H"], s["W"], False, False)
    r = {"num_rendered": R, "out": out.cpu().numpy(), "radii": radii.cpu().numpy()}
    for nm in ("keys", "point_list", "ranges", "n_contrib"):
        r[nm] = _C.debug_state(geom, binning, img, nm).numpy()
    g = _C.rasterize_gaussians_backward(c["means3D"], radii, c["scales"], c["rotations"],
                                        scale_modifier, empty, c["viewmatrix"], c["projmatrix"],
                                        c["dL_dout"], geom, R, binning, img, False)
    for nm, t in zip(("means2D", "opacity", "means3D", "transMat", "scales", "rotations"), g):
        r["grad_" + nm] = t.cpu().numpy()
    return r


def _compare(r, g, HW):
    assert r["num_rendered"] == g["num_rendered"]
    assert np.array_equal(r["radii"], g["radii"])
    for nm in ("keys", "point_list", "ranges"):
        a, b = (np.ascontiguousarray(x).ravel().view(np.uint8) for x in (r[nm], g[nm]))
        assert np.array_equal(a, b), nm
    assert np.array_equal(bits(r["out"]), bits(g["out"]))
    assert_n_contrib_equal(r["n_contrib"], g["n_contrib"], HW)
    for k in GRADS:
        if "grad_" + k in g:
            assert rel_l2(r["grad_" + k], g["grad_" + k]) < 1e-4, k
            assert cosine(r["grad_" + k], g["grad_" + k]) > 0.9999, k


@pytest.mark.parametrize("scale_modifier", [0.6, 1.7])
def test_scale_modifier_matches_reference(native_lib, scale_modifier):
    """The forward scales the surfels (forward.cu:122-126); the backward does not apply the
    modifier to dL_dscales (quirk Q5) - both as the reference."""
    s = scenes.make_scene("C1")
    _compare(_product(s, scale_modifier), _reference(s, scale_modifier), s["H"] * s["W"])


def test_two_tile_digit_passes_match_reference(native_lib):
    """128x2048 pixels = 1024 tiles is the last single-pass size; 144x2064 needs 11 tile bits,
    i.e. two tile passes, identifyTileRanges and the tile_order kernel."""
    s = scenes.make_scene((144, 2064, 45.0, 60000))
    r, g = _product(s), _reference(s)
    assert ((s["H"] + 15) // 16) * ((s["W"] + 15) // 16) > 1024
    _compare(r, g, s["H"] * s["W"])


def test_non_default_stream_and_interleaving(native_lib):
    """Everything is enqueued on the caller's current stream: two frames on two streams give
    the results of the same frames run alone."""
    frames = [scenes.make_scene("C0", keyframe=k) for k in range(2)]
    alone = [run_product(f, state=()) for f in frames]
    streams = [torch.cuda.Stream() for _ in frames]
    res = [None, None]
    for rep in range(3):
        for i, (f, st) in enumerate(zip(frames, streams)):
            with torch.cuda.stream(st):
                res[i] = run_product(f, state=())
    for a, b in zip(alone, res):
        assert np.array_equal(bits(a["out"]), bits(b["out"]))
        for k in GRADS:
            assert rel_l2(b["grad_" + k], a["grad_" + k]) < 1e-5, k


def test_two_host_threads(native_lib):
    """Per-thread host state (pinned count word, event, last error): two threads render
    different frames concurrently on their own streams."""
    frames = [scenes.make_scene("C0", keyframe=k) for k in range(2)]
    alone = [run_product(f, state=()) for f in frames]
    out, err = [None, None], []

    def work(i):
        try:
            torch.cuda.set_device(0)
            with torch.cuda.stream(torch.cuda.Stream()):
                for _ in range(5):
                    out[i] = run_product(frames[i], state=())
        except Exception as e:   # pragma: no cover
            err.append(e)

    ts = [threading.Thread(target=work, args=(i,)) for i in range(2)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not err, err
    for a, b in zip(alone, out):
        assert np.array_equal(bits(a["out"]), bits(b["out"]))


def test_prefiltered_with_a_culled_surfel_is_an_error(native_lib):
    """The reference __trap()s the device (forward.cu:131-137); here it is a RuntimeError with
    the reference's message and the process lives on."""
    from diff_surfel_spherical_rasterization import _C
    s = scenes.make_scene("C0")
    c = to_cuda(s)
    c["means3D"] = c["means3D"].clone()
    c["means3D"][0] = torch.tensor([1e4, 0.0, 0.0], device="cuda")     # beyond the 120 m range
    empty = torch.empty(0, device="cuda")
    with pytest.raises(RuntimeError, match="filtered although prefiltered"):
        _C.rasterize_gaussians(c["means3D"], c["opacities"], c["scales"], c["rotations"], 1.0, empty,
                               c["viewmatrix"], c["projmatrix"], s["H"], s["W"], True, False)
    r = run_product(s, state=())      # and the next frame is fine
    assert np.isfinite(r["out"]).all()


def _stress_scene(seed, H=48, W=384, vfov=150.0, P=6000):
    """Geometry the conservative sub-tile masks find hardest: grazing tilts up to 89 degrees,
    scales from hair-thin to metres, ranges from 0.3 m, elevations up to the image's +-75
    degrees (footprints that wrap around the vertical axis), opacities down to 1/255."""
    import math
    rng = np.random.default_rng(seed)
    vf = math.radians(vfov)
    az = rng.uniform(-math.pi, math.pi, P)
    el = np.clip(rng.normal(0, vf / 3, P), -vf / 2 - 0.05, vf / 2 + 0.05)
    r = np.exp(rng.uniform(math.log(0.3), math.log(80.0), P))
    d = np.stack([np.cos(az) * np.cos(el), np.sin(az) * np.cos(el), np.sin(el)], 1)
    axis = rng.normal(size=(P, 3))
    axis -= (axis * d).sum(1, keepdims=True) * d
    axis /= np.linalg.norm(axis, axis=1, keepdims=True)
    ang = np.radians(rng.choice([0, 30, 60, 80, 87, 89], P) + rng.uniform(-1, 1, P))
    nrm = -d * np.cos(ang)[:, None] + np.cross(axis, -d) * np.sin(ang)[:, None]
    t = np.cross(nrm, axis)
    t /= np.linalg.norm(t, axis=1, keepdims=True)
    phi = rng.uniform(0, 2 * math.pi, P)
    e0 = t * np.cos(phi)[:, None] + np.cross(nrm, t) * np.sin(phi)[:, None]
    R = np.stack([e0, np.cross(nrm, e0), nrm], 2)
    scales = np.exp(rng.uniform(math.log(0.003), math.log(2.0), (P, 2)))
    opa = rng.choice([1.0 / 255.0, 0.004, 0.05, 0.5, 0.99, 1.0], P) * rng.uniform(0.97, 1.0, P)
    s = {"means3D": (d * r[:, None]).astype(np.float32), "scales": scales.astype(np.float32),
         "rotations": scenes._rotmat_to_quat(R).astype(np.float32),
         "opacities": np.minimum(opa, 1.0).astype(np.float32)[:, None], "H": H, "W": W,
         "viewmatrix": scenes.viewmatrix(seed % 5), "projmatrix": scenes.intrinsics(H, W, vfov),
         "dL_dout": scenes.upstream_grad(H, W, seed)}
    return s


@pytest.mark.parametrize("seed", [1, 2, 3, 4, 5, 6])
def test_stress_geometry_matches_reference(native_lib, seed):
    """Every map, n_contrib, the sort and the gradients against the unmodified reference build:
    a sub-tile mask that dropped a contributing surfel anywhere would change a pixel."""
    s = _stress_scene(seed)
    _compare(_product(s), _reference(s), s["H"] * s["W"])


@pytest.mark.parametrize("P", [300000])
def test_sort_with_16_keys_per_thread_matches_reference(native_lib, P):
    """The sort picks 8, 12 or 16 keys per thread from the expected instance count (C0/C1 run
    with 8, C2 with 12); ~1.05 M instances is the range where 16 fills one wave."""
    s = scenes.make_scene((64, 1024, 33.2, P))
    r, g = _product(s), _reference(s)
    assert 909000 < r["num_rendered"] < 1210000, r["num_rendered"]
    _compare(r, g, s["H"] * s["W"])
    r2 = _product(s)           # second frame: sized from the previous frame's count (hint path)
    _compare(r2, g, s["H"] * s["W"])


def test_tile_ranges_longer_than_shared_memory_match_reference(native_lib):
    """The tile-first sort keeps a tile's range in shared memory up to 12 288 instances and
    sorts longer ranges through global memory: 8 tiles with ~40 000 instances each, plus the
    hinted second frame."""
    s = scenes.make_scene((32, 64, 45.0, 150000))
    r, g = _product(s), _reference(s, backward=False)
    lens = (g["ranges"].reshape(-1, 2).astype(np.int64) @ np.array([-1, 1]))
    assert lens.max() > 12288 and lens.min() > 0, lens
    _compare(r, g, s["H"] * s["W"])
    _compare(_product(s), g, s["H"] * s["W"])
