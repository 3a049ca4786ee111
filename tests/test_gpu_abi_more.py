"""GPU: corner cases of the C-ABI that a binding other than ours can hit - gradient outputs
placed anywhere in memory, a too-small capacity hint on an image with more than 1024 tiles,
one host thread driving two devices, a peer that never arrives at the gradient exchange, and
the launch counter."""
import ctypes
import os

import numpy as np
import pytest
import torch

import scenes
from conftest import cosine, rel_l2
from gpu_helpers import bits, run_product, to_cuda

pytestmark = pytest.mark.gpu


def _forward(c, s, hint=None):
    from diff_surfel_spherical_rasterization import _C
    empty = torch.empty(0, device=c["means3D"].device)
    if hint is not None:
        _C._hints[(c["means3D"].device.index, s["H"], s["W"])] = hint
    return _C.rasterize_gaussians(c["means3D"], c["opacities"], c["scales"], c["rotations"], 1.0,
                                  empty, c["viewmatrix"], c["projmatrix"], s["H"], s["W"], False,
                                  False)


@pytest.mark.parametrize("layout", ["scattered", "reversed_tight", "misaligned"])
def test_gradient_outputs_anywhere_in_memory(native_lib, layout):
    """dssr_backward clears exactly the four accumulation arrays it is given: outputs that a
    caller allocated separately (the seven torch::zeros of the rasterize_points.cu rebinding)
    can have other live data between them, which must survive."""
    import splat_b200_native as native
    s = scenes.make_scene("C0")
    c = to_cuda(s)
    P, H, W = s["means3D"].shape[0], s["H"], s["W"]
    R, out, radii, geom, binning, img = _forward(c, s)
    want = run_product(s, state=())
    # carve the seven outputs out of one arena with sentinel words between them
    sizes = {"means2D": 3 * P, "normal": 3 * P, "opacity": P, "means3D": 3 * P,
             "transMat": 12 * P, "scales": 2 * P, "rotations": 4 * P}
    order = list(sizes)
    gap = {"scattered": 1024, "reversed_tight": 4, "misaligned": 4}[layout]
    if layout != "scattered":
        order = ["transMat", "rotations", "opacity", "normal", "means2D", "scales", "means3D"]
    arena = torch.full((sum(sizes.values()) + gap * 9 + 64,), 12345.0, device="cuda")
    off, view = gap + (1 if layout == "misaligned" else 0), {}
    for k in order:
        # the vector accesses of the per-surfel kernel keep their documented alignment
        al = {"transMat": 4, "rotations": 4, "scales": 2}.get(k, 1)
        off = (off + al - 1) // al * al
        view[k] = (off, sizes[k])
        off += sizes[k] + gap
    arena_ptr = arena.data_ptr()
    ptr = {k: arena_ptr + 4 * o for k, (o, n) in view.items()}
    a = native.DssrBackwardArgs(
        P, W, H, int(R), 1.0, 0, c["means3D"].data_ptr(), c["scales"].data_ptr(),
        c["rotations"].data_ptr(), None, c["viewmatrix"].data_ptr(), c["projmatrix"].data_ptr(),
        radii.data_ptr(), geom.data_ptr(), binning.data_ptr(), img.data_ptr(),
        c["dL_dout"].data_ptr(), ptr["means2D"], ptr["normal"], ptr["opacity"], ptr["means3D"],
        ptr["transMat"], ptr["scales"], ptr["rotations"], None, 0, None,
        torch.cuda.current_stream().cuda_stream)
    native.check(native_lib.dssr_backward(ctypes.byref(a)))
    torch.cuda.synchronize()
    host = arena.cpu().numpy()
    used = np.zeros(host.shape, dtype=bool)
    for k, (o, n) in view.items():
        used[o:o + n] = True
        if k == "normal":
            continue
        got = host[o:o + n]
        ref = want["grad_" + k].ravel()
        assert rel_l2(got, ref) < 1e-5 or not ref.any(), k
    assert np.all(host[~used] == 12345.0)          # nothing outside the seven arrays was touched


@pytest.mark.parametrize("dims", [(144, 2064, 45.0, 60000), (64, 1024, 33.2, 60000)],
                         ids=["global_sort_two_tile_passes", "tile_first_sort"])
def test_small_hint_with_two_tile_passes_on_poisoned_buffers(native_lib, dims):
    """A capacity hint far below the instance count, the buffers the allocator hands out
    pre-filled with 0xFF: the hinted attempt must stay inside its arrays and the exact-size
    re-run give the usual result.  With more than 1024 tiles the tile ranges come from
    identify_ranges_kernel (global sort path); with fewer, from the tile histogram, and the
    per-tile sort and the block lists run over ranges with holes of stale bytes."""
    from diff_surfel_spherical_rasterization import _C
    s = scenes.make_scene(dims)
    c = to_cuda(s)
    _C._hints.clear()
    ref = _forward(c, s)                       # exact-size path
    ref_pl = _C.debug_state(ref[3], ref[4], ref[5], "point_list").numpy()
    ref_rg = _C.debug_state(ref[3], ref[4], ref[5], "ranges").numpy()
    del ref
    for hint in (1000, 20000):
        # poison what the caching allocator will hand out next
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
        junk = [torch.full((n,), 0xFF, dtype=torch.uint8, device="cuda")
                for n in (64 << 20, 16 << 20, 8 << 20, 4 << 20, 2 << 20, 1 << 20)]
        guard = torch.full((8 << 20,), 0x5A, dtype=torch.uint8, device="cuda")
        del junk
        got = _forward(c, s, hint=hint)
        torch.cuda.synchronize()
        assert bool((guard == 0x5A).all())
        assert np.array_equal(_C.debug_state(got[3], got[4], got[5], "point_list").numpy(), ref_pl)
        assert np.array_equal(_C.debug_state(got[3], got[4], got[5], "ranges").numpy(), ref_rg)
        _C._hints.clear()
    r = run_product(s, state=())
    assert np.isfinite(r["out"]).all()


def test_one_host_thread_two_devices(native_lib):
    """The per-thread host state of dssr_forward (count event, pinned word, profiling events)
    is kept per device: the same thread renders on cuda:0, then cuda:1, then cuda:0 again."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import splat_b200_native as native
    s = scenes.make_scene("C1")
    want = run_product(s, state=())
    native.profile_enable(True)
    try:
        for dev in (0, 1, 0, 1):
            with torch.cuda.device(dev):
                c = {k: (torch.from_numpy(v).to(f"cuda:{dev}") if isinstance(v, np.ndarray) else v)
                     for k, v in s.items()}
                R, out, *_ = _forward(c, s)
                torch.cuda.synchronize(dev)
                assert R == want["num_rendered"]
                assert np.array_equal(bits(out.cpu().numpy()), bits(want["out"]))
                assert native.profile_read().get("render_forward", 0.0) > 0.0
    finally:
        native.profile_enable(False)


def test_a_peer_that_never_arrives_poisons_every_rank(native_lib, monkeypatch):
    """Rank 1 never pushes: rank 0's reduce gives up after the (shortened) timeout, marks both
    arenas, and both ranks report the fault - rank 1 as soon as it completes an exchange - and
    refuse further work."""
    import splat_b200_native as native
    monkeypatch.setenv("SPLAT_B200_XCHG_TIMEOUT_MS", "50")
    P, world = 1000, 2
    hs = []
    for r in range(world):
        h = ctypes.c_void_p()
        native.check(native_lib.splat_xchg_create(P, world, r, ctypes.byref(h)))
        hs.append(h)
    try:
        arr = (ctypes.c_void_p * world)(*[h.value for h in hs])
        for h in hs:
            native.check(native_lib.splat_xchg_connect_local(h, arr))
        st = torch.cuda.current_stream().cuda_stream
        buf = torch.ones((P, 10), device="cuda")
        assert native_lib.splat_xchg_fault(hs[0]) == 0
        native.check(native_lib.splat_xchg_push(hs[0], buf.data_ptr(), st))
        native.check(native_lib.splat_xchg_reduce(hs[0], st))
        native.check(native_lib.splat_xchg_wait(hs[0], st))
        torch.cuda.synchronize()
        code = native_lib.splat_xchg_fault(hs[0])
        assert code != 0 and (code - 1) // 16 == 1          # it was rank 1 that did not come
        assert native_lib.splat_xchg_push(hs[0], buf.data_ptr(), st) != 0
        assert b"invalid on every rank" in native_lib.splat_last_error()
        # the late rank goes through its exchange and finds the fault in its own arena
        assert native_lib.splat_xchg_fault(hs[1]) == 0
        native.check(native_lib.splat_xchg_push(hs[1], buf.data_ptr(), st))
        native.check(native_lib.splat_xchg_reduce(hs[1], st))
        native.check(native_lib.splat_xchg_wait(hs[1], st))
        torch.cuda.synchronize()
        assert native_lib.splat_xchg_fault(hs[1]) != 0
        assert native_lib.splat_xchg_push(hs[1], buf.data_ptr(), st) != 0
    finally:
        for h in hs:
            native_lib.splat_xchg_destroy(h)


def test_launch_counter(native_lib):
    import splat_b200_native as native
    s = scenes.make_scene("C0")
    run_product(s, state=())
    n0 = native.launch_count()
    run_product(s, state=())
    per_frame = native.launch_count() - n0
    assert 9 <= per_frame <= 16, per_frame        # 6 forward + 3 backward kernels (tile-first sort)
