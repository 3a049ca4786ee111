"""GPU test of sknn_dist2_static, the distCUDA2 entry point for callers that cannot allocate
device memory (a real-time sensor thread, a process whose GPU memory is owned by another
framework's allocator, a plug-in barred from allocating CUDA calls).  It keeps its scratch in
a fixed device array inside the library instead of taking a workspace from the caller.

Called straight through the C-ABI (ctypes, raw device pointers) and compared with the
existing entry point sknn_dist2 on two calls in a row; the outputs must be bit-identical
(the tolerance the other kNN tests use).  The second call uses a different, larger cloud so
that anything stale left in the shared array by the first call would show up.  This file
sorts last on purpose: with `pytest -x` every other test has run before it."""
import ctypes

import numpy as np
import pytest
import torch

import scenes

pytestmark = pytest.mark.gpu

MAX_POINTS = 65536   # SKNN_STATIC_MAX_POINTS in include/splat_b200.h


def _both(lib, pts):
    # load the kernels first (CUDA loads a module lazily at its first launch, which is when
    # the driver places the code and the static array), so that the free-memory comparison
    # below only sees what the call itself does
    warm = torch.zeros((8, 3), device="cuda")
    wout = torch.empty(8, device="cuda")
    assert lib.sknn_dist2_static(8, warm.data_ptr(), wout.data_ptr(), None) == 0
    d_pts = torch.from_numpy(pts).cuda()
    P = pts.shape[0]
    out_static = torch.full((P,), -1.0, device="cuda")
    out_ws = torch.full((P,), -2.0, device="cuda")
    ws = torch.empty(int(lib.sknn_workspace_bytes(P)), dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    rc = lib.sknn_dist2_static(P, d_pts.data_ptr(), out_static.data_ptr(), stream)
    assert rc == 0, lib.splat_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free_before      # the driver handed out nothing
    rc = lib.sknn_dist2(P, d_pts.data_ptr(), out_ws.data_ptr(), ws.data_ptr(), ws.numel(), stream)
    assert rc == 0, lib.splat_last_error()
    torch.cuda.synchronize()
    return out_static.cpu().numpy(), out_ws.cpu().numpy()


def test_static_scratch_entry_point_matches_workspace_entry_point(native_lib):
    lib = native_lib
    first = scenes.knn_cloud(5000, seed=3)
    a, b = _both(lib, first)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    second = scenes.knn_cloud(60000, seed=4)          # second call in a row, other cloud
    a, b = _both(lib, second)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    assert np.isfinite(a).all() and a.min() >= 0
    a, b = _both(lib, first[:7].copy())               # and a tiny one after the large one
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_static_scratch_limit_is_enforced(native_lib):
    lib = native_lib
    P = MAX_POINTS + 1
    pts = torch.zeros((P, 3), device="cuda")
    out = torch.full((P,), 7.0, device="cuda")
    rc = lib.sknn_dist2_static(P, pts.data_ptr(), out.data_ptr(), None)
    assert rc != 0
    assert b"exceed the static scratch limit" in lib.splat_last_error()
    torch.cuda.synchronize()
    assert bool((out == 7.0).all())                   # refused before anything was launched
    # the limit itself is accepted
    pts = torch.from_numpy(scenes.knn_cloud(MAX_POINTS, seed=5)).cuda()
    out = torch.empty(MAX_POINTS, device="cuda")
    assert lib.sknn_dist2_static(MAX_POINTS, pts.data_ptr(), out.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert bool(torch.isfinite(out).all())
    assert lib.sknn_dist2_static(0, None, None, None) == 0
