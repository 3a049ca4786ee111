"""Keyframe-window gradient exchange over peer memory (C-ABI splat_xchg_*), on ONE GPU:
several ranks' exchanges live in this process and are wired with splat_xchg_connect_local,
so the protocol (push -> reduce+gather -> wait, flags, epochs) runs exactly as it does
between GPUs, only the "peer" stores land in the same device's memory.  All of one step's
kernels are issued before the next step's (push for every rank, then reduce for every rank,
then wait for every rank), so no kernel ever waits for a later launch.  The real
multi-process / multi-GPU run is tests/test_gpu_exchange_mp.py."""
import ctypes

import numpy as np
import pytest
import torch

import scenes
from gpu_helpers import to_cuda

pytestmark = pytest.mark.gpu


class LocalWindow:
    def __init__(self, lib, native, P, world):
        self.lib, self.n, self.P, self.world = lib, native, P, world
        self.h = []
        for r in range(world):
            h = ctypes.c_void_p()
            native.check(lib.splat_xchg_create(P, world, r, ctypes.byref(h)))
            self.h.append(h)
        arr = (ctypes.c_void_p * world)(*[h.value for h in self.h])
        for h in self.h:
            native.check(lib.splat_xchg_connect_local(h, arr))
        self.stream = torch.cuda.current_stream().cuda_stream

    def result(self, r):
        from keyframe_window import _DeviceArray
        ptr = self.lib.splat_xchg_result(self.h[r])
        return torch.as_tensor(_DeviceArray(ptr, (self.P, 10), self), device="cuda")

    def finish(self):
        for h in self.h:
            self.n.check(self.lib.splat_xchg_reduce(h, self.stream))
        for h in self.h:
            self.n.check(self.lib.splat_xchg_wait(h, self.stream))
        torch.cuda.synchronize()
        return [self.result(r).clone() for r in range(self.world)]

    def close(self):
        for h in self.h:
            self.lib.splat_xchg_destroy(h)


def ordered_sum(bufs):
    acc = bufs[0].clone()
    for b in bufs[1:]:
        acc += b
    return acc


@pytest.mark.parametrize("world,P", [(1, 777), (2, 1000), (3, 4097), (8, 50001), (16, 333)])
def test_push_reduce_gather_is_the_rank_ordered_sum(native_lib, world, P):
    import splat_b200_native as native
    w = LocalWindow(native_lib, native, P, world)
    try:
        g = torch.Generator(device="cuda").manual_seed(world * 1000 + P)
        for step in range(3):      # epochs: buffers and flags are reused
            bufs = [torch.randn((P, 10), device="cuda", generator=g) for _ in range(world)]
            for r in range(world):
                src = None if (step == 1 and r == world - 1 and world > 1) else bufs[r]
                native.check(native_lib.splat_xchg_push(
                    w.h[r], src.data_ptr() if src is not None else None, w.stream))
                if src is None:
                    bufs[r] = torch.zeros_like(bufs[r])
            res = w.finish()
            want = ordered_sum(bufs)
            for r in range(world):
                assert torch.equal(res[r], want), f"rank {r} step {step}"   # bit-identical
    finally:
        w.close()


def test_fused_push_from_the_vjp_kernel(native_lib):
    """Two ranks render two keyframes of C0; the per-surfel VJP kernel pushes its rows.  The
    exchange result equals the sum of the packed gradients computed without an exchange, and
    the second local keyframe of a rank is accumulated before the push."""
    import splat_b200_native as native
    from diff_surfel_spherical_rasterization import _C
    frames = [to_cuda(scenes.make_scene("C0", keyframe=k)) for k in range(3)]
    P, H, W = frames[0]["means3D"].shape[0], frames[0]["H"], frames[0]["W"]
    empty = torch.empty(0, device="cuda")

    def backward(c, packed, accumulate, xchg):
        R, out, radii, geom, binning, img = _C.rasterize_gaussians(
            c["means3D"], c["opacities"], c["scales"], c["rotations"], 1.0, empty,
            c["viewmatrix"], c["projmatrix"], H, W, False, False)
        _C.rasterize_gaussians_backward(
            c["means3D"], radii, c["scales"], c["rotations"], 1.0, empty, c["viewmatrix"],
            c["projmatrix"], c["dL_dout"], geom, R, binning, img, False, packed=packed,
            packed_accumulate=accumulate, xchg=xchg)

    plain = []
    for c in frames:
        b = torch.empty((P, 10), device="cuda")
        backward(c, b, False, None)
        plain.append(b)
    w = LocalWindow(native_lib, native, P, 2)
    try:
        # rank 0 owns keyframes 0 and 2, rank 1 owns keyframe 1
        b0 = torch.empty((P, 10), device="cuda")
        backward(frames[0], b0, False, None)
        backward(frames[2], b0, True, w.h[0].value)      # accumulates, then pushes the sum
        backward(frames[1], None, False, w.h[1].value)   # no packed buffer at all: push only
        res = w.finish()
        # atomics make a frame's gradients vary in the last bits from run to run
        want = (plain[0] + plain[2]) + plain[1]
        for r in range(2):
            assert torch.equal(res[0], res[r])
            err = (res[r] - want).norm() / want.norm()
            assert err < 1e-5, err
        assert want.abs().max() > 0
    finally:
        w.close()


def test_exchange_error_behaviour(native_lib):
    import splat_b200_native as native
    h = ctypes.c_void_p()
    assert native_lib.splat_xchg_create(100, 17, 0, ctypes.byref(h)) != 0      # world > 16
    assert b"world" in native_lib.splat_last_error()
    assert native_lib.splat_xchg_create(100, 2, 2, ctypes.byref(h)) != 0       # rank >= world
    native.check(native_lib.splat_xchg_create(100, 2, 0, ctypes.byref(h)))
    try:
        st = torch.cuda.current_stream().cuda_stream
        assert native_lib.splat_xchg_push(h, None, st) != 0                    # not connected
        assert b"not connected" in native_lib.splat_last_error()
    finally:
        native_lib.splat_xchg_destroy(h)
    native.check(native_lib.splat_xchg_create(100, 1, 0, ctypes.byref(h)))
    try:
        assert native_lib.splat_xchg_reduce(h, st) != 0                        # nothing pushed
        native.check(native_lib.splat_xchg_push(h, None, st))
        assert native_lib.splat_xchg_push(h, None, st) != 0                    # push twice
        native.check(native_lib.splat_xchg_allreduce(h, st))
        torch.cuda.synchronize()
    finally:
        native_lib.splat_xchg_destroy(h)
