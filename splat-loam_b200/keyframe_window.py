"""Keyframe-sharded mapping window (BASELINE.json north_star).

The reference optimises ONE randomly sampled keyframe per iteration on one GPU
(slam/mapper.py:150-156).  Here a window of K keyframes is partitioned round-robin over the
ranks of a torch.distributed group (one process per GPU); every rank renders and
back-propagates its keyframes against a full replica of the surfel parameters, accumulating
into ONE contiguous [P,10] fp32 buffer (dL_dmeans3D 3 | dL_dopacity 1 | dL_dscales 2 |
dL_drotations 4), and the buffers are summed over the ranks:
  * on GPUs of one NVLink domain by `PeerExchange` - the per-surfel VJP kernel of a rank's last
    keyframe pushes every row to the rank that owns it while it computes (reduce-scatter half),
    the owners sum and store the result into every rank's memory (all-gather half);
  * otherwise by one all-reduce of the process group (NCCL; gloo in the CPU tests).
dL_dmeans2D is identically zero and dL_dtransMat is not a leaf gradient, so neither is
exchanged.  The result equals the sum over the window of the reference's single-keyframe
gradients.
"""
import ctypes
import os

import torch
import torch.distributed as dist

GRAD_COLS = 10  # 3 + 1 + 2 + 4


class _DeviceArray:
    """Minimal __cuda_array_interface__ holder so torch can view library-owned memory."""

    def __init__(self, ptr, shape, owner):
        self.owner = owner
        self.__cuda_array_interface__ = {"shape": shape, "typestr": "<f4", "data": (ptr, False),
                                         "version": 2, "strides": None}


class PeerExchange:
    """Sum of the ranks' [P,10] buffers over NVLink peer memory (C-ABI splat_xchg_*,
    csrc/peer_exchange.cu).  One instance per (P, process group); all ranks must make the same
    calls.  `result` is a [P,10] tensor view of library-owned device memory that holds the sum
    after `allreduce()` (ordered on the current stream)."""

    def __init__(self, P, device, group=None):
        import splat_b200_native as native
        self._n, self._lib = native, native.load()
        self.group, self.P, self.device = group, int(P), torch.device(device)
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        self.handle = ctypes.c_void_p()
        self.result = None
        with torch.cuda.device(self.device):
            # every step is agreed on by all ranks before the next collective call: a rank whose
            # device cannot export or map peer memory must not leave the others waiting
            err, mine = "", ctypes.create_string_buffer(native.XCHG_HANDLE_BYTES)
            try:
                native.check(self._lib.splat_xchg_create(self.P, world, rank,
                                                         ctypes.byref(self.handle)))
                native.check(self._lib.splat_xchg_export(self.handle, mine))
            except RuntimeError as e:
                err = str(e)
            handles = [None] * world
            dist.all_gather_object(handles, (err, mine.raw), group=group)
            errs = [h[0] for h in handles if h[0]]
            if not errs:
                try:
                    native.check(self._lib.splat_xchg_connect(
                        self.handle, b"".join(h[1] for h in handles)))
                except RuntimeError as e:
                    err = str(e)
                flags = [None] * world
                dist.all_gather_object(flags, err, group=group)
                errs = [f for f in flags if f]
            if errs:
                self.close()
                raise RuntimeError("peer-memory exchange unavailable: " + errs[0])
            ptr = self._lib.splat_xchg_result(self.handle)
            self.result = torch.as_tensor(_DeviceArray(ptr, (self.P, GRAD_COLS), self),
                                          device=self.device)
        dist.barrier(group=group)        # every rank has mapped every arena

    def push(self, packed=None):
        """Unfused push of a finished [P,10] buffer (None: this rank contributes zeros)."""
        with torch.cuda.device(self.device):
            self._n.check(self._lib.splat_xchg_push(
                self.handle, packed.data_ptr() if packed is not None else None,
                torch.cuda.current_stream(self.device).cuda_stream))

    def allreduce(self):
        with torch.cuda.device(self.device):
            self._n.check(self._lib.splat_xchg_allreduce(
                self.handle, torch.cuda.current_stream(self.device).cuda_stream))
        return self.result

    def check(self):
        """Synchronise the current stream and raise if this or ANY peer rank gave up waiting
        during an exchange: a rank that times out marks every rank's arena before it goes on, so
        all ranks fail here together and none applies the stale sums."""
        torch.cuda.current_stream(self.device).synchronize()
        code = int(self._lib.splat_xchg_fault(self.handle))
        if code:
            raise RuntimeError(
                f"peer-memory exchange: a wait for rank {(code - 1) // 16} gave up in step "
                f"{(code - 1) % 16}; the summed gradients of this window are invalid on every rank")

    def close(self):
        if self.handle:
            self.result = None
            self._lib.splat_xchg_destroy(self.handle)
            self.handle = ctypes.c_void_p()

    def __del__(self):      # the arena (device memory + IPC mappings) must not outlive us
        try:
            self.close()
        except Exception:
            pass


def shard_keyframes(num_keyframes, rank, world_size):
    """Keyframes owned by `rank`: k with k mod world_size == rank."""
    return list(range(rank, num_keyframes, world_size))


def cuda_frame_grads(params, camera, dL_dout_fn, out=None, xchg=None, accumulate=None):
    """One keyframe through the CUDA rasterizer (drop-in `_C`): returns [P,10] gradients and
    the rendered map.  `camera` = (viewmatrix, projmatrix, H, W); `dL_dout_fn(allmap)` gives the
    upstream gradient of the 7xHxW map.  With `out` ([P,10]) the per-surfel VJP kernel adds
    its result straight into that buffer (no concatenation or add kernels); with `xchg` (a
    PeerExchange) it also pushes the rows to their owner ranks."""
    from diff_surfel_spherical_rasterization import _C
    means3D, opacities, scales, rotations = params
    view, proj, H, W = camera
    empty = torch.empty(0, device=means3D.device)
    R, allmap, radii, geom, binning, img = _C.rasterize_gaussians(
        means3D, opacities, scales, rotations, 1.0, empty, view, proj, H, W, False, False)
    dL = dL_dout_fn(allmap)
    packed = out if out is not None else torch.empty((means3D.shape[0], GRAD_COLS),
                                                     dtype=torch.float32, device=means3D.device)
    _C.rasterize_gaussians_backward(means3D, radii, scales, rotations, 1.0, empty, view, proj, dL,
                                    geom, R, binning, img, False, packed=packed,
                                    packed_accumulate=(out is not None) if accumulate is None
                                    else accumulate,
                                    xchg=xchg.handle.value if xchg is not None else None)
    return packed, allmap


class KeyframeWindow:
    """Gradient of a multi-keyframe window, sharded by keyframe across the process group."""

    def __init__(self, frame_fn=cuda_frame_grads, group=None, exchange="auto"):
        """exchange: "auto" (peer memory when the frames run on CUDA with more than one rank
        and SPLAT_B200_XCHG != "nccl"), "peer", or "collective" (process-group all-reduce)."""
        self.frame_fn = frame_fn
        self.group = group
        self.exchange = exchange
        self._buf = None
        self._xchg = None

    def _peer_exchange(self, P, device):
        _, world = self._world()
        want = self.exchange
        if want == "auto":
            want = "peer" if (self.frame_fn is cuda_frame_grads and world > 1 and
                              os.environ.get("SPLAT_B200_XCHG", "peer") != "nccl") else "collective"
        if want != "peer":
            return None
        if self._xchg is None or self._xchg.P != P or self._xchg.device != device:
            if self._xchg is not None:
                self._xchg.close()
                self._xchg = None
            try:
                self._xchg = PeerExchange(P, device, self.group)
            except RuntimeError as e:
                if self.exchange != "auto":
                    raise
                import sys
                print(f"[keyframe_window] {e}; using the process group's all-reduce", file=sys.stderr)
                self.exchange = "collective"       # agreed by all ranks (see PeerExchange)
                return None
        return self._xchg

    def _world(self):
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(self.group), dist.get_world_size(self.group)
        return 0, 1

    def accumulate(self, params, cameras, dL_dout_fn, local_only=False):
        """Local part: sum of the [P,10] gradients of this rank's keyframes (no communication).
        `local_only`: of ALL keyframes, i.e. the whole window on this GPU alone."""
        rank, world = (0, 1) if local_only else self._world()
        P = params[0].shape[0]
        if self._buf is None or self._buf.shape[0] != P or self._buf.device != params[0].device:
            self._buf = torch.empty((P, GRAD_COLS), dtype=torch.float32, device=params[0].device)
        buf = self._buf.zero_()
        mine = shard_keyframes(len(cameras), rank, world)
        for k in mine:
            if self.frame_fn is cuda_frame_grads:   # accumulates in place
                self.frame_fn(params, cameras[k], lambda m, k=k: dL_dout_fn(k, m), out=buf)
            else:
                g, _ = self.frame_fn(params, cameras[k], lambda m, k=k: dL_dout_fn(k, m))
                buf += g
        return buf, mine

    def gradients(self, params, cameras, dL_dout_fn, check=True):
        """Summed [P,10] gradient of the whole window, identical on every rank.  With `check`
        (peer-memory path) the stream is synchronised and an exchange in which any rank gave
        up waiting raises on every rank, before anybody applies the result; pass False to keep
        the step asynchronous and call `exchange_check()` before the optimizer step instead."""
        rank, world = self._world()
        xchg = self._peer_exchange(params[0].shape[0], params[0].device) if world > 1 else None
        if xchg is None:
            buf, _ = self.accumulate(params, cameras, dL_dout_fn)
            if world > 1:
                dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=self.group)
            return buf
        # peer-memory path: local keyframes accumulate in `buf`; the VJP kernel of the LAST one
        # adds what `buf` holds and pushes the rows to their owners
        P = params[0].shape[0]
        if self._buf is None or self._buf.shape[0] != P or self._buf.device != params[0].device:
            self._buf = torch.empty((P, GRAD_COLS), dtype=torch.float32, device=params[0].device)
        buf = self._buf
        mine = shard_keyframes(len(cameras), rank, world)
        for j, k in enumerate(mine):
            self.frame_fn(params, cameras[k], lambda m, k=k: dL_dout_fn(k, m), out=buf,
                          accumulate=j > 0, xchg=xchg if j == len(mine) - 1 else None)
        if not mine:
            xchg.push(None)
        res = xchg.allreduce()
        if check:
            xchg.check()
        return res

    def exchange_check(self):
        if self._xchg is not None:
            self._xchg.check()

    @staticmethod
    def split(buf):
        """[P,10] -> (dL_dmeans3D [P,3], dL_dopacity [P,1], dL_dscales [P,2], dL_drotations [P,4])."""
        return buf[:, 0:3], buf[:, 3:4], buf[:, 4:6], buf[:, 6:10]
