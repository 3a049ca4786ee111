"""Keyframe-sharded mapping window (BASELINE.json north_star).

The reference optimises ONE randomly sampled keyframe per iteration on one GPU
(slam/mapper.py:150-156).  Here a window of K keyframes is partitioned round-robin over the
ranks of a torch.distributed group (one process per GPU); every rank renders and
back-propagates its keyframes against a full replica of the surfel parameters, accumulating
into ONE contiguous [P,10] fp32 buffer (dL_dmeans3D 3 | dL_dopacity 1 | dL_dscales 2 |
dL_drotations 4), and a single all-reduce (NCCL over NVLink on the GPU box, gloo in the CPU
tests) sums the buffers.  dL_dmeans2D is identically zero and dL_dtransMat is not a leaf
gradient, so neither is exchanged.  The result equals the sum over the window of the
reference's single-keyframe gradients.
"""
import torch
import torch.distributed as dist

GRAD_COLS = 10  # 3 + 1 + 2 + 4


def shard_keyframes(num_keyframes, rank, world_size):
    """Keyframes owned by `rank`: k with k mod world_size == rank."""
    return list(range(rank, num_keyframes, world_size))


def cuda_frame_grads(params, camera, dL_dout_fn, out=None):
    """One keyframe through the CUDA rasterizer (drop-in `_C`): returns [P,10] gradients and
    the rendered map.  `camera` = (viewmatrix, projmatrix, H, W); `dL_dout_fn(allmap)` gives the
    upstream gradient of the 7xHxW map.  With `out` ([P,10]) the per-surfel VJP kernel adds
    its result straight into that buffer (no concatenation or add kernels)."""
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
                                    packed_accumulate=out is not None)
    return packed, allmap


class KeyframeWindow:
    """Gradient of a multi-keyframe window, sharded by keyframe across the process group."""

    def __init__(self, frame_fn=cuda_frame_grads, group=None):
        self.frame_fn = frame_fn
        self.group = group
        self._buf = None

    def _world(self):
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(self.group), dist.get_world_size(self.group)
        return 0, 1

    def accumulate(self, params, cameras, dL_dout_fn):
        """Local part: sum of the [P,10] gradients of this rank's keyframes (no communication)."""
        rank, world = self._world()
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

    def gradients(self, params, cameras, dL_dout_fn):
        """All-reduced [P,10] gradient of the whole window, identical on every rank."""
        buf, _ = self.accumulate(params, cameras, dL_dout_fn)
        _, world = self._world()
        if world > 1:
            dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=self.group)
        return buf

    @staticmethod
    def split(buf):
        """[P,10] -> (dL_dmeans3D [P,3], dL_dopacity [P,1], dL_dscales [P,2], dL_drotations [P,4])."""
        return buf[:, 0:3], buf[:, 3:4], buf[:, 4:6], buf[:, 6:10]
