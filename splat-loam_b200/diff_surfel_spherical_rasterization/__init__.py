"""Drop-in replacement for the reference's `diff_surfel_spherical_rasterization` package.

Public surface (names, argument order, return values and error behaviour) follows
submodules/diff-surfel-spherical-rasterization/diff_surfel_spherical_rasterization/
__init__.py:18-258 so that gaussian_renderer/__init__.py, slam/mapper.py and
slam/tracker.py run unchanged:

    GaussianRasterizationSettings   NamedTuple (:18-25)
    GaussianRasterizer              nn.Module with forward / markVisible (:203-258)
    rasterize_gaussians             functional form (:36-53)
    _C                              rasterize_gaussians, rasterize_gaussians_backward,
                                    mark_visible

The native side is hand-written sm_100a CUDA behind a C-ABI (include/splat_b200.h);
there is no CPU or PyTorch fallback.
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _C

__all__ = [
    "GaussianRasterizationSettings",
    "GaussianRasterizer",
    "rasterize_gaussians",
]


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    prefiltered: bool
    debug: bool


def cpu_deep_copy_tuple(input_tuple):
    """CPU clones of every tensor in a tuple (used for the debug snapshots)."""
    return tuple(x.cpu().clone() if isinstance(x, torch.Tensor) else x for x in input_tuple)


def _call_with_snapshot(fn, args, debug, dump_name, what):
    """Reference debug behaviour (:86-104, :161-179): with debug=True keep CPU copies of
    the arguments and dump them if the native call throws."""
    if not debug:
        return fn(*args)
    saved = cpu_deep_copy_tuple(args)
    try:
        return fn(*args)
    except Exception:
        torch.save(saved, dump_name)
        print(f"\nAn error occured in {what}. Please forward {dump_name} for debugging.")
        raise


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        rs = raster_settings
        args = (means3D, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.image_height, rs.image_width, rs.prefiltered,
                rs.debug)
        num_rendered, depth, radii, geomBuffer, binningBuffer, imgBuffer = _call_with_snapshot(
            _C.rasterize_gaussians, args, rs.debug, "snapshot_fw.dump", "forward")
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        # `depth` (the 7xHxW map) is NOT saved: render() post-processes it in place
        ctx.save_for_backward(means3D, scales, rotations, cov3Ds_precomp, radii, geomBuffer,
                              binningBuffer, imgBuffer)
        return radii, depth

    @staticmethod
    def backward(ctx, grad_radii, grad_depth):
        rs = ctx.raster_settings
        (means3D, scales, rotations, cov3Ds_precomp, radii, geomBuffer, binningBuffer,
         imgBuffer) = ctx.saved_tensors
        args = (means3D, radii, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, grad_depth, geomBuffer, ctx.num_rendered,
                binningBuffer, imgBuffer, rs.debug)
        (grad_means2D, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_scales,
         grad_rotations) = _call_with_snapshot(
            _C.rasterize_gaussians_backward, args, rs.debug, "snapshot_bw.dump", "backward")
        # gradients in the order of forward()'s inputs; settings get None
        return (grad_means3D, grad_means2D, grad_opacities, grad_scales, grad_rotations,
                grad_cov3Ds_precomp, None)


def rasterize_gaussians(means3D, means2D, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """bool[P]: 0.1 < |V p| <= 120 (checkFrustum)."""
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, scales=None, rotations=None,
                cov3D_precomp=None):
        have_sr = scales is not None or rotations is not None
        missing_sr = scales is None or rotations is None
        if (missing_sr and cov3D_precomp is None) or (have_sr and cov3D_precomp is not None):
            raise Exception("Please provide exactly one of either scale/rotation "
                            "pair or precomputed 3D covariance!")
        dev = means3D.device
        empty = torch.empty(0, dtype=torch.float32, device=dev if dev.type == "cuda" else "cuda")
        if scales is None:
            scales = empty
        if rotations is None:
            rotations = empty
        if cov3D_precomp is None:
            cov3D_precomp = empty
        return rasterize_gaussians(means3D, means2D, opacities, scales, rotations,
                                   cov3D_precomp, self.raster_settings)
