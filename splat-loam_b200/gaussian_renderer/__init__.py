"""Drop-in for the reference's `gaussian_renderer` package (gaussian_renderer/__init__.py):
`render(camera, model, depth_ratio)` with the same arguments and the same result dictionary,
built from the CUDA rasterizer of this repository and ONE fused post-processing kernel per
direction (csrc/render_post.cu, C-ABI dssr_post_forward / dssr_post_backward) instead of the
~25 forward / ~50 backward PyTorch kernels the reference spends on the 7xHxW map.

`camera` needs image_height, image_width, world_view_transform, projection_matrix
(scene/cameras.py); `model` needs get_xyz, get_opacity, get_scaling, get_rotation
(scene/gaussian_model.py).  There is no CPU path.
"""
import ctypes

import torch

import splat_b200_native as _n
from diff_surfel_spherical_rasterization import (GaussianRasterizationSettings,
                                                 GaussianRasterizer)

_lib = _n.load()


def _f32c(t):
    t = t.contiguous()
    return t if t.dtype == torch.float32 else t.float()


class _PostProcess(torch.autograd.Function):
    """allmap [7,H,W] -> (rend_normal [3,H,W], surf_depth [1,H,W], surf_normal [3,H,W])."""

    @staticmethod
    def forward(ctx, allmap, viewmatrix, projmatrix, depth_ratio):
        if not allmap.is_cuda:
            raise RuntimeError("allmap must be a CUDA tensor")
        allmap, viewmatrix, projmatrix = _f32c(allmap), _f32c(viewmatrix), _f32c(projmatrix)
        _, H, W = allmap.shape
        dev = allmap.device
        out = torch.empty((7, H, W), dtype=torch.float32, device=dev)   # 3 | 1 | 3 channels
        with torch.cuda.device(dev):
            a = _n.DssrPostForwardArgs(
                W, H, float(depth_ratio), allmap.data_ptr(), viewmatrix.data_ptr(),
                projmatrix.data_ptr(), out[0:3].data_ptr(), out[3:4].data_ptr(),
                out[4:7].data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
            _n.check(_lib.dssr_post_forward(ctypes.byref(a)))
        ctx.save_for_backward(allmap, viewmatrix, projmatrix)
        ctx.depth_ratio = float(depth_ratio)
        return out[0:3], out[3:4], out[4:7]

    @staticmethod
    def backward(ctx, g_rend_normal, g_surf_depth, g_surf_normal):
        allmap, viewmatrix, projmatrix = ctx.saved_tensors
        _, H, W = allmap.shape
        dev = allmap.device
        grads = [None if g is None else _f32c(g) for g in (g_rend_normal, g_surf_depth, g_surf_normal)]
        dallmap = torch.empty_like(allmap)
        with torch.cuda.device(dev):
            a = _n.DssrPostBackwardArgs(
                W, H, ctx.depth_ratio, allmap.data_ptr(), viewmatrix.data_ptr(),
                projmatrix.data_ptr(), *[None if g is None else g.data_ptr() for g in grads],
                dallmap.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
            _n.check(_lib.dssr_post_backward(ctypes.byref(a)))
        return dallmap, None, None, None


def postprocess(allmap, world_view_transform, projection_matrix, depth_ratio=0.0):
    """The maps `render()` derives from the rasterizer output (differentiable w.r.t. allmap)."""
    rend_normal, surf_depth, surf_normal = _PostProcess.apply(
        allmap, world_view_transform, projection_matrix, depth_ratio)
    return {"rend_alpha": allmap[1:2], "rend_normal": rend_normal, "rend_dist": allmap[6:7],
            "surf_depth": surf_depth, "surf_normal": surf_normal}


def render(camera, model, depth_ratio: float = 0.0):
    """Same contract as the reference's render() (gaussian_renderer/__init__.py:11-93)."""
    raster_settings = GaussianRasterizationSettings(
        image_height=int(camera.image_height), image_width=int(camera.image_width),
        scale_modifier=1.0, viewmatrix=camera.world_view_transform,
        projmatrix=camera.projection_matrix, prefiltered=False, debug=False)
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)
    means3D = model.get_xyz
    means2D = torch.zeros_like(means3D, dtype=torch.float32, device=means3D.device)
    radii, allmap = rasterizer(means3D=means3D, means2D=means2D, opacities=model.get_opacity,
                               scales=model.get_scaling, rotations=model.get_rotation,
                               cov3D_precomp=None)
    rets = {"viewspace_points": means2D, "visibility_filter": radii > 0, "radii": radii}
    rets.update(postprocess(allmap, camera.world_view_transform, camera.projection_matrix,
                            depth_ratio))
    return rets
