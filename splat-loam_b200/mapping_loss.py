"""The mapper's per-keyframe image losses, fused with render()'s map post-processing.

The reference evaluates them inline in `Mapper.optimize` (slam/mapper.py:158-195) with ~15
PyTorch kernels on top of the ~25 of render()'s post-processing, and twice that through
autograd.  Here the rasterizer's [7,H,W] output goes through ONE kernel that produces the
loss values (C-ABI dssr_post_loss_forward) and ONE that produces dL/dallmap
(dssr_post_loss_backward); the maps themselves are never written unless asked for.

    losses = mapping_image_losses(allmap, camera, depth_ratio, lambda_normal, lambda_alpha)
    total = losses["total"] + reg_scales            # mapper.py:186-195; reg_scales stays torch
    total.backward()

replaces mapper.py:154-190 (the `render()` call included; use `render_and_losses` to keep
one call).  `camera` needs world_view_transform, projection_matrix, image_depth [1,H,W]
float32 and image_valid [1,H,W] uint8 (scene/cameras.py).  No CPU path.
"""
import ctypes

import torch

import splat_b200_native as _n

_lib = _n.load()
_workspaces = {}


def _workspace(device):
    ws = _workspaces.get(device)
    if ws is None:
        ws = torch.zeros(_n.POST_LOSS_WORKSPACE_BYTES, dtype=torch.uint8, device=device)
        _workspaces[device] = ws
    return ws


def _f32c(t):
    t = t.contiguous()
    return t if t.dtype == torch.float32 else t.float()


class _ImageLosses(torch.autograd.Function):
    """allmap [7,H,W] -> losses [5] = (geom_l1, normal, alpha, total, n_valid)."""

    @staticmethod
    def forward(ctx, allmap, viewmatrix, projmatrix, gt_depth, gt_valid, depth_ratio,
                lambda_normal, lambda_alpha, maps):
        if not allmap.is_cuda:
            raise RuntimeError("allmap must be a CUDA tensor")
        allmap, viewmatrix, projmatrix = _f32c(allmap), _f32c(viewmatrix), _f32c(projmatrix)
        gt_depth = _f32c(gt_depth)
        gt_valid = gt_valid.contiguous()
        if gt_valid.dtype != torch.uint8:
            gt_valid = gt_valid.to(torch.uint8)
        _, H, W = allmap.shape
        if gt_depth.numel() != H * W or gt_valid.numel() != H * W:
            raise RuntimeError("gt_depth / gt_valid must have H*W elements")
        dev = allmap.device
        losses = torch.empty(5, dtype=torch.float32, device=dev)
        ptrs = [None, None, None]
        if maps is not None:       # caller-provided [7,H,W] buffer: 3 | 1 | 3 channels
            ptrs = [maps[0:3].data_ptr(), maps[3:4].data_ptr(), maps[4:7].data_ptr()]
        with torch.cuda.device(dev):
            a = _n.DssrPostLossForwardArgs(
                W, H, float(depth_ratio), allmap.data_ptr(), viewmatrix.data_ptr(),
                projmatrix.data_ptr(), gt_depth.data_ptr(), gt_valid.data_ptr(),
                float(lambda_normal), float(lambda_alpha), *ptrs, losses.data_ptr(),
                _workspace(dev).data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
            _n.check(_lib.dssr_post_loss_forward(ctypes.byref(a)))
        ctx.save_for_backward(allmap, viewmatrix, projmatrix, gt_depth, gt_valid, losses)
        ctx.scalars = (float(depth_ratio), float(lambda_normal), float(lambda_alpha))
        return losses

    @staticmethod
    def backward(ctx, g_losses):
        allmap, viewmatrix, projmatrix, gt_depth, gt_valid, losses = ctx.saved_tensors
        depth_ratio, lambda_normal, lambda_alpha = ctx.scalars
        _, H, W = allmap.shape
        dev = allmap.device
        g = _f32c(g_losses)
        dallmap = torch.empty_like(allmap)
        with torch.cuda.device(dev):
            a = _n.DssrPostLossBackwardArgs(
                W, H, depth_ratio, allmap.data_ptr(), viewmatrix.data_ptr(), projmatrix.data_ptr(),
                gt_depth.data_ptr(), gt_valid.data_ptr(), lambda_normal, lambda_alpha,
                losses.data_ptr(), g.data_ptr(), dallmap.data_ptr(),
                torch.cuda.current_stream(dev).cuda_stream)
            _n.check(_lib.dssr_post_loss_backward(ctypes.byref(a)))
        return dallmap, None, None, None, None, None, None, None, None


def mapping_image_losses(allmap, camera, depth_ratio=0.0, lambda_normal=0.1, lambda_alpha=0.1,
                         maps=None):
    """geom_l1 + normal + alpha losses of one keyframe (slam/mapper.py:166-190) from the
    rasterizer output.  `maps`: optional preallocated float32 [7,H,W] CUDA tensor that receives
    rend_normal | surf_depth | surf_normal (not differentiable; for logging / densification)."""
    losses = _ImageLosses.apply(allmap, camera.world_view_transform, camera.projection_matrix,
                                camera.image_depth, camera.image_valid, depth_ratio,
                                lambda_normal, lambda_alpha, maps)
    return {"geom_l1": losses[0], "normal_loss": losses[1], "alpha_loss": losses[2],
            "total": losses[3], "n_valid": losses[4].detach()}


def render_and_losses(camera, model, depth_ratio=0.0, lambda_normal=0.1, lambda_alpha=0.1):
    """Rasterize `model` into `camera` and evaluate the image losses: the body of one mapping
    iteration up to `loss_total` (mapper.py:154-190) in 10 + 1 kernels forward, 1 + 2 backward.
    Returns (losses dict, radii)."""
    from diff_surfel_spherical_rasterization import (GaussianRasterizationSettings,
                                                     GaussianRasterizer)
    settings = GaussianRasterizationSettings(
        image_height=int(camera.image_height), image_width=int(camera.image_width),
        scale_modifier=1.0, viewmatrix=camera.world_view_transform,
        projmatrix=camera.projection_matrix, prefiltered=False, debug=False)
    means3D = model.get_xyz
    radii, allmap = GaussianRasterizer(raster_settings=settings)(
        means3D=means3D, means2D=torch.zeros_like(means3D), opacities=model.get_opacity,
        scales=model.get_scaling, rotations=model.get_rotation, cov3D_precomp=None)
    return mapping_image_losses(allmap, camera, depth_ratio, lambda_normal, lambda_alpha), radii
