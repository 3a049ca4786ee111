"""TEST INFRASTRUCTURE ONLY - torch restatement of the map post-processing that the
reference's `render()` applies to the rasterizer's 7xHxW output
(gaussian_renderer/__init__.py:51-92, utils/graphic_utils.py:26-88).  Only tests/ may import
this; the product is the fused CUDA pair in splat-loam_b200/csrc/render_post.cu.

Pinned against outputs of the UNMODIFIED reference `render()` (tests/golden/post_*.npz, made
by tests/golden/make_golden_post.py, which imports the reference in this container).

Written functionally (no in-place edits of the input) and dtype-generic, so the same code
serves as the fp64 "truth" both the reference's fp32 result and the CUDA result are
measured against, and - through autograd - as the gradient reference.

Channels of allmap (DSSR/cuda_rasterizer/auxiliary.h offsets): 0 expected depth * alpha,
1 alpha, 2-4 view-space normal * alpha, 5 median depth, 6 distortion.
"""
import torch


def pixel_rays(world_view_transform, projection_matrix, H, W):
    """World-frame unit ray of every pixel and the sensor origin (graphic_utils.py:36-63).
    The image is an equirectangular range image: K maps (azimuth, elevation) to pixels."""
    dt, dev = world_view_transform.dtype, world_view_transform.device
    sensor_to_world = torch.linalg.inv(world_view_transform.T)          # :36-38
    K_inv = torch.linalg.inv(projection_matrix[:3, :3].T)               # :39-40
    xs = torch.arange(W, dtype=dt, device=dev) - 0.5                    # :41-49
    ys = torch.arange(H, dtype=dt, device=dev) - 0.5
    pix = torch.stack([xs[None, :].expand(H, W), ys[:, None].expand(H, W),
                       torch.ones((H, W), dtype=dt, device=dev)], dim=-1)
    ang = pix @ K_inv.T                                                 # :50
    az, el = ang[..., 0], ang[..., 1]
    local = torch.stack([torch.cos(az) * torch.cos(el), torch.sin(az) * torch.cos(el),
                         torch.sin(el)], dim=-1)                        # :51-59
    return local @ sensor_to_world[:3, :3].T, sensor_to_world[:3, 3]    # :61-62


def normals_from_depth(depth, rays, origin):
    """[3,H,W] normal of the surface a [H,W] depth map describes: normalised cross product of
    the vertical and horizontal central differences of the back-projected points; zero on the
    one-pixel border (graphic_utils.py:69-88)."""
    pts = depth[..., None] * rays + origin                              # :63
    H, W = depth.shape
    out = torch.zeros((H, W, 3), dtype=depth.dtype, device=depth.device)
    if H > 2 and W > 2:
        dv = pts[2:, 1:-1] - pts[:-2, 1:-1]                             # :80-81
        dh = pts[1:-1, 2:] - pts[1:-1, :-2]                             # :82-83
        out[1:-1, 1:-1] = torch.nn.functional.normalize(torch.linalg.cross(dv, dh), dim=-1)
    return out.permute(2, 0, 1)


def post_forward(allmap, world_view_transform, projection_matrix, depth_ratio=0.0):
    """The five maps `render()` returns besides the rasterizer's own outputs."""
    H, W = allmap.shape[1:]
    alpha = allmap[1]
    hit = alpha > 0                                                     # __init__.py:52-53
    safe = torch.where(hit, alpha, torch.ones_like(alpha))
    # view -> world rotation of the normal map, then un-premultiply where alpha > 0 (:56-63)
    n_world = torch.einsum("ij,jhw->ihw", world_view_transform[:3, :3], allmap[2:5])
    n_world = torch.where(hit, n_world / safe, n_world)
    d_expected = torch.where(hit, allmap[0] / safe, allmap[0])          # :70-72
    surf_depth = d_expected * (1 - depth_ratio) + allmap[5] * depth_ratio   # :79-80
    rays, origin = pixel_rays(world_view_transform, projection_matrix, H, W)
    surf_normal = normals_from_depth(surf_depth, rays, origin) * alpha  # :82-83
    return {"rend_alpha": allmap[1:2], "rend_normal": n_world, "rend_dist": allmap[6:7],
            "surf_depth": surf_depth[None], "surf_normal": surf_normal}


def mapper_image_losses(pkg, gt_depth, gt_alpha, lambda_normal, lambda_alpha):
    """The per-keyframe image losses of the mapping loop, as written in slam/mapper.py:158-195
    (inline code there; `gt_alpha` is camera.image_valid, uint8 [1,H,W]; `gt_depth`
    camera.image_depth [1,H,W]).  Returns (geom_l1, normal_loss, alpha_loss, total)."""
    est_alpha, est_depth = pkg["rend_alpha"], pkg["surf_depth"]
    est_normal, surf_normal = pkg["rend_normal"], pkg["surf_normal"]
    valid = gt_alpha[0] == 1.0                                               # :166
    geom_l1 = torch.abs(valid * (est_depth - gt_depth)).mean()               # :174-175
    cos = (est_normal[..., valid] * surf_normal[..., valid]).sum(dim=0)      # :177-179
    normal_loss = (1 - cos).mean() * lambda_normal                           # :177-180
    alpha_loss = torch.nn.functional.binary_cross_entropy(                   # :182-186
        est_alpha[..., valid], gt_alpha[..., valid].to(est_alpha.dtype), reduction="mean")
    alpha_loss = alpha_loss * lambda_alpha                                   # :187
    return geom_l1, normal_loss, alpha_loss, geom_l1 + alpha_loss + normal_loss   # :193-195


OUTPUTS = ("rend_alpha", "rend_normal", "rend_dist", "surf_depth", "surf_normal")


def post_backward(allmap, world_view_transform, projection_matrix, depth_ratio, grads):
    """dL/dallmap for upstream gradients `grads` (dict over OUTPUTS), by autograd."""
    a = allmap.detach().clone().requires_grad_(True)
    out = post_forward(a, world_view_transform, projection_matrix, depth_ratio)
    total = sum((out[k] * grads[k]).sum() for k in OUTPUTS)
    (g,) = torch.autograd.grad(total, a)
    return g
