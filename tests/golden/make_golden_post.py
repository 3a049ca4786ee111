"""Generate tests/golden/post_*.npz: inputs and outputs of the UNMODIFIED reference `render()`
post-processing (gaussian_renderer/__init__.py:11-93 with utils/graphic_utils.py), imported
from /root/reference in this container and run on the CPU in fp32:

    python tests/golden/make_golden_post.py            # writes into tests/golden/

Only the rasterizer call inside `render()` is replaced - by a stub that returns a prepared
7xHxW map - and the two class imports `render()` uses as type annotations are stubbed so
that the reference's heavier dependencies (open3d, plyfile, simple_knn) are not needed.
Everything after the rasterizer call, including autograd through it, is the reference's.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import scenes  # noqa: E402
from oracle import oracle  # noqa: E402


def import_reference_render(state):
    def stub(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    class Settings:
        def __init__(self, **kw):
            self.__dict__.update(kw)

    class Rasterizer:
        def __init__(self, raster_settings):
            self.raster_settings = raster_settings

        def __call__(self, means3D, means2D, opacities, scales, rotations, cov3D_precomp=None):
            # a non-leaf copy: render() edits views of it in place
            return state["radii"], state["allmap"] * 1.0

    stub("diff_surfel_spherical_rasterization", GaussianRasterizer=Rasterizer,
         GaussianRasterizationSettings=Settings)
    stub("scene")
    stub("scene.cameras", Camera=object)
    stub("scene.gaussian_model", GaussianModel=object)
    sys.path.insert(0, REF)
    from gaussian_renderer import render   # the reference's
    return render


class Cam:
    def __init__(self, view, proj, H, W):
        self.world_view_transform = torch.from_numpy(view)
        self.projection_matrix = torch.from_numpy(proj)
        self.image_height, self.image_width = H, W


class Model:
    def __init__(self, P):
        self.get_xyz = torch.zeros((P, 3))
        self.get_opacity = torch.zeros((P, 1))
        self.get_scaling = torch.zeros((P, 2))
        self.get_rotation = torch.zeros((P, 4))


OUTPUTS = ("rend_alpha", "rend_normal", "rend_dist", "surf_depth", "surf_normal")


def case(render, state, allmap, view, proj, depth_ratio, seed):
    H, W = allmap.shape[1:]
    state["allmap"] = torch.from_numpy(allmap.copy()).requires_grad_(True)
    state["radii"] = torch.ones(4, dtype=torch.int32)
    pkg = render(Cam(view, proj, H, W), Model(4), depth_ratio)
    rng = np.random.default_rng(seed)
    grads = {k: rng.standard_normal(tuple(pkg[k].shape)).astype(np.float32) for k in OUTPUTS}
    total = sum((pkg[k] * torch.from_numpy(grads[k])).sum() for k in OUTPUTS)
    total.backward()
    g = {"allmap": allmap, "viewmatrix": view, "projmatrix": proj,
         "depth_ratio": np.float32(depth_ratio), "dL_dallmap": state["allmap"].grad.numpy()}
    for k in OUTPUTS:
        g[k] = pkg[k].detach().numpy()
        g["grad_" + k] = grads[k]
    return g


def loss_case(render, state, allmap, view, proj, depth_ratio, seed, lam_n, lam_a):
    """render() is the reference's; the 12 loss lines of slam/mapper.py:158-195 have no callable
    boundary there (inline in Mapper.optimize), so they are evaluated through the restatement
    oracle/render_post.mapper_image_losses (the torch primitives are the reference's)."""
    from oracle import render_post as rp
    H, W = allmap.shape[1:]
    state["allmap"] = torch.from_numpy(allmap.copy()).requires_grad_(True)
    state["radii"] = torch.ones(4, dtype=torch.int32)
    pkg = render(Cam(view, proj, H, W), Model(4), depth_ratio)
    rng = np.random.default_rng(seed)
    gt_depth = (pkg["surf_depth"].detach().numpy() + 0.2 * rng.standard_normal((1, H, W))).astype(np.float32)
    gt_depth[0, rng.uniform(size=(H, W)) < 0.02] = pkg["surf_depth"].detach().numpy()[0].mean()
    gt_valid = (rng.uniform(size=(1, H, W)) < 0.8).astype(np.uint8)
    gt_valid[0][allmap[1] <= 0] = 0        # BCE of alpha == 0 is the -100 clamp: kept out of the
    gt_valid[0, :2, :5] = 1                # mean except for this deliberate corner
    losses = rp.mapper_image_losses(pkg, torch.from_numpy(gt_depth), torch.from_numpy(gt_valid),
                                    lam_n, lam_a)
    losses[3].backward()
    return {"allmap": allmap, "viewmatrix": view, "projmatrix": proj,
            "depth_ratio": np.float32(depth_ratio), "gt_depth": gt_depth, "gt_valid": gt_valid,
            "lambda_normal": np.float32(lam_n), "lambda_alpha": np.float32(lam_a),
            "losses": np.array([float(x.detach()) for x in losses], dtype=np.float64),
            "dL_dallmap": state["allmap"].grad.numpy()}


def rasterized_map(cfg, keyframe):
    s = scenes.make_scene(cfg, keyframe=keyframe)
    oracle.build()
    f = oracle.rasterize_forward(s["means3D"], s["scales"], s["rotations"], s["opacities"],
                                 s["viewmatrix"], s["projmatrix"], s["H"], s["W"])
    return np.ascontiguousarray(f["out"], dtype=np.float32), s["viewmatrix"], s["projmatrix"]


def synthetic_map(H, W, seed):
    """Smooth depth with holes (alpha == 0), random normals: exercises every branch."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    depth = 8.0 + 3.0 * np.sin(xx / 17.0) + 2.0 * np.cos(yy / 5.0) + 0.05 * rng.standard_normal((H, W))
    alpha = np.clip(rng.uniform(-0.3, 1.2, (H, W)), 0.0, 0.999).astype(np.float32)
    n = rng.standard_normal((3, H, W)).astype(np.float32)
    n /= np.linalg.norm(n, axis=0, keepdims=True)
    m = np.zeros((7, H, W), dtype=np.float32)
    m[0] = depth * alpha
    m[1] = alpha
    m[2:5] = n * alpha
    m[5] = depth + 0.1 * rng.standard_normal((H, W))
    m[6] = rng.uniform(0, 0.01, (H, W))
    return m.astype(np.float32)


def main(outdir):
    state = {}
    render = import_reference_render(state)
    m, view, proj = rasterized_map("C0", 1)
    np.savez_compressed(os.path.join(outdir, "post_c0.npz"), **case(render, state, m, view, proj, 0.0, 1))
    np.savez_compressed(os.path.join(outdir, "post_c0_ratio.npz"),
                        **case(render, state, m, view, proj, 0.3, 2))
    np.savez_compressed(os.path.join(outdir, "post_loss_c0.npz"),
                        **loss_case(render, state, m, view, proj, 0.0, 11, 0.1, 0.1))
    H, W = 24, 80
    proj = scenes.intrinsics(H, W, 40.0)
    np.savez_compressed(os.path.join(outdir, "post_loss_synth.npz"),
                        **loss_case(render, state, synthetic_map(H, W, 8), scenes.viewmatrix(3),
                                    proj, 0.4, 12, 0.05, 0.5))
    view = scenes.viewmatrix(3)
    np.savez_compressed(os.path.join(outdir, "post_synth.npz"),
                        **case(render, state, synthetic_map(H, W, 5), view, proj, 1.0, 3))
    np.savez_compressed(os.path.join(outdir, "post_synth_mix.npz"),
                        **case(render, state, synthetic_map(H, W, 6), view, proj, 0.5, 4))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else HERE)
