"""GPU: the fused post-processing + mapper image losses (C-ABI dssr_post_loss_forward /
dssr_post_loss_backward through splat-loam_b200/mapping_loss.py) against goldens made with
the reference's render() and against the fp64 oracle.  Tolerances: loss values 1e-5 relative,
gradients 1e-4 relative L2 with cosine > 0.9999."""
import numpy as np
import pytest
import torch

from conftest import cosine, load_golden, rel_l2
from oracle import render_post as rp

pytestmark = pytest.mark.gpu


class Cam:
    def __init__(self, V, Pm, gt_depth, gt_valid):
        self.world_view_transform = torch.from_numpy(V).cuda()
        self.projection_matrix = torch.from_numpy(Pm).cuda()
        self.image_depth = torch.from_numpy(gt_depth).cuda()
        self.image_valid = torch.from_numpy(gt_valid).cuda()


def product(g, which="total", maps=None):
    import mapping_loss as ml
    a = torch.from_numpy(g["allmap"]).cuda().requires_grad_(True)
    cam = Cam(g["viewmatrix"], g["projmatrix"], g["gt_depth"], g["gt_valid"])
    out = ml.mapping_image_losses(a, cam, float(g["depth_ratio"]), float(g["lambda_normal"]),
                                  float(g["lambda_alpha"]), maps=maps)
    out[which].backward()
    return {k: float(v.detach()) for k, v in out.items()}, a.grad.cpu().numpy()


def oracle64(g, which=3):
    t = lambda k: torch.from_numpy(np.asarray(g[k])).double()  # noqa: E731
    a = t("allmap").requires_grad_(True)
    pkg = rp.post_forward(a, t("viewmatrix"), t("projmatrix"), float(g["depth_ratio"]))
    losses = rp.mapper_image_losses(pkg, t("gt_depth"), torch.from_numpy(g["gt_valid"]),
                                    float(g["lambda_normal"]), float(g["lambda_alpha"]))
    losses[which].backward()
    return [float(x.detach()) for x in losses], a.grad.numpy()


@pytest.mark.parametrize("name", ("post_loss_c0", "post_loss_synth"))
def test_against_golden(native_lib, name):
    g = load_golden(name)
    for _ in range(2):     # the workspace must come back zeroed
        vals, grad = product(g)
        for k, want in zip(("geom_l1", "normal_loss", "alpha_loss", "total"), g["losses"]):
            assert abs(vals[k] - want) <= 1e-5 * max(1.0, abs(want)), k
        assert vals["n_valid"] == float((g["gt_valid"] == 1).sum())
        assert rel_l2(grad, g["dL_dallmap"]) < 1e-4 and cosine(grad, g["dL_dallmap"]) > 0.9999
        for c in range(6):
            if np.abs(g["dL_dallmap"][c]).max() > 0:
                assert rel_l2(grad[c], g["dL_dallmap"][c]) < 1e-4, c
        assert not grad[6].any()


@pytest.mark.parametrize("which,idx", [("geom_l1", 0), ("normal_loss", 1), ("alpha_loss", 2)])
def test_each_component_backpropagates_alone(native_lib, which, idx):
    g = load_golden("post_loss_synth")
    _, grad = product(g, which)
    _, ref = oracle64(g, idx)
    assert rel_l2(grad, ref) < 1e-4


@pytest.mark.parametrize("H,W", [(64, 1024), (3, 3), (17, 50)])
def test_against_fp64_oracle_at_other_sizes(native_lib, H, W):
    import os, sys
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
    from make_golden_post import synthetic_map
    import scenes
    rng = np.random.default_rng(H + W)
    m = synthetic_map(H, W, 3 * H + W)
    g = {"allmap": m, "viewmatrix": scenes.viewmatrix(1), "projmatrix": scenes.intrinsics(max(H, 4), max(W, 4), 45.0),
         "depth_ratio": 0.3, "lambda_normal": 0.1, "lambda_alpha": 0.1,
         "gt_depth": (m[5:6] + 0.3 * rng.standard_normal((1, H, W))).astype(np.float32),
         "gt_valid": ((rng.uniform(size=(1, H, W)) < 0.7) & (m[1:2] > 0)).astype(np.uint8)}
    g["gt_valid"][0, 0, 0] = 1
    g["allmap"][1, 0, 0] = 0.5
    vals, grad = product(g)
    want, ref = oracle64(g)
    for k, w in zip(("geom_l1", "normal_loss", "alpha_loss", "total"), want):
        assert abs(vals[k] - w) <= 1e-5 * max(1.0, abs(w)), k
    assert rel_l2(grad, ref) < 1e-4 and cosine(grad, ref) > 0.9999


def test_maps_on_request_and_no_valid_pixel(native_lib):
    g = dict(load_golden("post_loss_synth"))
    H, W = g["allmap"].shape[1:]
    maps = torch.empty((7, H, W), device="cuda")
    product(g, maps=maps)
    t = lambda k: torch.from_numpy(np.asarray(g[k]))  # noqa: E731
    ref = rp.post_forward(t("allmap"), t("viewmatrix"), t("projmatrix"), float(g["depth_ratio"]))
    assert float((maps[0:3].cpu() - ref["rend_normal"]).abs().max()) < 1e-5
    assert float((maps[3:4].cpu() - ref["surf_depth"]).abs().max()) < 1e-5
    assert float((maps[4:7].cpu() - ref["surf_normal"]).abs().max()) < 2e-5
    g["gt_valid"] = np.zeros_like(g["gt_valid"])
    vals, _ = product(g)
    assert vals["geom_l1"] == 0.0 and np.isnan(vals["normal_loss"]) and np.isnan(vals["alpha_loss"])
    g["gt_valid"] = load_golden("post_loss_synth")["gt_valid"]
    vals2, _ = product(g)      # and the op recovers on the next call
    assert np.isfinite(vals2["total"])


def test_render_and_losses_reaches_the_parameters(native_lib):
    import scenes
    import mapping_loss as ml
    s = scenes.make_scene("C0", keyframe=1)

    class Model:
        pass
    model = Model()
    for nm, k in (("get_xyz", "means3D"), ("get_opacity", "opacities"), ("get_scaling", "scales"),
                  ("get_rotation", "rotations")):
        setattr(model, nm, torch.from_numpy(s[k]).cuda().requires_grad_(True))
    rng = np.random.default_rng(0)
    H, W = s["H"], s["W"]
    cam = Cam(s["viewmatrix"], s["projmatrix"], (10 + rng.standard_normal((1, H, W))).astype(np.float32),
              (rng.uniform(size=(1, H, W)) < 0.9).astype(np.uint8))
    cam.image_height, cam.image_width = H, W
    losses, radii = ml.render_and_losses(cam, model, 0.0, 0.1, 0.1)
    losses["total"].backward()
    assert radii.shape[0] == s["means3D"].shape[0]
    for p in (model.get_xyz, model.get_opacity, model.get_scaling, model.get_rotation):
        assert p.grad is not None and torch.isfinite(p.grad).all() and p.grad.abs().max() > 0
