"""CPU: the torch restatement of render()'s map post-processing (oracle/render_post.py) against
golden outputs of the UNMODIFIED reference render() (tests/golden/post_*.npz)."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import render_post as rp

CASES = ("post_c0", "post_c0_ratio", "post_synth", "post_synth_mix")


def _inputs(g, dtype):
    t = lambda k: torch.from_numpy(np.asarray(g[k])).to(dtype)  # noqa: E731
    return t("allmap"), t("viewmatrix"), t("projmatrix"), float(g["depth_ratio"])


@pytest.mark.parametrize("name", CASES)
def test_restatement_reproduces_the_reference_in_fp32(name):
    g = load_golden(name)
    a, V, Pm, r = _inputs(g, torch.float32)
    out = rp.post_forward(a, V, Pm, r)
    for k in rp.OUTPUTS:
        ref = torch.from_numpy(g[k])
        assert out[k].shape == ref.shape
        assert float((out[k] - ref).abs().max()) <= 1e-6 * max(1.0, float(ref.abs().max())), k
    grads = {k: torch.from_numpy(g["grad_" + k]) for k in rp.OUTPUTS}
    gb = rp.post_backward(a, V, Pm, r, grads)
    ref = torch.from_numpy(g["dL_dallmap"])
    assert float((gb - ref).norm() / ref.norm()) < 1e-6


@pytest.mark.parametrize("name", CASES)
def test_fp64_truth_is_within_the_stated_tolerance_of_the_reference(name):
    """The tolerance the CUDA kernels are held to (tests/test_gpu_post.py) is the distance
    between the reference's own fp32 result and the fp64 evaluation of the same formulas."""
    g = load_golden(name)
    a, V, Pm, r = _inputs(g, torch.float64)
    out = rp.post_forward(a, V, Pm, r)
    for k in rp.OUTPUTS:
        ref = torch.from_numpy(g[k]).double()
        assert float((out[k] - ref).abs().max()) <= 2e-5 * max(1.0, float(ref.abs().max())), k
    grads = {k: torch.from_numpy(g["grad_" + k]).double() for k in rp.OUTPUTS}
    gb = rp.post_backward(a, V, Pm, r, grads)
    ref = torch.from_numpy(g["dL_dallmap"]).double()
    assert float((gb - ref).norm() / ref.norm()) < 1e-5


def test_border_and_empty_pixels():
    g = load_golden("post_synth")
    sn = g["surf_normal"]
    assert not sn[:, 0].any() and not sn[:, -1].any() and not sn[:, :, 0].any() and not sn[:, :, -1].any()
    hole = g["allmap"][1] == 0
    assert hole.any() and not sn[:, hole].any()
    lens = np.linalg.norm(sn[:, 1:-1, 1:-1], axis=0)
    np.testing.assert_allclose(lens, g["allmap"][1][1:-1, 1:-1], atol=1e-5)


@pytest.mark.parametrize("name", ("post_loss_c0", "post_loss_synth"))
def test_loss_restatement_against_golden(name):
    """The loss goldens come from the reference's render() + the restated loss lines in fp32;
    re-evaluated here in fp32 (must reproduce) and fp64 (the tolerance the CUDA path gets)."""
    g = load_golden(name)
    for dtype, tol_l, tol_g in ((torch.float32, 1e-6, 1e-6), (torch.float64, 1e-5, 1e-4)):
        a, V, Pm, r = _inputs(g, dtype)
        a.requires_grad_(True)
        pkg = rp.post_forward(a, V, Pm, r)
        losses = rp.mapper_image_losses(pkg, torch.from_numpy(g["gt_depth"]).to(dtype),
                                        torch.from_numpy(g["gt_valid"]),
                                        float(g["lambda_normal"]), float(g["lambda_alpha"]))
        for got, want in zip(losses, g["losses"]):
            assert abs(float(got.detach()) - want) <= tol_l * max(1.0, abs(want))
        losses[3].backward()
        ref = torch.from_numpy(g["dL_dallmap"]).to(dtype)
        assert float((a.grad - ref).norm() / ref.norm()) < tol_g
