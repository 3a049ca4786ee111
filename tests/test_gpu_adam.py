"""GPU: the fused activation-backward + Adam step (C-ABI splat_activation_adam_step through
splat-loam_b200/fused_adam.py) against what the reference runs: autograd through the
activations of scene/gaussian_model.py:58-74 plus the scale regulariser of
slam/mapper.py:186-191, then torch.optim.Adam(eps=1e-15).step().  torch IS the reference
implementation here, so it is the oracle.  Tolerance: 2e-6 relative L2 per tensor after 1 step,
1e-5 after 10 (fp32 rounding differs in the last bit per operation)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
LRS = {"xyz": 1.6e-4, "opacity": 5e-2, "scaling": 5e-3, "rotation": 1e-3}
THR, PEN = 0.5, 0.2


def make_model(P, seed, ties=True):
    g = torch.Generator(device="cuda").manual_seed(seed)
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)  # noqa: E731
    raw = {"xyz": 5 * r(P, 3), "opacity": 2 * r(P, 1), "scaling": -1.0 + r(P, 2), "rotation": r(P, 4)}
    if ties and P >= 8:       # create_from_pcd starts with equal scales; some above the threshold
        raw["scaling"][: P // 4, 1] = raw["scaling"][: P // 4, 0]
        raw["scaling"][:3] = 0.1
    return raw


def make_opt(raw):
    params = {k: torch.nn.Parameter(v.clone()) for k, v in raw.items()}
    opt = torch.optim.Adam([{"params": [params[k]], "lr": LRS[k], "name": k} for k in params],
                           lr=0.0, eps=1e-15)
    return params, opt


def torch_step(params, opt, G, thr=THR, pen=PEN):
    opt.zero_grad(set_to_none=True)
    act = (params["xyz"], torch.sigmoid(params["opacity"]), torch.exp(params["scaling"]),
           torch.nn.functional.normalize(params["rotation"]))
    loss = sum((a * g).sum() for a, g in zip(act, (G[:, 0:3], G[:, 3:4], G[:, 4:6], G[:, 6:10])))
    smax = act[2].max(dim=1).values
    reg = (pen * (smax[smax >= thr] - thr)).sum()
    (loss + reg).backward()
    opt.step()
    return float(reg.detach())


def rel(a, b):
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.mark.parametrize("P", [1, 255, 4097, 100000])
def test_matches_autograd_plus_torch_adam(native_lib, P):
    from fused_adam import activation_adam_step
    # no exactly tied scales here: after one Adam step tied columns differ by a few ulp, and
    # which of them is "the max" (and receives the regulariser's gradient) is then decided by
    # the last bit of the arithmetic - in the reference as much as here.  Ties: next test.
    raw = make_model(P, P, ties=False)
    pa, oa = make_opt(raw)
    pb, ob = make_opt(raw)
    g = torch.Generator(device="cuda").manual_seed(99)
    reg_out = torch.zeros(1, device="cuda")
    for it in range(10):
        G = torch.randn(P, 10, device="cuda", generator=g) * (10.0 ** torch.randint(-3, 2, (1,)).item())
        reg = torch_step(pa, oa, G)
        reg_out.zero_()
        activation_adam_step(ob, G, THR, PEN, reg_out=reg_out)
        tol = 2e-6 if it == 0 else 1e-5
        for k in raw:
            assert rel(pb[k].data, pa[k].data) < tol, (k, it)
            sa, sb = oa.state[pa[k]], ob.state[pb[k]]
            assert rel(sb["exp_avg"], sa["exp_avg"]) < tol and rel(sb["exp_avg_sq"], sa["exp_avg_sq"]) < tol
            assert float(sa["step"]) == float(sb["step"]) == it + 1
        assert abs(float(reg_out) - reg) <= 1e-5 * max(1.0, abs(reg))


def test_tied_scales_follow_torch_max(native_lib):
    """create_from_pcd starts every surfel with two EQUAL scales (gaussian_model.py:84): the
    regulariser's gradient must go to the first column, as torch.max(dim=1) does."""
    from fused_adam import activation_adam_step
    P = 4096
    raw = make_model(P, 7, ties=False)
    raw["scaling"][:, 1] = raw["scaling"][:, 0]
    pa, oa = make_opt(raw)
    pb, ob = make_opt(raw)
    G = 1e-3 * torch.randn(P, 10, device="cuda")      # small next to the penalty: it decides signs
    torch_step(pa, oa, G)
    activation_adam_step(ob, G, THR, PEN)
    assert (torch.exp(raw["scaling"][:, 0]) >= THR).sum() > 100
    for k in raw:
        assert rel(pb[k].data, pa[k].data) < 2e-6, k
        assert rel(ob.state[pb[k]]["exp_avg"], oa.state[pa[k]]["exp_avg"]) < 2e-6, k


def test_state_is_torchs_own(native_lib):
    """torch steps, fused steps and torch steps again on the same optimizer object; pruning the
    state tensors like the reference's densification does keeps working."""
    from fused_adam import activation_adam_step
    P = 3000
    raw = make_model(P, 5, ties=False)
    pa, oa = make_opt(raw)
    pb, ob = make_opt(raw)
    g = torch.Generator(device="cuda").manual_seed(3)
    for it in range(6):
        G = torch.randn(P, 10, device="cuda", generator=g)
        torch_step(pa, oa, G)
        if it % 2 == 0:
            torch_step(pb, ob, G)
        else:
            activation_adam_step(ob, G, THR, PEN)
    for k in raw:
        assert rel(pb[k].data, pa[k].data) < 1e-5
    # prune half of the surfels the way gaussian_model.py:_prune_optimizer does
    keep = torch.arange(P, device="cuda") % 2 == 0
    for opt, params in ((oa, pa), (ob, pb)):
        for grp in opt.param_groups:
            old = grp["params"][0]
            st = opt.state.pop(old)
            st["exp_avg"], st["exp_avg_sq"] = st["exp_avg"][keep], st["exp_avg_sq"][keep]
            new = torch.nn.Parameter(old.data[keep].clone())
            grp["params"][0] = new
            opt.state[new] = st
            params[grp["name"]] = new
    G = torch.randn(int(keep.sum()), 10, device="cuda", generator=g)
    torch_step(pa, oa, G)
    activation_adam_step(ob, G, THR, PEN)
    for k in raw:
        assert rel(pb[k].data, pa[k].data) < 1e-5


def test_argument_checks(native_lib):
    from fused_adam import activation_adam_step
    raw = make_model(16, 1)
    p, o = make_opt(raw)
    with pytest.raises(RuntimeError, match="CUDA"):
        activation_adam_step(o, torch.zeros(16, 10))
    with pytest.raises(RuntimeError, match=r"\[P,10\]"):
        activation_adam_step(o, torch.zeros(16, 9, device="cuda"))
    with pytest.raises(RuntimeError, match="must be contiguous float32"):
        activation_adam_step(o, torch.zeros(17, 10, device="cuda"))
    bad = torch.optim.Adam([{"params": [p["xyz"]], "name": "xyz"}], lr=1e-3)
    with pytest.raises(RuntimeError, match="no param group named"):
        activation_adam_step(bad, torch.zeros(16, 10, device="cuda"))
    activation_adam_step(o, torch.zeros(16, 10, device="cuda"))    # and a valid call
