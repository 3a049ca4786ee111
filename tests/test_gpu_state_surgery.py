"""GPU: pruning / growing the parameters and their Adam state (splat_compact_* /
splat_append_rows through state_surgery.py) against the reference's own GaussianModel methods
(scene/gaussian_model.py:239-316, staged unmodified into oracle/_ref/py/scene/ by
oracle/Makefile; its heavier imports - plyfile, the config and logging modules - are stubbed,
none of them is touched by the methods under test).  A reference model and a twin that uses
the fused kernels go through the same densify -> step -> prune -> step sequences; parameters,
moments, step counters and the layout of optimizer.state must stay equal (bit for bit: the
surgery only moves floats)."""
import importlib.util
import os
import sys
import types

import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu
PY = os.path.join(ROOT, "oracle", "_ref", "py")
NAMES = ("xyz", "opacity", "scaling", "rotation")
WIDTH = {"xyz": 3, "opacity": 1, "scaling": 2, "rotation": 4}


def _reference_model_class():
    path = os.path.join(PY, "scene", "gaussian_model.py")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/py/scene/gaussian_model.py not staged")
    saved = {}

    def stub(name, **attrs):
        saved[name] = sys.modules.get(name)
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m

    f = lambda *a, **k: None  # noqa: E731
    stub("utils")
    stub("utils.general_utils", build_scaling_rotation=f, inverse_sigmoid=f,
         matrix_to_quaternion=f, create_rotation_matrix_from_direction_vector_batch=f)
    stub("utils.logging_utils", get_logger=lambda name: None)
    stub("utils.graphic_utils", BasicPointCloud=object)
    stub("utils.config_utils", Configuration=object)
    stub("plyfile", PlyData=object, PlyElement=object)
    try:
        import simple_knn._C  # noqa: F401  (the drop-in's distCUDA2 satisfies the import)
        spec = importlib.util.spec_from_file_location("ref_gaussian_model", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return mod.GaussianModel


class _Cfg:
    class opt:
        position_lr, opacity_lr, scaling_lr, rotation_lr = 1.6e-4, 5e-2, 5e-3, 1e-3


def _make(cls, P, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    m = cls(device="cuda")
    m._xyz = torch.nn.Parameter(torch.randn(P, 3, device="cuda", generator=g))
    m._opacity = torch.nn.Parameter(torch.randn(P, 1, device="cuda", generator=g))
    m._scaling = torch.nn.Parameter(torch.randn(P, 2, device="cuda", generator=g))
    m._rotation = torch.nn.Parameter(torch.randn(P, 4, device="cuda", generator=g))
    m.training_setup(_Cfg)
    return m


def _params(m):
    return {"xyz": m._xyz, "opacity": m._opacity, "scaling": m._scaling, "rotation": m._rotation}


def _step(m, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    m.optimizer.zero_grad(set_to_none=True)
    for p in _params(m).values():
        p.grad = torch.randn(p.shape, device="cuda", generator=g)
    m.optimizer.step()


def _assert_same(a, b):
    pa, pb = _params(a), _params(b)
    for k in NAMES:
        assert pa[k].shape == pb[k].shape, k
        assert torch.equal(pa[k], pb[k]), k
        assert isinstance(pb[k], torch.nn.Parameter) and pb[k].requires_grad and pb[k].is_leaf
    for ga, gb in zip(a.optimizer.param_groups, b.optimizer.param_groups):
        assert ga["name"] == gb["name"] and ga["lr"] == gb["lr"]
        assert ga["params"][0] is pa[ga["name"]] and gb["params"][0] is pb[gb["name"]]
        sa, sb = a.optimizer.state.get(ga["params"][0]), b.optimizer.state.get(gb["params"][0])
        assert (sa is None or len(sa) == 0) == (sb is None or len(sb) == 0), ga["name"]
        if sa:
            assert float(sa["step"]) == float(sb["step"])
            for nm in ("exp_avg", "exp_avg_sq"):
                assert torch.equal(sa[nm], sb[nm]), (ga["name"], nm)
    # the reference leaves an entry keyed by the group NAME behind after a prune (:250)
    ka = sorted(k for k in a.optimizer.state if isinstance(k, str))
    kb = sorted(k for k in b.optimizer.state if isinstance(k, str))
    assert ka == kb


def _extension(n, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return {k: torch.randn(n, WIDTH[k], device="cuda", generator=g) for k in NAMES}


@pytest.mark.parametrize("P,n_new", [(5000, 700), (1, 3), (70001, 1), (3000, 0)])
def test_densify_step_prune_sequences_match_the_reference(native_lib, P, n_new):
    from state_surgery import GaussianModelSurgery
    Ref = _reference_model_class()

    class Fast(GaussianModelSurgery, Ref):
        pass

    a, b = _make(Ref, P, 1), _make(Fast, P, 1)
    _assert_same(a, b)
    # growing BEFORE the first step: the optimizer has no state yet
    e = _extension(n_new, 2)
    a.densification_postfix(e["xyz"], e["opacity"], e["scaling"], e["rotation"])
    b.densification_postfix(e["xyz"], e["opacity"], e["scaling"], e["rotation"])
    _assert_same(a, b)
    for s in (3, 4):
        _step(a, s), _step(b, s)
    _assert_same(a, b)
    # growing with state: moments get zero rows
    e = _extension(n_new + 5, 5)
    a.densification_postfix(e["xyz"], e["opacity"], e["scaling"], e["rotation"])
    b.densification_postfix(e["xyz"], e["opacity"], e["scaling"], e["rotation"])
    _assert_same(a, b)
    _step(a, 6), _step(b, 6)
    _assert_same(a, b)
    # pruning (mask = True where a point goes away)
    n = a._xyz.shape[0]
    g = torch.Generator(device="cuda").manual_seed(7)
    mask = torch.rand(n, device="cuda", generator=g) < 0.3
    a.prune_points(mask), b.prune_points(mask)
    assert a._xyz.shape[0] == int((~mask).sum())
    _assert_same(a, b)
    for s in (8, 9):
        _step(a, s), _step(b, s)
    _assert_same(a, b)
    # prune everything but one row, then nothing, then grow again
    n = a._xyz.shape[0]
    mask = torch.ones(n, dtype=torch.bool, device="cuda")
    mask[n // 2] = False
    a.prune_points(mask), b.prune_points(mask)
    _assert_same(a, b)
    mask = torch.zeros(1, dtype=torch.bool, device="cuda")
    a.prune_points(mask), b.prune_points(mask)
    _assert_same(a, b)
    e = _extension(11, 10)
    a.densification_postfix(e["xyz"], e["opacity"], e["scaling"], e["rotation"])
    b.densification_postfix(e["xyz"], e["opacity"], e["scaling"], e["rotation"])
    _step(a, 11), _step(b, 11)
    _assert_same(a, b)


def test_compact_and_append_primitives(native_lib):
    """The kernels themselves: 12 tensors at once, widths 1-4, against torch indexing / cat;
    wrong arguments are errors."""
    from state_surgery import append_rows, compact_rows
    P = 200003
    g = torch.Generator(device="cuda").manual_seed(0)
    ts = [torch.randn(P, w, device="cuda", generator=g) for w in (3, 1, 2, 4) * 3] + \
         [torch.randn(P, device="cuda", generator=g)]          # 13 tensors: two kernel calls
    keep = torch.rand(P, device="cuda", generator=g) < 0.61
    for got, t in zip(compact_rows(ts, keep), ts):
        assert torch.equal(got, t[keep])
    assert compact_rows(ts, torch.zeros(P, dtype=torch.bool, device="cuda"))[0].shape == (0, 3)
    full = compact_rows(ts[:2], torch.ones(P, dtype=torch.bool, device="cuda"))
    assert torch.equal(full[0], ts[0]) and torch.equal(full[1], ts[1])
    ext = [torch.randn((777,) + tuple(t.shape[1:]), device="cuda", generator=g) for t in ts]
    for got, t, e in zip(append_rows(ts, ext), ts, ext):
        assert torch.equal(got, torch.cat((t, e)))
    for got, t in zip(append_rows(ts, [5] * len(ts)), ts):
        assert torch.equal(got[:P], t) and not got[P:].any() and got.shape[0] == P + 5
    with pytest.raises(RuntimeError):
        compact_rows([ts[0].cpu()], keep.cpu())
    with pytest.raises(RuntimeError):
        compact_rows(ts[:2], keep[:-1])
    with pytest.raises(RuntimeError):
        append_rows(ts[:2], [ext[0], ext[1][:-1]])
