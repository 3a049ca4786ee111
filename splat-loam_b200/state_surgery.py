"""Pruning and growing the surfel parameters together with their Adam state (C-ABI
splat_compact_* / splat_append_rows, csrc/state_surgery.cu).

Drop-in counterparts of the optimizer surgery in the reference's GaussianModel
(scene/gaussian_model.py:239-316), which runs once per keyframe (slam/mapper.py:50-138,
216-233): same arguments, same effect on `optimizer.param_groups` / `optimizer.state`, same
returned {group name: new nn.Parameter} dictionary - but every tensor of the model (four
parameters and, where they exist, eight moment tensors) goes through ONE kernel instead of a
boolean-index or torch.cat per tensor.

    optimizable = prune_optimizer(model.optimizer, valid_points_mask)        # _prune_optimizer
    optimizable = cat_tensors_to_optimizer(model.optimizer, {"xyz": ..})     # same name

`GaussianModelSurgery` bundles them as the mix-in methods prune_points /
densification_postfix with the reference's bodies.  No CPU path.
"""
import ctypes

import torch
import torch.nn as nn

import splat_b200_native as _n

_lib = _n.load()


def _ptr_array(tensors):
    return (ctypes.c_void_p * len(tensors))(*[t.data_ptr() if t is not None and t.numel() else None
                                              for t in tensors])


def _check(t, what):
    if not t.is_cuda:
        raise RuntimeError(f"{what} must be a CUDA tensor")
    if t.dtype != torch.float32 or not t.is_contiguous():
        raise RuntimeError(f"{what} must be a contiguous float32 tensor")


def _width(t):
    w = 1
    for d in t.shape[1:]:
        w *= int(d)
    return w


def compact_rows(tensors, keep):
    """[t[keep] for t in tensors] for row-aligned float32 CUDA tensors, in one gather kernel
    (plus the scan of the mask and one host read of the surviving row count)."""
    if not tensors:
        return []
    P, dev = tensors[0].shape[0], tensors[0].device
    for t in tensors:
        _check(t, "tensor")
        if t.shape[0] != P or t.device != dev:
            raise RuntimeError("all tensors must share the row count and the device")
    if keep.dim() != 1 or keep.shape[0] != P or keep.device != dev:
        raise RuntimeError("mask must be a 1-D tensor with one entry per row on the same device")
    keep_u8 = keep.to(torch.bool).contiguous().view(torch.uint8)
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        ws = torch.empty(_lib.splat_compact_workspace_bytes(P), dtype=torch.uint8, device=dev)
        n = ctypes.c_int(0)
        _n.check(_lib.splat_compact_plan(P, keep_u8.data_ptr() if P else None, ws.data_ptr(),
                                         ctypes.byref(n), stream))
        out = [torch.empty((n.value,) + tuple(t.shape[1:]), dtype=torch.float32, device=dev)
               for t in tensors]
        if P and n.value:
            for k in range(0, len(tensors), _n.SURGERY_MAX_TENSORS):
                src, dst = tensors[k:k + _n.SURGERY_MAX_TENSORS], out[k:k + _n.SURGERY_MAX_TENSORS]
                widths = (ctypes.c_int * len(src))(*[_width(t) for t in src])
                _n.check(_lib.splat_compact_rows(P, keep_u8.data_ptr(), ws.data_ptr(), len(src),
                                                 _ptr_array(src), _ptr_array(dst), widths, stream))
    return out


def append_rows(tensors, extensions):
    """[cat(t, e) for t, e in zip(tensors, extensions)] in one kernel; an extension given as an
    int appends that many zero rows (what a new point's Adam moments start from)."""
    if not tensors:
        return []
    dev = tensors[0].device
    n_old = tensors[0].shape[0]
    n_new = None
    for t, e in zip(tensors, extensions):
        _check(t, "tensor")
        m = e if isinstance(e, int) else e.shape[0]
        if not isinstance(e, int):
            _check(e, "extension")
            if tuple(e.shape[1:]) != tuple(t.shape[1:]) or e.device != dev:
                raise RuntimeError("an extension must match its tensor's row shape and device")
        if t.shape[0] != n_old or (n_new is not None and m != n_new):
            raise RuntimeError("all tensors / extensions must share their row counts")
        n_new = m
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        out = [torch.empty((n_old + n_new,) + tuple(t.shape[1:]), dtype=torch.float32, device=dev)
               for t in tensors]
        for k in range(0, len(tensors), _n.SURGERY_MAX_TENSORS):
            sl = slice(k, k + _n.SURGERY_MAX_TENSORS)
            src, dst = tensors[sl], out[sl]
            ext = [None if isinstance(e, int) else e for e in extensions[sl]]
            widths = (ctypes.c_int * len(src))(*[_width(t) for t in src])
            _n.check(_lib.splat_append_rows(n_old, n_new, len(src), _ptr_array(src),
                                            _ptr_array(ext), _ptr_array(dst), widths, stream))
    return out


def prune_optimizer(optimizer, mask):
    """GaussianModel._prune_optimizer (scene/gaussian_model.py:239-257): keep the rows where
    `mask` is True.  Reproduced as written, including what it does to the state: the pruned
    moments are computed but then NOT re-attached to the new parameter (the reference stores
    the new parameter under the group's NAME in optimizer.state, :250), so Adam starts the
    surviving points from fresh moments at its next step.  Hence only the parameters need
    to move."""
    groups = optimizer.param_groups
    new = compact_rows([g["params"][0].detach() for g in groups], mask)
    optimizable = {}
    for g, t in zip(groups, new):
        old = g["params"][0]
        had_state = optimizer.state.get(old, None) is not None
        if had_state:
            del optimizer.state[old]
        g["params"][0] = nn.Parameter(t.requires_grad_(True))
        if had_state:
            optimizer.state[g["name"]] = g["params"][0]
        optimizable[g["name"]] = g["params"][0]
    return optimizable


def cat_tensors_to_optimizer(optimizer, tensors_dict):
    """GaussianModel.cat_tensors_to_optimizer (scene/gaussian_model.py:268-300): append the new
    points' rows to every parameter and zero rows to the moments that exist."""
    groups = optimizer.param_groups
    tensors, exts, slots = [], [], []
    for g in groups:
        assert len(g["params"]) == 1
        p = g["params"][0]
        e = tensors_dict[g["name"]]
        tensors.append(p.detach())
        exts.append(e.detach().float().contiguous())
        slots.append((g, "param"))
        st = optimizer.state.get(p)
        if st is not None:
            for nm in ("exp_avg", "exp_avg_sq"):
                tensors.append(st[nm])
                exts.append(int(e.shape[0]))
                slots.append((g, nm))
    new = append_rows(tensors, exts)
    optimizable = {}
    fresh = {}
    for (g, what), t in zip(slots, new):
        fresh.setdefault(id(g), {})[what] = t
    for g in groups:
        p = g["params"][0]
        st = optimizer.state.get(p)
        f = fresh[id(g)]
        if st is not None:
            st["exp_avg"], st["exp_avg_sq"] = f["exp_avg"], f["exp_avg_sq"]
            del optimizer.state[p]
            g["params"][0] = nn.Parameter(f["param"].requires_grad_(True))
            optimizer.state[g["params"][0]] = st
        else:
            g["params"][0] = nn.Parameter(f["param"].requires_grad_(True))
        optimizable[g["name"]] = g["params"][0]
    return optimizable


class GaussianModelSurgery:
    """Mix-in with the reference's method names and bodies (scene/gaussian_model.py:259-316)
    over the fused kernels: `class FastModel(GaussianModelSurgery, GaussianModel): pass`."""

    def _prune_optimizer(self, mask):
        return prune_optimizer(self.optimizer, mask)

    def cat_tensors_to_optimizer(self, tensors_dict):
        return cat_tensors_to_optimizer(self.optimizer, tensors_dict)

    def prune_points(self, mask):
        optimizable = self._prune_optimizer(~mask)
        self._xyz, self._opacity = optimizable["xyz"], optimizable["opacity"]
        self._scaling, self._rotation = optimizable["scaling"], optimizable["rotation"]

    def densification_postfix(self, new_xyz, new_opacity, new_scaling, new_rotation):
        optimizable = self.cat_tensors_to_optimizer(
            {"xyz": new_xyz, "opacity": new_opacity, "scaling": new_scaling,
             "rotation": new_rotation})
        self._xyz, self._opacity = optimizable["xyz"], optimizable["opacity"]
        self._scaling, self._rotation = optimizable["scaling"], optimizable["rotation"]
