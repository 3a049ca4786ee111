"""Activation backward + Adam for the surfel parameters in one kernel (C-ABI
splat_activation_adam_step, csrc/adam_step.cu).

The reference's mapping iteration ends with `loss_total.backward(); optimizer.step()`
(slam/mapper.py:196-204): autograd walks back through the four activations of
scene/gaussian_model.py:58-74 and torch.optim.Adam (eps=1e-15, four groups,
gaussian_model.py:97-121) updates the raw parameters - some 30 kernels.  When the gradients
w.r.t. the ACTIVATED parameters are already in the packed [P,10] buffer (what dssr_backward
writes and KeyframeWindow.gradients returns), this does all of it in one pass:

    g = window.gradients((m.get_xyz, m.get_opacity, m.get_scaling, m.get_rotation), cams, up)
    activation_adam_step(m.optimizer, g, scaling_max=0.5, scaling_max_penalty=0.2)

The optimizer object stays a torch.optim.Adam: its `state[p]["exp_avg"|"exp_avg_sq"|"step"]`
entries are created and updated exactly where torch keeps them, so the reference's
densification / pruning code (gaussian_model.py:223-316) keeps working on it.  No CPU path.
"""
import ctypes

import torch

import splat_b200_native as _n

_lib = _n.load()
GROUPS = ("xyz", "opacity", "scaling", "rotation")     # gaussian_model.py:99-119
_WIDTH = {"xyz": 3, "opacity": 1, "scaling": 2, "rotation": 4}


def _group(optimizer, name):
    for g in optimizer.param_groups:
        if g.get("name") == name:
            if len(g["params"]) != 1:
                raise RuntimeError(f"param group '{name}' must hold exactly one tensor")
            return g
    raise RuntimeError(f"optimizer has no param group named '{name}'")


def activation_adam_step(optimizer, packed_grads, scaling_max=0.0, scaling_max_penalty=0.0,
                         reg_out=None):
    """One Adam step of the four surfel parameter groups from `packed_grads` ([P,10] float32
    CUDA: dL/d(xyz 3 | opacity 1 | scaling 2 | rotation 4), activated values).
    scaling_max / scaling_max_penalty: the regulariser of slam/mapper.py:186-191 (its gradient
    is added inside; `reg_out`, a 1-element float32 CUDA tensor, is incremented by its value)."""
    if not isinstance(optimizer, torch.optim.Adam):
        raise RuntimeError("activation_adam_step drives a torch.optim.Adam instance")
    if not packed_grads.is_cuda:
        raise RuntimeError("packed_grads must be a CUDA tensor")
    if packed_grads.dtype != torch.float32 or not packed_grads.is_contiguous() or \
            packed_grads.dim() != 2 or packed_grads.shape[1] != 10:
        raise RuntimeError("packed_grads must be a contiguous float32 tensor of shape [P,10]")
    P, dev = packed_grads.shape[0], packed_grads.device
    a = _n.SplatAdamArgs()
    a.P = P
    a.packed_grads = packed_grads.data_ptr() if P else None
    betas = eps = None
    keep = []
    for k, name in enumerate(GROUPS):
        g = _group(optimizer, name)
        if g.get("weight_decay", 0) or g.get("amsgrad", False) or g.get("maximize", False):
            raise RuntimeError("weight_decay / amsgrad / maximize are not supported")
        if betas is None:
            betas, eps = tuple(g["betas"]), g["eps"]
        elif tuple(g["betas"]) != betas or g["eps"] != eps:
            raise RuntimeError("all four groups must share betas and eps")
        p = g["params"][0]
        if p.device != dev or p.dtype != torch.float32 or not p.is_contiguous() or \
                p.numel() != P * _WIDTH[name]:
            raise RuntimeError(f"parameter '{name}' must be contiguous float32 [{P},{_WIDTH[name]}] "
                               "on the gradients' device")
        st = optimizer.state[p]
        if len(st) == 0:                        # torch's lazy state initialisation
            st["step"] = torch.tensor(0.0, dtype=torch.float32)
            st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        st["step"] += 1          # a CPU scalar tensor unless the optimizer is capturable
        step = int(st["step"])
        for nm in ("exp_avg", "exp_avg_sq"):
            if not st[nm].is_contiguous() or st[nm].shape != p.shape:
                raise RuntimeError(f"optimizer state '{nm}' of '{name}' must match the parameter")
        a.param[k] = p.data_ptr() if P else None
        a.exp_avg[k] = st["exp_avg"].data_ptr() if P else None
        a.exp_avg_sq[k] = st["exp_avg_sq"].data_ptr() if P else None
        a.lr[k] = float(g["lr"])
        a.step[k] = step
        keep.append((p, st))
    a.beta1, a.beta2, a.eps = float(betas[0]), float(betas[1]), float(eps)
    a.scaling_max, a.scaling_max_penalty = float(scaling_max), float(scaling_max_penalty)
    a.reg_out = reg_out.data_ptr() if reg_out is not None else None
    with torch.cuda.device(dev):
        a.stream = torch.cuda.current_stream(dev).cuda_stream
        if P:
            _n.check(_lib.splat_activation_adam_step(ctypes.byref(a)))
