"""`_C` of the drop-in package: same three functions, argument order and tuple layouts
as the reference's pybind module (submodules/diff-surfel-spherical-rasterization/
ext.cpp:5-9, rasterize_points.cu:32-165), implemented as a thin ctypes layer over the
C-ABI (include/splat_b200.h).  PyTorch is used for device memory and streams only.
"""
import ctypes
import os

import torch

import splat_b200_native as _n

_lib = _n.load()

# Capacity hints for the binning buffer, keyed by (device, H, W): with a hint the
# whole frame is enqueued before the host reads the instance count, so the per-forward
# host<->device round trip of the reference (rasterizer_impl.cu:216-219) does not stall
# the GPU.  SPLAT_B200_HINT=0 restores the exact-size two-phase behaviour.
_USE_HINT = os.environ.get("SPLAT_B200_HINT", "1") != "0"
_hints = {}


def _check_cuda(t, name):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")


def _f32(t):
    t = t.contiguous()
    if t.dtype != torch.float32:
        t = t.float()
    return t


def _ptr(t):
    return t.data_ptr() if t.numel() else None


# Allocator callbacks (resizeFunctional, rasterize_points.cu:24-30).  ONE module-level
# C callback; `ctx` is a key into _slots.  (A per-call bound-method callback would form a
# reference cycle and keep the buffers alive until the cyclic GC runs, which forces the
# caching allocator to cudaMalloc on every frame.)
_slots = {}
_next_slot = [1]


def _alloc(ctx, nbytes):
    slot = _slots[ctx]
    t = slot[0]
    if t.numel() < nbytes:
        t = torch.empty(int(nbytes), dtype=torch.uint8, device=t.device)
        slot[0] = t
    return t.data_ptr()


_ALLOC_CB = _n.ALLOC_FN(_alloc)


def _new_slot(device):
    key = _next_slot[0]
    _next_slot[0] = key + 1 if key < (1 << 40) else 1
    _slots[key] = [torch.empty(0, dtype=torch.uint8, device=device)]
    return key


def rasterize_gaussians(means3D, opacity, scales, rotations, scale_modifier, transMat_precomp,
                        viewmatrix, projmatrix, image_height, image_width, prefiltered, debug):
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    for t, nm in ((means3D, "means3D"), (opacity, "opacity"), (scales, "scales"),
                  (rotations, "rotations"), (transMat_precomp, "transMat_precomp"),
                  (viewmatrix, "viewmatrix"), (projmatrix, "projmatrix")):
        _check_cuda(t, nm)
    P, H, W = int(means3D.size(0)), int(image_height), int(image_width)
    dev = means3D.device
    with torch.cuda.device(dev):
        out = torch.empty((7, H, W), dtype=torch.float32, device=dev)
        radii = torch.empty((P,), dtype=torch.int32, device=dev)
        kg, kb, ki = _new_slot(dev), _new_slot(dev), _new_slot(dev)
        means3D_c, opacity_c = _f32(means3D), _f32(opacity)
        scales_c, rot_c, pre_c = _f32(scales), _f32(rotations), _f32(transMat_precomp)
        view_c, proj_c = _f32(viewmatrix), _f32(projmatrix)
        key = (dev.index, H, W)     # (P changes with every densification: not part of the key)
        hint = _hints.get(key, 0) if _USE_HINT else 0
        a = _n.DssrForwardArgs(
            P, W, H, float(scale_modifier), int(bool(prefiltered)), int(bool(debug)),
            _ptr(means3D_c), _ptr(opacity_c), _ptr(scales_c), _ptr(rot_c), _ptr(pre_c),
            _ptr(view_c), _ptr(proj_c), _ptr(out), _ptr(radii),
            _ALLOC_CB, kg, _ALLOC_CB, kb, _ALLOC_CB, ki, hint,
            torch.cuda.current_stream(dev).cuda_stream)
        n = ctypes.c_int(0)
        try:
            rc = _lib.dssr_forward(ctypes.byref(a), ctypes.byref(n))
        finally:
            geom, binning, img = _slots.pop(kg)[0], _slots.pop(kb)[0], _slots.pop(ki)[0]
        _n.check(rc)
        if _USE_HINT:
            _hints[key] = int(n.value * 1.25) + 4096
    return n.value, out, radii, geom, binning, img


def rasterize_gaussians_backward(means3D, radii, scales, rotations, scale_modifier,
                                 transMat_precomp, viewmatrix, projmatrix, dL_dout, geomBuffer,
                                 R, binningBuffer, imageBuffer, debug, packed=None,
                                 packed_accumulate=False, xchg=None):
    """Reference signature (rasterize_points.cu:90-147).  Extension: `packed`, a contiguous
    float32 [P,10] tensor that additionally receives (or, with packed_accumulate, is
    incremented by) dL_dmeans3D | dL_dopacity | dL_dscales | dL_drotations - the message of
    the keyframe-window all-reduce; `xchg`, the handle of a connected peer exchange
    (keyframe_window.PeerExchange): the same rows are pushed to their owner ranks by the VJP
    kernel itself."""
    for t, nm in ((means3D, "means3D"), (radii, "radii"), (scales, "scales"),
                  (rotations, "rotations"), (transMat_precomp, "transMat_precomp"),
                  (viewmatrix, "viewmatrix"), (projmatrix, "projmatrix"),
                  (binningBuffer, "binningBuffer"), (imageBuffer, "imageBuffer"),
                  (geomBuffer, "geomBuffer")):
        _check_cuda(t, nm)
    P = int(means3D.size(0))
    H, W = int(dL_dout.size(1)), int(dL_dout.size(2))
    dev = means3D.device
    if packed is not None and (packed.dtype != torch.float32 or not packed.is_contiguous()
                               or tuple(packed.shape) != (P, 10) or packed.device != dev):
        raise RuntimeError("packed must be a contiguous float32 CUDA tensor of shape [P,10]")
    with torch.cuda.device(dev):
        # The reference does seven torch::zeros (rasterize_points.cu:116-122); the kernels
        # here write every element themselves, so torch.empty.  What autograd hands to a leaf
        # as .grad gets a tensor of its own (a view of one big block would keep the whole block
        # alive for as long as the leaf keeps its .grad); the two per-surfel intermediates
        # (dL_dnormal, dL_dtransMat) share one.
        f32 = dict(dtype=torch.float32, device=dev)
        dmeans3D, dmeans2D = torch.empty((P, 3), **f32), torch.empty((P, 3), **f32)
        dopacity, dscales = torch.empty((P, 1), **f32), torch.empty((P, 2), **f32)
        drot = torch.empty((P, 4), **f32)
        inner = torch.empty(15 * P, **f32)
        dtrans, dnormal = inner[:12 * P].view(P, 12), inner[12 * P:].view(P, 3)
        if P:
            means3D_c, scales_c, rot_c = _f32(means3D), _f32(scales), _f32(rotations)
            pre_c, view_c, proj_c = _f32(transMat_precomp), _f32(viewmatrix), _f32(projmatrix)
            dout_c = _f32(dL_dout)
            radii_c = radii.contiguous()
            a = _n.DssrBackwardArgs(
                P, W, H, int(R), float(scale_modifier), int(bool(debug)),
                _ptr(means3D_c), _ptr(scales_c), _ptr(rot_c), _ptr(pre_c),
                _ptr(view_c), _ptr(proj_c), _ptr(radii_c),
                _ptr(geomBuffer), _ptr(binningBuffer), _ptr(imageBuffer), _ptr(dout_c),
                _ptr(dmeans2D), _ptr(dnormal), _ptr(dopacity), _ptr(dmeans3D), _ptr(dtrans),
                _ptr(dscales), _ptr(drot),
                _ptr(packed) if packed is not None else None, int(bool(packed_accumulate)),
                xchg, torch.cuda.current_stream(dev).cuda_stream)
            _n.check(_lib.dssr_backward(ctypes.byref(a)))
    return dmeans2D, dopacity, dmeans3D, dtrans, dscales, drot


def mark_visible(means3D, viewmatrix, projmatrix):
    P = int(means3D.size(0))
    present = torch.zeros((P,), dtype=torch.bool, device=means3D.device)
    if P:
        with torch.cuda.device(means3D.device):
            m, v, p = _f32(means3D), _f32(viewmatrix), _f32(projmatrix)
            _n.check(_lib.dssr_mark_visible(P, _ptr(m), _ptr(v), _ptr(p), present.data_ptr(),
                                            torch.cuda.current_stream().cuda_stream))
    return present


def debug_state(geomBuffer, binningBuffer, imageBuffer, name):
    """Parity-test accessor (dssr_debug_state): returns a CPU copy of an internal array."""
    import numpy as np
    dev = geomBuffer.device
    with torch.cuda.device(dev):
        hdr = geomBuffer.cpu().numpy()
        # P lives in the header; scratch for the unpacked transMat / normal_opacity
        off = (-geomBuffer.data_ptr()) % 256
        P = int(np.frombuffer(hdr[off + 4:off + 8].tobytes(), dtype=np.int32)[0])
        scratch = torch.empty(16 * max(P, 1), dtype=torch.float32, device=dev)
        ptr, cnt = ctypes.c_void_p(), ctypes.c_size_t()
        _n.check(_lib.dssr_debug_state(
            _ptr(geomBuffer), _ptr(binningBuffer) if binningBuffer is not None else None,
            _ptr(imageBuffer), name.encode(), scratch.data_ptr(), ctypes.byref(ptr),
            ctypes.byref(cnt), torch.cuda.current_stream().cuda_stream))
        dtype = {"depths": torch.float32, "means2D": torch.float32, "transMat": torch.float32,
                 "normal_opacity": torch.float32, "final_T": torch.float32,
                 "keys": torch.int64, "keys_unsorted": torch.int64}.get(name, torch.int32)
        n = cnt.value
        out = torch.empty(n, dtype=dtype, device=dev)
        if n:
            torch.cuda.synchronize()
            _n.copy_from_device_ptr(out, ptr.value, n * out.element_size())
        return out.cpu()
