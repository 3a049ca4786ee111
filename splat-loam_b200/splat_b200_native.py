"""ctypes loader for the in-tree C-ABI library (splat-loam_b200/lib/libsplat_b200.so).

There is deliberately NO fallback: if the CUDA library is missing the import fails.
Declarations mirror include/splat_b200.h one to one.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libsplat_b200.so")

c_float_p = ctypes.c_void_p  # raw device pointers travel as integers
ALLOC_FN = ctypes.CFUNCTYPE(ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t)


class DssrForwardArgs(ctypes.Structure):
    _fields_ = [
        ("P", ctypes.c_int), ("W", ctypes.c_int), ("H", ctypes.c_int),
        ("scale_modifier", ctypes.c_float),
        ("prefiltered", ctypes.c_int), ("debug", ctypes.c_int),
        ("means3D", c_float_p), ("opacities", c_float_p), ("scales", c_float_p),
        ("rotations", c_float_p), ("transMat_precomp", c_float_p),
        ("viewmatrix", c_float_p), ("projmatrix", c_float_p),
        ("out_maps", c_float_p), ("radii", ctypes.c_void_p),
        ("geom_alloc", ALLOC_FN), ("geom_ctx", ctypes.c_void_p),
        ("binning_alloc", ALLOC_FN), ("binning_ctx", ctypes.c_void_p),
        ("image_alloc", ALLOC_FN), ("image_ctx", ctypes.c_void_p),
        ("binning_hint", ctypes.c_int64),
        ("stream", ctypes.c_void_p),
    ]


class DssrBackwardArgs(ctypes.Structure):
    _fields_ = [
        ("P", ctypes.c_int), ("W", ctypes.c_int), ("H", ctypes.c_int), ("R", ctypes.c_int),
        ("scale_modifier", ctypes.c_float), ("debug", ctypes.c_int),
        ("means3D", c_float_p), ("scales", c_float_p), ("rotations", c_float_p),
        ("transMat_precomp", c_float_p),
        ("viewmatrix", c_float_p), ("projmatrix", c_float_p),
        ("radii", ctypes.c_void_p),
        ("geom_buffer", ctypes.c_void_p), ("binning_buffer", ctypes.c_void_p),
        ("image_buffer", ctypes.c_void_p),
        ("dL_dout", c_float_p),
        ("dL_dmeans2D", c_float_p), ("dL_dnormal", c_float_p), ("dL_dopacity", c_float_p),
        ("dL_dmeans3D", c_float_p), ("dL_dtransMat", c_float_p), ("dL_dscales", c_float_p),
        ("dL_drotations", c_float_p),
        ("packed_grads", c_float_p), ("packed_accumulate", ctypes.c_int),
        ("xchg", ctypes.c_void_p),
        ("stream", ctypes.c_void_p),
    ]


class DssrPostForwardArgs(ctypes.Structure):
    _fields_ = [
        ("W", ctypes.c_int), ("H", ctypes.c_int), ("depth_ratio", ctypes.c_float),
        ("allmap", c_float_p), ("viewmatrix", c_float_p), ("projmatrix", c_float_p),
        ("rend_normal", c_float_p), ("surf_depth", c_float_p), ("surf_normal", c_float_p),
        ("stream", ctypes.c_void_p),
    ]


class DssrPostBackwardArgs(ctypes.Structure):
    _fields_ = [
        ("W", ctypes.c_int), ("H", ctypes.c_int), ("depth_ratio", ctypes.c_float),
        ("allmap", c_float_p), ("viewmatrix", c_float_p), ("projmatrix", c_float_p),
        ("dL_drend_normal", c_float_p), ("dL_dsurf_depth", c_float_p),
        ("dL_dsurf_normal", c_float_p), ("dL_dallmap", c_float_p),
        ("stream", ctypes.c_void_p),
    ]


class DssrPostLossForwardArgs(ctypes.Structure):
    _fields_ = [
        ("W", ctypes.c_int), ("H", ctypes.c_int), ("depth_ratio", ctypes.c_float),
        ("allmap", c_float_p), ("viewmatrix", c_float_p), ("projmatrix", c_float_p),
        ("gt_depth", c_float_p), ("gt_valid", ctypes.c_void_p),
        ("lambda_normal", ctypes.c_float), ("lambda_alpha", ctypes.c_float),
        ("rend_normal", c_float_p), ("surf_depth", c_float_p), ("surf_normal", c_float_p),
        ("losses", c_float_p), ("workspace", ctypes.c_void_p), ("stream", ctypes.c_void_p),
    ]


class DssrPostLossBackwardArgs(ctypes.Structure):
    _fields_ = [
        ("W", ctypes.c_int), ("H", ctypes.c_int), ("depth_ratio", ctypes.c_float),
        ("allmap", c_float_p), ("viewmatrix", c_float_p), ("projmatrix", c_float_p),
        ("gt_depth", c_float_p), ("gt_valid", ctypes.c_void_p),
        ("lambda_normal", ctypes.c_float), ("lambda_alpha", ctypes.c_float),
        ("losses", c_float_p), ("dL_dlosses", c_float_p), ("dL_dallmap", c_float_p),
        ("stream", ctypes.c_void_p),
    ]


class SplatAdamArgs(ctypes.Structure):
    _fields_ = [
        ("P", ctypes.c_int), ("packed_grads", c_float_p),
        ("param", c_float_p * 4), ("exp_avg", c_float_p * 4), ("exp_avg_sq", c_float_p * 4),
        ("lr", ctypes.c_double * 4), ("step", ctypes.c_longlong * 4),
        ("beta1", ctypes.c_double), ("beta2", ctypes.c_double), ("eps", ctypes.c_double),
        ("scaling_max", ctypes.c_double), ("scaling_max_penalty", ctypes.c_double),
        ("reg_out", c_float_p), ("stream", ctypes.c_void_p),
    ]


class GsaPhotoParams(ctypes.Structure):
    _fields_ = [
        ("H", ctypes.c_int), ("W", ctypes.c_int),
        ("fx", ctypes.c_float), ("fy", ctypes.c_float), ("cx", ctypes.c_float), ("cy", ctypes.c_float),
        ("min_depth", ctypes.c_float), ("max_depth", ctypes.c_float),
        ("omega_depth", ctypes.c_float), ("depth_rejection_threshold", ctypes.c_float),
        ("kernel_chi_threshold", ctypes.c_float), ("stream", ctypes.c_void_p),
    ]


POST_LOSS_WORKSPACE_BYTES = 64

# every symbol include/splat_b200.h declares
EXPORTS = (
    "dssr_forward", "dssr_backward", "dssr_mark_visible", "dssr_geometry_bytes",
    "dssr_image_bytes", "dssr_binning_bytes", "dssr_debug_state", "sknn_workspace_bytes",
    "sknn_dist2", "splat_last_error", "splat_abi_version", "splat_profile_enable",
    "splat_profile_read", "splat_profile_select", "sknn_dist2_static",
    "splat_xchg_create", "splat_xchg_export", "splat_xchg_connect", "splat_xchg_connect_local",
    "splat_xchg_push", "splat_xchg_reduce", "splat_xchg_wait", "splat_xchg_allreduce",
    "splat_xchg_result", "splat_xchg_destroy", "dssr_post_forward", "dssr_post_backward",
    "dssr_post_loss_forward", "dssr_post_loss_backward", "splat_activation_adam_step",
    "splat_xchg_fault", "splat_launch_count", "splat_compact_workspace_bytes",
    "splat_compact_plan", "splat_compact_rows", "splat_append_rows",
    "gsa_photo_workspace_bytes", "gsa_photo_set_reference", "gsa_photo_set_query",
    "gsa_photo_align", "gsa_photo_debug_reference",
    "rimg_workspace_bytes", "rimg_spherical_intrinsics", "rimg_project",
)
XCHG_HANDLE_BYTES = 64
SURGERY_MAX_TENSORS = 12

_lib = None


def load():
    """Load the library once; raise if it has not been built (no CPU fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python splat-loam_b200/build.py` "
            "(nvcc, sm_100a). There is no CPU or PyTorch fallback for this path.")
    lib = ctypes.CDLL(LIB_PATH)
    lib.dssr_forward.argtypes = [ctypes.POINTER(DssrForwardArgs), ctypes.POINTER(ctypes.c_int)]
    lib.dssr_forward.restype = ctypes.c_int
    lib.dssr_backward.argtypes = [ctypes.POINTER(DssrBackwardArgs)]
    lib.dssr_backward.restype = ctypes.c_int
    lib.dssr_mark_visible.argtypes = [ctypes.c_int, c_float_p, c_float_p, c_float_p,
                                      ctypes.c_void_p, ctypes.c_void_p]
    lib.dssr_mark_visible.restype = ctypes.c_int
    lib.dssr_geometry_bytes.argtypes = [ctypes.c_int]
    lib.dssr_geometry_bytes.restype = ctypes.c_size_t
    lib.dssr_image_bytes.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.dssr_image_bytes.restype = ctypes.c_size_t
    lib.dssr_binning_bytes.argtypes = [ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                       ctypes.c_int]
    lib.dssr_binning_bytes.restype = ctypes.c_size_t
    lib.dssr_debug_state.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                     ctypes.c_char_p, c_float_p,
                                     ctypes.POINTER(ctypes.c_void_p),
                                     ctypes.POINTER(ctypes.c_size_t), ctypes.c_void_p]
    lib.dssr_debug_state.restype = ctypes.c_int
    lib.sknn_workspace_bytes.argtypes = [ctypes.c_int]
    lib.sknn_workspace_bytes.restype = ctypes.c_size_t
    lib.sknn_dist2.argtypes = [ctypes.c_int, c_float_p, c_float_p, ctypes.c_void_p,
                               ctypes.c_size_t, ctypes.c_void_p]
    lib.sknn_dist2.restype = ctypes.c_int
    lib.sknn_dist2_static.argtypes = [ctypes.c_int, c_float_p, c_float_p, ctypes.c_void_p]
    lib.sknn_dist2_static.restype = ctypes.c_int
    lib.splat_last_error.argtypes = []
    lib.splat_last_error.restype = ctypes.c_char_p
    lib.splat_abi_version.argtypes = []
    lib.splat_abi_version.restype = ctypes.c_int
    lib.splat_profile_enable.argtypes = [ctypes.c_int]
    lib.splat_profile_enable.restype = ctypes.c_int
    lib.splat_profile_select.argtypes = [ctypes.c_char_p]
    lib.splat_profile_select.restype = ctypes.c_int
    lib.splat_profile_read.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_char_p),
                                       ctypes.POINTER(ctypes.c_float)]
    lib.splat_profile_read.restype = ctypes.c_int
    lib.dssr_post_forward.argtypes = [ctypes.POINTER(DssrPostForwardArgs)]
    lib.dssr_post_forward.restype = ctypes.c_int
    lib.dssr_post_backward.argtypes = [ctypes.POINTER(DssrPostBackwardArgs)]
    lib.dssr_post_backward.restype = ctypes.c_int
    lib.dssr_post_loss_forward.argtypes = [ctypes.POINTER(DssrPostLossForwardArgs)]
    lib.dssr_post_loss_forward.restype = ctypes.c_int
    lib.dssr_post_loss_backward.argtypes = [ctypes.POINTER(DssrPostLossBackwardArgs)]
    lib.dssr_post_loss_backward.restype = ctypes.c_int
    lib.splat_activation_adam_step.argtypes = [ctypes.POINTER(SplatAdamArgs)]
    lib.splat_activation_adam_step.restype = ctypes.c_int
    vp = ctypes.c_void_p
    lib.splat_xchg_create.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(vp)]
    lib.splat_xchg_export.argtypes = [vp, ctypes.c_char_p]
    lib.splat_xchg_connect.argtypes = [vp, ctypes.c_char_p]
    lib.splat_xchg_connect_local.argtypes = [vp, ctypes.POINTER(vp)]
    lib.splat_xchg_push.argtypes = [vp, c_float_p, vp]
    for f in (lib.splat_xchg_reduce, lib.splat_xchg_wait, lib.splat_xchg_allreduce):
        f.argtypes = [vp, vp]
    lib.splat_xchg_result.argtypes = [vp]
    lib.splat_xchg_result.restype = vp
    lib.splat_xchg_destroy.argtypes = [vp]
    lib.splat_xchg_fault.argtypes = [vp]
    lib.splat_xchg_fault.restype = ctypes.c_int
    pp = ctypes.POINTER(ctypes.c_void_p)
    ip = ctypes.POINTER(ctypes.c_int)
    lib.splat_compact_workspace_bytes.argtypes = [ctypes.c_int]
    lib.splat_compact_workspace_bytes.restype = ctypes.c_size_t
    lib.splat_compact_plan.argtypes = [ctypes.c_int, vp, vp, ip, vp]
    lib.splat_compact_plan.restype = ctypes.c_int
    lib.splat_compact_rows.argtypes = [ctypes.c_int, vp, vp, ctypes.c_int, pp, pp, ip, vp]
    lib.splat_compact_rows.restype = ctypes.c_int
    lib.splat_append_rows.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, pp, pp, pp, ip, vp]
    lib.splat_append_rows.restype = ctypes.c_int
    fl = ctypes.c_float
    lib.gsa_photo_workspace_bytes.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.gsa_photo_workspace_bytes.restype = ctypes.c_size_t
    lib.gsa_photo_set_reference.argtypes = [vp, ctypes.c_int, ctypes.c_int, vp, fl, ctypes.c_int, vp]
    lib.gsa_photo_set_reference.restype = ctypes.c_int
    lib.gsa_photo_set_query.argtypes = [vp, ctypes.c_int, ctypes.c_int, vp, fl, fl, fl, fl, fl, fl, vp]
    lib.gsa_photo_set_query.restype = ctypes.c_int
    lib.gsa_photo_debug_reference.argtypes = [vp, ctypes.c_int, ctypes.c_int, vp, vp]
    lib.gsa_photo_debug_reference.restype = ctypes.c_int
    lib.gsa_photo_align.argtypes = [vp, ctypes.POINTER(GsaPhotoParams), ctypes.POINTER(fl),
                                    ctypes.c_int, ctypes.POINTER(fl), ctypes.POINTER(fl),
                                    ctypes.POINTER(ctypes.c_double)]
    lib.gsa_photo_align.restype = ctypes.c_int
    lib.rimg_workspace_bytes.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.rimg_workspace_bytes.restype = ctypes.c_size_t
    lib.rimg_spherical_intrinsics.argtypes = [ctypes.c_int, vp, ctypes.c_int, ctypes.c_int, vp, vp, vp]
    lib.rimg_spherical_intrinsics.restype = ctypes.c_int
    lib.rimg_project.argtypes = [ctypes.c_int, vp, ctypes.c_int, ctypes.c_int, vp, fl, fl, vp, vp, vp,
                                 vp, vp, vp]
    lib.rimg_project.restype = ctypes.c_int
    lib.splat_launch_count.argtypes = []
    lib.splat_launch_count.restype = ctypes.c_ulonglong
    for f in (lib.splat_xchg_create, lib.splat_xchg_export, lib.splat_xchg_connect,
              lib.splat_xchg_connect_local, lib.splat_xchg_push, lib.splat_xchg_reduce,
              lib.splat_xchg_wait, lib.splat_xchg_allreduce, lib.splat_xchg_destroy):
        f.restype = ctypes.c_int
    _lib = lib
    return lib


def profile_enable(on=True, only=None):
    """Bracket every stage (or just `only`) of the calling thread's frames with CUDA events."""
    check(load().splat_profile_select(only.encode() if only else None))
    check(load().splat_profile_enable(int(bool(on))))


def profile_read():
    """{stage: milliseconds summed over the calls since the last read} (synchronise first)."""
    names = (ctypes.c_char_p * 16)()
    ms = (ctypes.c_float * 16)()
    n = load().splat_profile_read(16, names, ms)
    return {names[i].decode(): float(ms[i]) for i in range(n)}


def launch_count():
    """Kernels launched by the library so far in this process."""
    return int(load().splat_launch_count())


def last_error():
    msg = load().splat_last_error()
    return msg.decode(errors="replace") if msg else "unknown error"


def check(rc):
    if rc != 0:
        raise RuntimeError(last_error())


_cudart = None


def copy_from_device_ptr(dst_tensor, src_ptr, nbytes):
    """cudaMemcpy(device -> device) from a raw pointer into a torch tensor (debug only)."""
    global _cudart
    if _cudart is None:
        for name in ("libcudart.so.12", "libcudart.so"):
            try:
                _cudart = ctypes.CDLL(name)
                break
            except OSError:
                continue
        if _cudart is None:
            raise RuntimeError("libcudart not found")
        _cudart.cudaMemcpy.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t,
                                       ctypes.c_int]
        _cudart.cudaMemcpy.restype = ctypes.c_int
    rc = _cudart.cudaMemcpy(dst_tensor.data_ptr(), src_ptr, nbytes, 3)
    if rc != 0:
        raise RuntimeError(f"cudaMemcpy failed with {rc}")
