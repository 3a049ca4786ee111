"""ctypes access to oracle/_ref/libref_splat.so — the UNMODIFIED reference CUDA rasterizer
and simple-knn built by oracle/Makefile (test infrastructure; never imported by the
product).  Only usable on a machine with a GPU.  The reference launches on the legacy
default stream and synchronises, so callers need no stream handling.
"""
import ctypes
import os

import torch

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_LIB = os.path.join(_ROOT, "oracle", "_ref", "libref_splat.so")

_STATE_DTYPES = {
    "depths": torch.float32, "means2D": torch.float32, "transMat": torch.float32,
    "normal_opacity": torch.float32, "tiles_touched": torch.int32,
    "point_offsets": torch.int32, "keys_unsorted": torch.int64, "keys": torch.int64,
    "point_list_unsorted": torch.int32, "point_list": torch.int32, "ranges": torch.int32,
    "final_T": torch.float32, "n_contrib": torch.int32,
}


def available():
    return os.path.exists(REF_LIB) and torch.cuda.is_available()


_lib = None
_cudart = None


def _copy(dst, src_ptr, nbytes):
    global _cudart
    if _cudart is None:
        for name in ("libcudart.so.12", "libcudart.so"):
            try:
                _cudart = ctypes.CDLL(name)
                break
            except OSError:
                continue
        _cudart.cudaMemcpy.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t,
                                       ctypes.c_int]
        _cudart.cudaMemcpy.restype = ctypes.c_int
    rc = _cudart.cudaMemcpy(dst.data_ptr(), src_ptr, nbytes, 3)
    if rc != 0:
        raise RuntimeError(f"cudaMemcpy failed with {rc}")


def _load():
    global _lib
    if _lib is None:
        lib = ctypes.CDLL(REF_LIB)
        vp = ctypes.c_void_p
        lib.ref_dssr_create.restype = vp
        lib.ref_dssr_destroy.argtypes = [vp]
        fwd_args = [vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, vp, vp, vp, ctypes.c_float,
                    vp, vp, vp, vp, ctypes.c_int, vp, vp]
        lib.ref_dssr_forward.argtypes = fwd_args
        lib.ref_dssr_forward.restype = ctypes.c_int
        lib.ref_dssr_forward_nosync.argtypes = fwd_args
        lib.ref_dssr_forward_nosync.restype = ctypes.c_int
        bwd_args = [vp, vp, vp, ctypes.c_float, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]
        lib.ref_dssr_backward.argtypes = bwd_args
        lib.ref_dssr_backward.restype = ctypes.c_int
        lib.ref_dssr_backward_nosync.argtypes = bwd_args
        lib.ref_dssr_backward_nosync.restype = ctypes.c_int
        lib.ref_dssr_state.argtypes = [vp, ctypes.c_char_p, ctypes.POINTER(vp),
                                       ctypes.POINTER(ctypes.c_size_t)]
        lib.ref_dssr_state.restype = ctypes.c_int
        lib.ref_dssr_mark_visible.argtypes = [ctypes.c_int, vp, vp, vp, vp]
        lib.ref_dssr_mark_visible.restype = ctypes.c_int
        lib.ref_sknn_dist2.argtypes = [ctypes.c_int, vp, vp]
        lib.ref_sknn_dist2.restype = ctypes.c_int
        lib.ref_sknn_dist2_nosync.argtypes = [ctypes.c_int, vp, vp]
        lib.ref_sknn_dist2_nosync.restype = ctypes.c_int
        _lib = lib
    return _lib


def _p(t):
    return t.data_ptr() if t is not None and t.numel() else None


class RefRasterizer:
    """One reference rasterizer instance (owns its geometry/binning/image chunks)."""

    def __init__(self):
        self.lib = _load()
        self.h = self.lib.ref_dssr_create()
        self.P = 0

    def __del__(self):
        try:
            self.lib.ref_dssr_destroy(self.h)
        except Exception:
            pass

    def forward(self, means3D, opacities, scales, rotations, viewmatrix, projmatrix, H, W,
                scale_modifier=1.0, precomp=None, prefiltered=False, sync=True):
        means3D, opacities, viewmatrix, projmatrix = (
            t.contiguous() for t in (means3D, opacities, viewmatrix, projmatrix))
        scales = scales.contiguous() if scales is not None else None
        rotations = rotations.contiguous() if rotations is not None else None
        precomp = precomp.contiguous() if precomp is not None else None
        P = means3D.shape[0]
        self.P, self.H, self.W = P, H, W
        dev = means3D.device
        out = torch.empty((7, H, W), dtype=torch.float32, device=dev)
        radii = torch.empty((P,), dtype=torch.int32, device=dev)
        self._keep = (means3D, opacities, scales, rotations, viewmatrix, projmatrix, precomp)
        fn = self.lib.ref_dssr_forward if sync else self.lib.ref_dssr_forward_nosync
        torch.cuda.current_stream().synchronize() if sync else None
        R = fn(self.h, P, W, H, _p(means3D), _p(opacities), _p(scales), float(scale_modifier),
               _p(rotations), _p(precomp), _p(viewmatrix), _p(projmatrix), int(prefiltered),
               _p(out), _p(radii))
        if R < 0:
            raise RuntimeError("reference forward failed")
        self.radii = radii
        return R, out, radii

    def backward(self, dL_dout, scale_modifier=1.0, sync=True):
        means3D, opacities, scales, rotations, viewmatrix, projmatrix, precomp = self._keep
        P, dev = self.P, means3D.device
        g = {k: torch.empty((P, c), dtype=torch.float32, device=dev)
             for k, c in (("means2D", 3), ("normal", 3), ("opacity", 1), ("means3D", 3),
                          ("transMat", 12), ("scales", 2), ("rotations", 4))}
        fn = self.lib.ref_dssr_backward if sync else self.lib.ref_dssr_backward_nosync
        rc = fn(self.h, _p(means3D), _p(scales), float(scale_modifier), _p(rotations),
                _p(precomp), _p(viewmatrix), _p(projmatrix), _p(self.radii), _p(dL_dout),
                _p(g["means2D"]), _p(g["normal"]), _p(g["opacity"]), _p(g["means3D"]),
                _p(g["transMat"]), _p(g["scales"]), _p(g["rotations"]))
        if rc != 0:
            raise RuntimeError("reference backward failed")
        return g

    def state(self, name):
        ptr, cnt = ctypes.c_void_p(), ctypes.c_size_t()
        rc = self.lib.ref_dssr_state(self.h, name.encode(), ctypes.byref(ptr), ctypes.byref(cnt))
        if rc != 0:
            raise RuntimeError(f"ref_dssr_state({name}) -> {rc}")
        out = torch.empty(cnt.value, dtype=_STATE_DTYPES[name], device="cuda")
        if cnt.value:
            torch.cuda.synchronize()
            _copy(out, ptr.value, cnt.value * out.element_size())
        return out.cpu()


def mark_visible(means3D, viewmatrix, projmatrix):
    P = means3D.shape[0]
    present = torch.zeros(P, dtype=torch.bool, device=means3D.device)
    torch.cuda.synchronize()
    rc = _load().ref_dssr_mark_visible(P, _p(means3D), _p(viewmatrix), _p(projmatrix),
                                       _p(present))
    if rc != 0:
        raise RuntimeError("reference mark_visible failed")
    return present


def dist2(points, sync=True):
    P = points.shape[0]
    out = torch.empty(P, dtype=torch.float32, device=points.device)
    torch.cuda.synchronize()
    fn = _load().ref_sknn_dist2 if sync else _load().ref_sknn_dist2_nosync
    rc = fn(P, _p(points), _p(out))
    if rc != 0:
        raise RuntimeError("reference distCUDA2 failed")
    return out


# ---- the reference's photometric registration (oracle/_ref/libref_photo.so) ----------------
REF_PHOTO_LIB = os.path.join(_ROOT, "oracle", "_ref", "libref_photo.so")
_photo = None


def photo_available():
    return os.path.exists(REF_PHOTO_LIB) and torch.cuda.is_available()


class RefPhotometric:
    """PhotometricFac + ImageSet of the reference's gsaligner submodule, configured as
    GSAligner configures them (oracle/ref_photo_shim.cu)."""

    def __init__(self, H, W, omega_depth=5.0, depth_rejection_threshold=0.8, rho_kernel=1.0,
                 min_depth=0.5, max_depth=80.0):
        global _photo
        if _photo is None:
            lib = ctypes.CDLL(REF_PHOTO_LIB)
            vp, fl = ctypes.c_void_p, ctypes.c_float
            lib.ref_photo_create.restype = vp
            lib.ref_photo_create.argtypes = [ctypes.c_int, ctypes.c_int, fl, fl, fl, fl, fl]
            lib.ref_photo_destroy.argtypes = [vp]
            lib.ref_photo_set_reference.argtypes = [vp, vp, ctypes.POINTER(fl)]
            lib.ref_photo_set_query.argtypes = [vp, vp]
            lib.ref_photo_align.argtypes = [vp, ctypes.POINTER(fl), ctypes.c_int,
                                            ctypes.POINTER(fl), ctypes.POINTER(fl)]
            lib.ref_photo_linearize.argtypes = [vp, ctypes.POINTER(fl), ctypes.POINTER(ctypes.c_double)]
            lib.ref_photo_reference_image.argtypes = [vp, ctypes.POINTER(fl)]
            _photo = lib
        self.lib, self.H, self.W = _photo, H, W
        self.h = self.lib.ref_photo_create(H, W, omega_depth, depth_rejection_threshold,
                                           rho_kernel, min_depth, max_depth)

    def __del__(self):
        try:
            self.lib.ref_photo_destroy(self.h)
        except Exception:
            pass

    def set_reference(self, depth, K):
        torch.cuda.synchronize()
        self._ref = depth.contiguous().float()
        k = (ctypes.c_float * 9)(*[float(x) for x in K.reshape(-1)])
        assert self.lib.ref_photo_set_reference(self.h, self._ref.data_ptr(), k) == 0

    def set_query(self, depth):
        torch.cuda.synchronize()
        self._query = depth.contiguous().float()
        assert self.lib.ref_photo_set_query(self.h, self._query.data_ptr()) == 0

    def align(self, T, iterations):
        t = (ctypes.c_float * 16)(*[float(x) for x in T.reshape(-1)])
        out, stats = (ctypes.c_float * 16)(), (ctypes.c_float * 2)()
        assert self.lib.ref_photo_align(self.h, t, iterations, out, stats) == 0
        return torch.tensor(list(out)).reshape(4, 4), float(stats[1]), float(stats[0])

    def linearize(self, T):
        t = (ctypes.c_float * 16)(*[float(x) for x in T.reshape(-1)])
        out = (ctypes.c_double * 30)()
        assert self.lib.ref_photo_linearize(self.h, t, out) == 0
        return list(out)

    def reference_image(self):
        out = (ctypes.c_float * (4 * self.H * self.W))()
        assert self.lib.ref_photo_reference_image(self.h, out) == 0
        return torch.tensor(list(out)).reshape(4, self.H, self.W)
