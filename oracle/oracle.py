"""oracle/oracle.py — TEST INFRASTRUCTURE: numpy front end of the plain-C CPU oracle
(oracle/oracle.c).  May be imported only by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / reference arm; the product path (splat-loam_b200/) never touches it.

    build()                      gcc -O2 -fopenmp -> oracle/_build/liboracle.so
    rasterize_forward(scene)     dict with every intermediate the parity tests compare
    rasterize_backward(fwd, dL)  dict of gradients
    dist2(points)                distCUDA2
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "oracle.c")
_LIB = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None


def build(force=False):
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(_SRC):
        os.makedirs(os.path.dirname(_LIB), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-fopenmp", "-shared", "-fPIC", "-std=gnu99",
                               _SRC, "-o", _LIB, "-lm"])
    return _LIB


def _load():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.orc_build_keys.restype = ctypes.c_long
        _lib.orc_higher_msb.restype = ctypes.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def rasterize_forward(means3D, scales, rotations, opacities, viewmatrix, projmatrix, H, W,
                      scale_modifier=1.0, precomp=None):
    """Reference forward (Rasterizer::forward, DSSR/rasterizer_impl.cu:170-262)."""
    L = _load()
    means3D, opac = _f(means3D), _f(opacities).reshape(-1)
    scales = _f(scales) if scales is not None else None
    rotations = _f(rotations) if rotations is not None else None
    precomp = _f(precomp) if precomp is not None else None
    V, Pm = _f(viewmatrix).reshape(-1), _f(projmatrix).reshape(-1)
    P = means3D.shape[0]
    gx, gy = (W + 15) // 16, (H + 15) // 16
    tiles = gx * gy
    depths = np.zeros(P, np.float32); radii = np.zeros(P, np.int32)
    xy = np.zeros((P, 2), np.float32); T = np.zeros((P, 12), np.float32)
    nrm_opa = np.zeros((P, 4), np.float32); tt = np.zeros(P, np.uint32)
    L.orc_preprocess(P, _p(means3D), _p(scales), _p(rotations), _p(opac), _p(precomp), _p(V),
                     _p(Pm), W, H, ctypes.c_float(scale_modifier), _p(depths), _p(radii), _p(xy),
                     _p(T), _p(nrm_opa), _p(tt))
    N = int(tt.sum(dtype=np.int64))
    keys = np.zeros(max(N, 1), np.uint64); vals = np.zeros(max(N, 1), np.uint32)
    n2 = L.orc_build_keys(P, _p(xy), _p(depths), _p(radii), W, H, _p(keys), _p(vals))
    assert n2 == N
    keys_unsorted, vals_unsorted = keys[:N].copy(), vals[:N].copy()
    ktmp, vtmp = np.empty_like(keys), np.empty_like(vals)
    nbits = 32 + L.orc_higher_msb(ctypes.c_uint32(tiles))
    L.orc_sort_pairs(ctypes.c_long(N), _p(keys), _p(vals), _p(ktmp), _p(vtmp), nbits)
    ranges = np.zeros((tiles, 2), np.uint32)
    L.orc_ranges(ctypes.c_long(N), _p(keys), tiles, _p(ranges))
    final_T = np.zeros((3, H, W), np.float32); n_contrib = np.zeros((2, H, W), np.uint32)
    out = np.zeros((7, H, W), np.float32)
    L.orc_forward(W, H, _p(ranges), _p(vals), _p(xy), _p(T), _p(nrm_opa), _p(Pm), _p(final_T),
                  _p(n_contrib), _p(out))
    return dict(P=P, H=H, W=W, num_rendered=N, depths=depths, radii=radii, means2D=xy, transMat=T,
                normal_opacity=nrm_opa, tiles_touched=tt, point_offsets=np.cumsum(tt, dtype=np.uint32),
                keys_unsorted=keys_unsorted, point_list_unsorted=vals_unsorted, keys=keys[:N],
                point_list=vals[:N], ranges=ranges, final_T=final_T, n_contrib=n_contrib, out=out,
                viewmatrix=V, projmatrix=Pm, scales=scales, rotations=rotations)


def rasterize_backward(fwd, dL_dout):
    """Reference backward (Rasterizer::backward, DSSR/rasterizer_impl.cu:266-313)."""
    L = _load()
    P, H, W = fwd["P"], fwd["H"], fwd["W"]
    dL = _f(dL_dout)
    dT = np.zeros((P, 12), np.float64); dN = np.zeros((P, 3), np.float64); dO = np.zeros(P, np.float64)
    L.orc_backward_render(W, H, _p(fwd["ranges"]), _p(fwd["point_list"]), _p(fwd["means2D"]),
                          _p(fwd["transMat"]), _p(fwd["normal_opacity"]), _p(fwd["projmatrix"]),
                          _p(fwd["final_T"]), _p(fwd["n_contrib"]), _p(dL), _p(dT), _p(dN), _p(dO))
    dTf, dNf = dT.astype(np.float32), dN.astype(np.float32)
    dmeans = np.zeros((P, 3), np.float32); dscales = np.zeros((P, 2), np.float32)
    drots = np.zeros((P, 4), np.float32)
    L.orc_backward_preprocess(P, _p(fwd["radii"]), _p(fwd["scales"]), _p(fwd["rotations"]),
                              _p(fwd["viewmatrix"]), _p(dTf), _p(dNf), _p(dmeans), _p(dscales),
                              _p(drots))
    return dict(means2D=np.zeros((P, 3), np.float32), opacity=dO.astype(np.float32)[:, None],
                means3D=dmeans, transMat=dTf, scales=dscales, rotations=drots, normal=dNf)


def mark_visible(means3D, viewmatrix):
    L = _load()
    m = _f(means3D)
    out = np.zeros(m.shape[0], np.uint8)
    L.orc_mark_visible(m.shape[0], _p(m), _p(_f(viewmatrix).reshape(-1)), _p(out))
    return out.astype(bool)


def dist2(points, q0=0, q1=None):
    """distCUDA2 for query points [q0, q1) against ALL points (brute force, exact)."""
    L = _load()
    pts = _f(points)
    P = pts.shape[0]
    q1 = P if q1 is None else q1
    out = np.zeros(P, np.float32)
    L.orc_knn_range(P, _p(pts), q0, q1, _p(out))
    return out
