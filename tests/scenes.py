"""Seeded synthetic inputs for the rasterizer and the kNN.

numpy only (Generator(PCG64) streams are identical on every machine), float32 outputs.
Used by the tests, by bench.py and by the golden-fixture generator.
"""
import math

import numpy as np

CONFIGS = {
    # name: (H, W, vfov_deg, P, keyframes)
    "C0": (16, 256, 30.0, 2000, 1),        # tiny: CPU-oracle sized
    "C1": (16, 1024, 30.0, 20000, 1),
    "C2": (64, 1024, 33.2, 500000, 1),
    "C3": (64, 2048, 26.9, 1000000, 8),
    "C4": (128, 2048, 45.0, 2000000, 32),
    "C5": (128, 2048, 45.0, 4000000, 1),   # dense depth/normal render stress (4 M surfels)
}


def intrinsics(H, W, vfov_deg):
    """K as in DSSR/tests/generate_input.py:27-47; projmatrix = eye(4) with [:3,:3]=K^T
    (scene/cameras.py:48-50)."""
    vfov = math.radians(vfov_deg)
    K = np.array([[-W / (2 * math.pi), 0, W * 0.5], [0, -H / vfov, H * 0.5], [0, 0, 1]],
                 dtype=np.float64)
    proj = np.eye(4)
    proj[:3, :3] = K.T
    return np.ascontiguousarray(proj, dtype=np.float32)


def pose(k):
    """world_T_lidar of keyframe k: 0.5 m steps along x, 2 degrees of yaw per step."""
    yaw = math.radians(2.0 * k)
    T = np.eye(4)
    T[:3, :3] = [[math.cos(yaw), -math.sin(yaw), 0], [math.sin(yaw), math.cos(yaw), 0], [0, 0, 1]]
    T[:3, 3] = [0.5 * k, 0, 0]
    return T


def viewmatrix(k=0):
    """inv(world_T_lidar) transposed (scene/cameras.py:43-46)."""
    return np.ascontiguousarray(np.linalg.inv(pose(k)).T, dtype=np.float32)


def yaw_viewmatrix(deg):
    """View matrix of a sensor at the origin rotated by `deg` about its vertical axis."""
    a = math.radians(deg)
    T = np.eye(4)
    T[:3, :3] = [[math.cos(a), -math.sin(a), 0], [math.sin(a), math.cos(a), 0], [0, 0, 1]]
    return np.ascontiguousarray(np.linalg.inv(T).T, dtype=np.float32)


def _rotmat_to_quat(R):
    """(w,x,y,z) of a batch of rotation matrices [n,3,3] (Shepperd's method)."""
    n = R.shape[0]
    q = np.empty((n, 4))
    tr = R[:, 0, 0] + R[:, 1, 1] + R[:, 2, 2]
    c0 = tr > 0
    c1 = (~c0) & (R[:, 0, 0] >= R[:, 1, 1]) & (R[:, 0, 0] >= R[:, 2, 2])
    c2 = (~c0) & (~c1) & (R[:, 1, 1] >= R[:, 2, 2])
    c3 = ~(c0 | c1 | c2)
    s = np.sqrt(np.maximum(tr[c0] + 1.0, 1e-12)) * 2
    q[c0] = np.stack([0.25 * s, (R[c0, 2, 1] - R[c0, 1, 2]) / s, (R[c0, 0, 2] - R[c0, 2, 0]) / s,
                      (R[c0, 1, 0] - R[c0, 0, 1]) / s], 1)
    s = np.sqrt(np.maximum(1.0 + R[c1, 0, 0] - R[c1, 1, 1] - R[c1, 2, 2], 1e-12)) * 2
    q[c1] = np.stack([(R[c1, 2, 1] - R[c1, 1, 2]) / s, 0.25 * s, (R[c1, 0, 1] + R[c1, 1, 0]) / s,
                      (R[c1, 0, 2] + R[c1, 2, 0]) / s], 1)
    s = np.sqrt(np.maximum(1.0 + R[c2, 1, 1] - R[c2, 0, 0] - R[c2, 2, 2], 1e-12)) * 2
    q[c2] = np.stack([(R[c2, 0, 2] - R[c2, 2, 0]) / s, (R[c2, 0, 1] + R[c2, 1, 0]) / s, 0.25 * s,
                      (R[c2, 1, 2] + R[c2, 2, 1]) / s], 1)
    s = np.sqrt(np.maximum(1.0 + R[c3, 2, 2] - R[c3, 0, 0] - R[c3, 1, 1], 1e-12)) * 2
    q[c3] = np.stack([(R[c3, 1, 0] - R[c3, 0, 1]) / s, (R[c3, 0, 2] + R[c3, 2, 0]) / s,
                      (R[c3, 1, 2] + R[c3, 2, 1]) / s, 0.25 * s], 1)
    return q / np.linalg.norm(q, axis=1, keepdims=True)


def surfels(P, H, W, vfov_deg, seed=0):
    """Random surfel cloud around the sensor; returns dict of float32 arrays."""
    rng = np.random.default_rng(seed)
    vfov = math.radians(vfov_deg)
    az = rng.uniform(-math.pi, math.pi, P)
    el = rng.uniform(-vfov / 2, vfov / 2, P)
    out = rng.random(P) < 0.10                      # 10 %: up to 5 deg outside the band
    el[out] = np.sign(el[out]) * (vfov / 2 + rng.uniform(0, math.radians(5.0), out.sum()))
    r = np.exp(rng.uniform(math.log(3.0), math.log(60.0), P))
    u = rng.random(P)
    near, far = u < 0.005, (u >= 0.005) & (u < 0.01)  # 1 %: outside (0.1, 120]
    r[near] = rng.uniform(0.05, 0.1, near.sum())
    r[far] = rng.uniform(120.0, 130.0, far.sum())
    d = np.stack([np.cos(az) * np.cos(el), np.sin(az) * np.cos(el), np.sin(el)], 1)
    means = d * r[:, None]
    # normal = -dir rotated by a random axis-angle of at most 60 deg; 5 % flipped
    axis = rng.normal(size=(P, 3))
    axis -= (axis * d).sum(1, keepdims=True) * d
    axis /= np.linalg.norm(axis, axis=1, keepdims=True)
    ang = rng.uniform(0, math.radians(60.0), P)
    nrm = -d * np.cos(ang)[:, None] + np.cross(axis, -d) * np.sin(ang)[:, None]
    flip = rng.random(P) < 0.05
    nrm[flip] *= -1
    # frame with third axis = normal, random in-plane rotation
    t = np.cross(nrm, axis)
    t /= np.linalg.norm(t, axis=1, keepdims=True)
    b = np.cross(nrm, t)
    phi = rng.uniform(0, 2 * math.pi, P)
    e0 = t * np.cos(phi)[:, None] + b * np.sin(phi)[:, None]
    e1 = np.cross(nrm, e0)
    R = np.stack([e0, e1, nrm], 2)                   # columns
    quat = _rotmat_to_quat(R)
    pix = 2 * math.pi / W
    scales = np.clip(r[:, None] * pix * rng.uniform(0.5, 2.0, (P, 2)), 0.02, 0.3)
    opa = rng.uniform(0.3, 0.99, P)
    opa[rng.random(P) < 0.02] = 1.0
    return {
        "means3D": means.astype(np.float32),
        "scales": scales.astype(np.float32),
        "rotations": quat.astype(np.float32),
        "opacities": opa.astype(np.float32)[:, None],
    }


def upstream_grad(H, W, seed=1):
    return np.random.default_rng(seed).normal(size=(7, H, W)).astype(np.float32)


def make_scene(name_or_dims, seed=0, keyframe=0, P=None):
    """Full input set for one frame: surfels + matrices + dL_dout."""
    if isinstance(name_or_dims, str):
        H, W, vfov, P0, _ = CONFIGS[name_or_dims]
    else:
        H, W, vfov, P0 = name_or_dims
    P = P0 if P is None else P
    s = surfels(P, H, W, vfov, seed)
    s.update(H=H, W=W, viewmatrix=viewmatrix(keyframe), projmatrix=intrinsics(H, W, vfov),
             dL_dout=upstream_grad(H, W, seed + 1))
    return s


def reference_test_scene():
    """The one fixture the reference ships: DSSR/tests/generate_input.py:5-68 — a 32x32
    grid of surfels on the z=0 plane, 2048x512 image, hfov 2*pi, vfov pi/2."""
    n, length = 32, 10.0
    x = np.linspace(-1, 1, n) * length
    xx, yy = np.meshgrid(x, x)
    means = np.stack([xx, yy, np.zeros_like(xx)], -1).reshape(-1, 3)
    P = means.shape[0]
    quat = np.zeros((P, 4)); quat[:, 0] = 1.0
    scales = np.full((P, 2), length / (n - 1))
    opa = np.ones((P, 1))
    W, H = 2048, 512
    hfov, vfov = 2 * math.pi, math.pi / 2
    K = np.array([[-W / hfov, 0, W * 0.5], [0, -H / vfov, H * 0.5], [0, 0, 1]], dtype=np.float32)
    proj = np.zeros((4, 4), dtype=np.float32)
    proj[:3, :3] = K
    proj = proj.T.copy()
    world_T_lidar = np.array([[0, -1, 0, 0], [0, 0, 1, 0.49], [-1, 0, 0, 5], [0, 0, 0, 1]],
                             dtype=np.float32)
    view = np.linalg.inv(world_T_lidar).T.astype(np.float32).copy()
    return {"means3D": means.astype(np.float32), "scales": scales.astype(np.float32),
            "rotations": quat.astype(np.float32), "opacities": opa.astype(np.float32),
            "H": H, "W": W, "viewmatrix": view, "projmatrix": proj,
            "dL_dout": upstream_grad(H, W, 1)}


def knn_cloud(P, seed=0, keyframes=64):
    """Accumulated cloud for distCUDA2: `keyframes` scans from the range model above, in
    the world frame, plus 0.1 % exact duplicates."""
    per = (P + keyframes - 1) // keyframes
    rng = np.random.default_rng(seed)
    chunks = []
    for k in range(keyframes):
        s = surfels(per, 64, 1024, 33.2, seed=seed * 1000 + k + 1)["means3D"].astype(np.float64)
        T = pose(k)
        chunks.append(s @ T[:3, :3].T + T[:3, 3])
    pts = np.concatenate(chunks)[:P]
    ndup = max(1, P // 1000)
    src = rng.integers(0, P, ndup)
    dst = rng.integers(0, P, ndup)
    pts[dst] = pts[src]
    return pts.astype(np.float32)


def edge_scene():
    """Hand-made edge cases on a 40x200 image (neither a multiple of
    16): azimuth-seam straddlers, elevation beyond the image, back-facing, range <= 0.1 and
    > 120, a giant near surfel whose tile rect is wider than the grid (duplicate tiles),
    opacity 1.0 (0.99 clamp), stacked opaque surfels (mid-list termination), tiny opacity."""
    H, W, vfov = 40, 200, 60.0
    rng = np.random.default_rng(7)
    rows = []   # (az, el, r, tilt_deg, sx, sy, opa, flip)

    def add(az, el, r, tilt=0.0, sx=0.05, sy=0.05, opa=0.8, flip=False):
        rows.append((az, el, r, tilt, sx, sy, opa, flip))

    for a in (math.pi - 0.003, -math.pi + 0.004, math.pi - 0.02, -math.pi + 0.015):
        add(a, 0.05, 6.0, 20.0, 0.12, 0.06)                      # straddle the +-pi seam
    add(0.3, math.radians(40.0), 5.0)                            # above the image
    add(0.3, math.radians(-33.0), 5.0, sx=0.3, sy=0.3)           # partly below
    add(1.0, 0.0, 5.0, flip=True)                                # back-facing
    add(1.2, 0.0, 0.08); add(1.3, 0.0, 0.1); add(1.4, 0.0, 125.0); add(1.5, 0.0, 120.0)
    add(-2.0, 0.0, 0.6, 10.0, 1.5, 1.2, 0.6)                     # giant: rect wider than grid
    add(2.0, 0.1, 4.0, 0.0, 0.2, 0.2, 1.0)                       # alpha clamp 0.99
    for k in range(12):                                           # opaque stack -> termination
        add(-1.0, -0.1, 3.0 + 0.25 * k, 5.0 * k, 0.25, 0.25, 0.97)
    add(-1.0, -0.1, 8.0, 0.0, 0.4, 0.4, 0.9)                     # behind the stack
    add(0.7, 0.2, 7.0, 0.0, 0.1, 0.1, 0.003)                     # below 1/255
    add(0.75, 0.2, 7.0, 0.0, 0.1, 0.1, 1.0 / 255.0)
    add(-0.5, 0.3, 9.0, 75.0, 0.3, 0.02, 0.9)                    # grazing, very anisotropic
    for _ in range(40):                                           # filler
        add(rng.uniform(-math.pi, math.pi), rng.uniform(-0.5, 0.5), rng.uniform(2, 30),
            rng.uniform(0, 70), rng.uniform(0.02, 0.4), rng.uniform(0.02, 0.4),
            rng.uniform(0.2, 1.0))
    a = np.array([r[:7] for r in rows], dtype=np.float64)
    flip = np.array([r[7] for r in rows])
    az, el, r, tilt = a[:, 0], a[:, 1], a[:, 2], np.radians(a[:, 3])
    d = np.stack([np.cos(az) * np.cos(el), np.sin(az) * np.cos(el), np.sin(el)], 1)
    axis = np.cross(d, np.array([0.0, 0.0, 1.0]) + 0.1 * d[:, ::-1])
    axis /= np.linalg.norm(axis, axis=1, keepdims=True)
    nrm = -d * np.cos(tilt)[:, None] + np.cross(axis, -d) * np.sin(tilt)[:, None]
    nrm[flip] *= -1
    t = np.cross(nrm, axis); t /= np.linalg.norm(t, axis=1, keepdims=True)
    R = np.stack([t, np.cross(nrm, t), nrm], 2)
    s = {"means3D": (d * r[:, None]).astype(np.float32), "scales": a[:, 4:6].astype(np.float32),
         "rotations": _rotmat_to_quat(R).astype(np.float32),
         "opacities": a[:, 6:7].astype(np.float32), "H": H, "W": W,
         "viewmatrix": viewmatrix(1), "projmatrix": intrinsics(H, W, vfov),
         "dL_dout": upstream_grad(H, W, 3)}
    # the scene is defined in the frame of keyframe 0, viewed from keyframe 1
    return s


def precomp_scene():
    """Inputs for the transMat_precomp path: the T matrices of scene C0 as the reference
    computes them are fed back as `cov3D_precomp` (filled in by the caller)."""
    return make_scene("C0", seed=5)


def input_digest(s):
    """Stable digest of a scene's arrays, stored in golden files to detect generator drift."""
    import hashlib
    h = hashlib.sha256()
    for k in ("means3D", "scales", "rotations", "opacities", "viewmatrix", "projmatrix", "dL_dout"):
        h.update(np.ascontiguousarray(s[k]).tobytes())
    return h.hexdigest()


def room_range_image(H, W, vfov_deg, pose=None, half=(9.0, 6.0, 2.5), holes=0.02, seed=0):
    """Range image of the inside of a box room (plus a pillar) seen by a spherical sensor at
    `pose` (4x4 world_T_sensor), with K as in `intrinsics`; `holes`: fraction of pixels without a
    return (0).  For the photometric-registration tests."""
    K = intrinsics(H, W, vfov_deg)[:3, :3].T.astype(np.float64)
    pose = np.eye(4) if pose is None else np.asarray(pose, dtype=np.float64)
    col, row = np.meshgrid(np.arange(W, dtype=np.float64), np.arange(H, dtype=np.float64))
    az, el = (col - K[0, 2]) / K[0, 0], (row - K[1, 2]) / K[1, 1]
    d = np.stack([np.cos(az) * np.cos(el), np.sin(az) * np.cos(el), np.sin(el)], -1) @ pose[:3, :3].T
    o = pose[:3, 3]
    t = np.full((H, W), np.inf)
    for ax in range(3):
        for sgn in (-1.0, 1.0):
            with np.errstate(divide="ignore", invalid="ignore"):
                tt = (sgn * half[ax] - o[ax]) / d[..., ax]
            tt[~(tt > 0)] = np.inf
            t = np.minimum(t, tt)
    # a vertical pillar of radius 0.7 at (4, 2)
    cxy, rad = np.array([4.0, 2.0]), 0.7
    oc = o[:2] - cxy
    a = d[..., 0] ** 2 + d[..., 1] ** 2
    b = 2 * (d[..., 0] * oc[0] + d[..., 1] * oc[1])
    c = oc @ oc - rad * rad
    disc = b * b - 4 * a * c
    with np.errstate(divide="ignore", invalid="ignore"):
        tp = (-b - np.sqrt(np.where(disc > 0, disc, np.nan))) / (2 * a)
    tp[~(tp > 0)] = np.inf
    t = np.minimum(t, tp)
    rng = np.random.default_rng(seed)
    t = t * (1 + 0.002 * rng.standard_normal(t.shape))
    t[rng.random(t.shape) < holes] = 0.0
    t[~np.isfinite(t)] = 0.0
    return t.astype(np.float32), K.astype(np.float32)


def small_motion():
    """A generic small rigid motion (13 mm, -21 mm, 4 mm; 0.17 / -0.11 / 0.23 degrees): the
    initial guess of the photometric-registration comparisons (see make_golden_photo.py)."""
    rx, ry, rz = 0.003, -0.002, 0.004
    Rx = np.array([[1, 0, 0], [0, math.cos(rx), -math.sin(rx)], [0, math.sin(rx), math.cos(rx)]])
    Ry = np.array([[math.cos(ry), 0, math.sin(ry)], [0, 1, 0], [-math.sin(ry), 0, math.cos(ry)]])
    Rz = np.array([[math.cos(rz), -math.sin(rz), 0], [math.sin(rz), math.cos(rz), 0], [0, 0, 1]])
    T = np.eye(4)
    T[:3, :3] = Rz @ Ry @ Rx
    T[:3, 3] = [0.013, -0.021, 0.004]
    return T.astype(np.float32)
