"""TEST INFRASTRUCTURE - numpy restatement of the range-image front end
(scene/preprocessing.py:24-83 of the reference).  Only tests/ may import this.

PARITY UNPINNED.  The reference delegates the projection and the intrinsics to the
`pyprojections` package (PyPI 0.0.3, pixi.toml:39), which is neither vendored under the
reference tree nor installed here, and it has no test or fixture for this step; so this file
follows the lines of preprocessing.py that ARE in the tree (:56-64, :112-113: the float32 norm,
the gathers, the zeroing of invalid pixels, normals = -p/|p|) and, for the two pyprojections
calls, the conventions documented in csrc/range_image.cu - the ones consistent with the two
artefacts of this step the reference ships (submodules/gsaligner/tests/data/K_ref.npy and
depth_query.npy; tests/golden/range_image_evidence.json holds what was read off them)."""
import numpy as np

f32 = np.float32


def fma(a, b, c):
    """float32 fused multiply-add (the product of two float32 is exact in float64)."""
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(f32)


def angles(cloud):
    x, y, z = cloud[:, 0].astype(f32), cloud[:, 1].astype(f32), cloud[:, 2].astype(f32)
    az = np.arctan2(y, x).astype(f32)
    el = np.arctan2(z, np.sqrt(fma(y, y, x * x))).astype(f32)
    return az, el


def spherical_intrinsics(cloud, H, W):
    """pyp.calculate_spherical_intrinsics (preprocessing.py:41-43), restated: K (3x3 float32)."""
    az, el = angles(cloud)
    ok = ~(np.isnan(az) | np.isnan(el))
    a0, a1, e0, e1 = az[ok].min(), az[ok].max(), el[ok].min(), el[ok].max()
    fx = f32(-(W - 1)) / f32(a1 - a0)
    fy = f32(-(H - 1)) / f32(e1 - e0)
    cx = f32(0.5) * f32(W) - fx * (f32(0.5) * f32(a0 + a1))
    cy = f32(0.5) * f32(H) - fy * (f32(0.5) * f32(e0 + e1))
    return np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], dtype=f32)


def project(cloud, K, H, W, depth_min, depth_max):
    """pyp.Camera(...).project + preprocessing.py:56-64 -> lut, range image, normal image,
    valid image."""
    cloud = cloud.astype(f32).reshape(-1, 3)
    if cloud.shape[0] == 0:
        return (np.full((H, W), -1, np.int64), np.zeros((1, H, W), f32), np.zeros((3, H, W), f32),
                np.zeros((1, H, W), bool))
    x, y, z = cloud[:, 0], cloud[:, 1], cloud[:, 2]
    r = np.linalg.norm(cloud, axis=1)                 # preprocessing.py:56, float32
    az, el = angles(cloud)
    u = fma(np.full_like(az, K[0, 0]), az, np.full_like(az, K[0, 2]))
    v = fma(np.full_like(el, K[1, 1]), el, np.full_like(el, K[1, 2]))
    fin = (np.abs(u) < 1e9) & (np.abs(v) < 1e9)
    u, v = np.where(fin, u, f32(-9)), np.where(fin, v, f32(-9))
    col, row = np.floor(u).astype(np.int64) + 1, np.floor(v).astype(np.int64) + 1
    ok = (r > f32(depth_min)) & (r <= f32(depth_max)) & (col >= 0) & (col < W) & (row >= 0) & (row < H)
    idx = np.flatnonzero(ok)
    pix = row[idx] * W + col[idx]
    key = (r[idx].view(np.uint32).astype(np.uint64) << np.uint64(32)) | idx.astype(np.uint64)
    order = np.argsort(key, kind="stable")
    first = np.unique(pix[order], return_index=True)
    lut = np.full(H * W, -1, dtype=np.int64)
    lut[first[0]] = idx[order][first[1]]
    valid = lut >= 0
    rng = np.where(valid, r[np.maximum(lut, 0)], f32(0)).astype(f32)
    p = cloud[np.maximum(lut, 0)]
    nrm = np.where(valid[:, None], -p / np.where(valid, rng, f32(1))[:, None], f32(0)).astype(f32)
    return (lut.reshape(H, W), rng.reshape(1, H, W), nrm.T.reshape(3, H, W).copy(),
            valid.reshape(1, H, W))
