"""TEST INFRASTRUCTURE - numpy restatement (float32 where the reference computes in float32)
of the photometric registration of the reference's gsaligner submodule, in the configuration
Splat-LOAM uses (spherical camera, depth channel only, identity sensor offset).  Only tests/
and bench.py's cpu_baseline leg may import this; the product never does.

Follows, under /root/reference/submodules/gsaligner/gsaligner/:
  aligner_utils.cu:3-16                      depth > 1e-5 unmasks a pixel
  tools/image_set_impl.cu:318-376            mask growing (flattened offsets), :269-300 derivatives
  tools/image_set_impl.cu:20-77              toCloud_kernel (spherical back-projection)
  tools/utils_photo.cuh:54-82                project (spherical)
  registration/photometric_cuda.cu:8-72      two-pass z-buffer projection
  tools/image_set.cuh:75-107                 getSubPixel (bilinear value + derivatives)
  registration/photometric_cuda.cu:99-311    errorAndJacobian / linearize_kernel
  registration/photometric_cuda.cu:313-395   sums, LDLT solve, pose update
  tools/lie_algebra.h:41-53                  expMapSO3
Pinned by tests/golden/photo_*.npz (outputs of the reference build, tests/golden/
make_golden_photo.py)."""
import numpy as np

f32 = np.float32


def prepare_reference(depth, threshold=0.5, radius=1):
    """-> (depth, d/dcol, d/drow, masked) of the reference image."""
    H, W = depth.shape
    d = depth.astype(f32)
    m0 = ~(d > 1e-5)
    flat = m0.ravel()
    grown = flat.copy()
    n = H * W
    idx = np.flatnonzero(flat)
    for r in range(-radius, radius + 1):
        for c in range(-radius, radius + 1):
            if r * r + c * c <= radius * radius:
                t = idx + (r * W + c)                  # flattened offsets, no row check
                t = t[(t >= 0) & (t < n)]
                grown[t] = True
    mask = grown.reshape(H, W)
    dc = np.zeros((H, W), f32)
    dr = np.zeros((H, W), f32)
    dr[1:-1, 1:-1] = (f32(0.5) * d[2:, 1:-1] - f32(0.5) * d[:-2, 1:-1])[:, :]
    dc[1:-1, 1:-1] = (f32(0.5) * d[1:-1, 2:] - f32(0.5) * d[1:-1, :-2])[:, :]
    dr[:, [0, -1]] = 0
    dc[[0, -1], :] = 0
    nrm = dc * dc + dr * dr
    clamp = (~mask) & ~(nrm < f32(threshold * threshold))
    clamp[[0, -1], :] = False
    clamp[:, [0, -1]] = False
    s = np.sqrt(f32(threshold * threshold) / np.where(clamp, nrm, f32(1))).astype(f32)
    dc = np.where(clamp, dc * s, dc).astype(f32)
    dr = np.where(clamp, dr * s, dr).astype(f32)
    return d, dc, dr, mask


def query_cloud(depth, K, min_depth, max_depth):
    H, W = depth.shape
    w = depth.astype(f32)
    ifx, ify, cx, cy = f32(1) / f32(K[0, 0]), f32(1) / f32(K[1, 1]), f32(K[0, 2]), f32(K[1, 2])
    col, row = np.meshgrid(np.arange(W, dtype=f32), np.arange(H, dtype=f32))
    az, el = ifx * (col - cx), ify * (row - cy)
    good = (w > 1e-5) & ~(w < f32(min_depth)) & ~(w > f32(max_depth))
    p = np.stack([np.cos(az) * np.cos(el) * w, np.sin(az) * np.cos(el) * w, np.sin(el) * w], -1)
    p = np.where(good[..., None], p, 0).astype(f32)
    return p.reshape(-1, 3)


def linearize(ref, cloud, T, K, min_depth, max_depth, omega_depth, d_threshold, e_threshold):
    """One projection + linearisation at pose T (4x4).  -> 30 sums: upper H | b | chi | good |
    inliers."""
    d, dc, dr, mask = ref
    H, W = d.shape
    R, t = T[:3, :3].astype(f32), T[:3, 3].astype(f32)
    fx, fy, cx, cy = f32(K[0, 0]), f32(K[1, 1]), f32(K[0, 2]), f32(K[1, 2])
    tp = (cloud @ R.T + t).astype(f32)
    rxy2 = tp[:, 0] * tp[:, 0] + tp[:, 1] * tp[:, 1]
    depth = np.sqrt(rxy2 + tp[:, 2] * tp[:, 2]).astype(f32)
    good = ~(depth < f32(min_depth)) & ~(depth > f32(max_depth))
    u = (fx * np.arctan2(tp[:, 1], tp[:, 0]).astype(f32) + cx).astype(f32)
    v = (fy * np.arctan2(tp[:, 2], np.sqrt(rxy2)).astype(f32) + cy).astype(f32)
    irow = (v + np.where(v >= 0, f32(0.5), f32(-0.5))).astype(np.int64)   # truncation
    icol = (u + np.where(u >= 0, f32(0.5), f32(-0.5))).astype(np.int64)
    good &= (irow >= 0) & (irow < H) & (icol >= 0) & (icol < W)
    idx = np.flatnonzero(good)
    pix = irow[idx] * W + icol[idx]
    keep = ~mask.ravel()[pix]
    idx, pix = idx[keep], pix[keep]
    # z-buffer: smallest (depth bits, index) per pixel
    key = (depth[idx].view(np.uint32).astype(np.uint64) << np.uint64(32)) | idx.astype(np.uint64)
    order = np.argsort(key, kind="stable")
    pix_s, idx_s = pix[order], idx[order]
    first = np.unique(pix_s, return_index=True)[1]
    win = idx_s[first]
    # getSubPixel
    uu, vv = u[win], v[win]
    r0, c0 = vv.astype(np.int64), uu.astype(np.int64)     # (int) truncates
    ok = (r0 >= 0) & (c0 >= 0) & (r0 + 1 < H) & (c0 + 1 < W)
    win, uu, vv, r0, c0 = win[ok], uu[ok], vv[ok], r0[ok], c0[ok]
    mk = mask[r0, c0] | mask[r0, c0 + 1] | mask[r0 + 1, c0] | mask[r0 + 1, c0 + 1]
    win, uu, vv, r0, c0 = win[~mk], uu[~mk], vv[~mk], r0[~mk], c0[~mk]
    drr, dcc = (vv - r0.astype(f32)).astype(f32), (uu - c0.astype(f32)).astype(f32)
    dr1, dc1 = f32(1) - drr, f32(1) - dcc

    def interp(a):
        return ((a[r0, c0] * dc1 + a[r0, c0 + 1] * dcc) * dr1 +
                (a[r0 + 1, c0] * dc1 + a[r0 + 1, c0 + 1] * dcc) * drr).astype(f32)

    meas, gc, gr = interp(d), interp(dc), interp(dr)
    wd = np.sqrt(f32(omega_depth)).astype(f32)
    e = ((depth[win] - meas) * wd).astype(f32)
    ok = ~(e * e > f32(d_threshold))
    win, e, gc, gr = win[ok], e[ok], gc[ok], gr[ok]
    p, q, z = cloud[win], tp[win], depth[win]
    ir = (f32(1) / z).astype(f32)
    ir2 = ir * ir
    rxy2 = (q[:, 0] * q[:, 0] + q[:, 1] * q[:, 1]).astype(f32)
    irxy2 = (1.0 / rxy2.astype(np.float64)).astype(f32)
    rxy = np.sqrt(rxy2).astype(f32)
    irxy = (1.0 / rxy.astype(np.float64)).astype(f32)
    p00, p01 = -fx * q[:, 1] * irxy2, fx * q[:, 0] * irxy2
    p10 = -fy * q[:, 0] * q[:, 2] * irxy * ir2
    p11 = -fy * q[:, 1] * q[:, 2] * irxy * ir2
    p12 = fy * rxy * ir2
    a = np.stack([q[:, 0] * ir - (gc * p00 + gr * p10), q[:, 1] * ir - (gc * p01 + gr * p11),
                  q[:, 2] * ir - gr * p12], 1).astype(f32)
    zero = np.zeros(len(win), f32)
    S = np.stack([np.stack([zero, -p[:, 2], p[:, 1]], 1), np.stack([p[:, 2], zero, -p[:, 0]], 1),
                  np.stack([-p[:, 1], p[:, 0], zero], 1)], 1)                 # [n,3,3]
    RS = (R[None] @ S).astype(f32)
    J = np.concatenate([a @ R, f32(-2) * np.einsum("nk,nkc->nc", a, RS)], 1).astype(f32) * wd
    chi = (e * e).astype(f32)
    out_l = chi > f32(e_threshold)
    lam = np.where(out_l, np.sqrt(f32(e_threshold) / np.where(out_l, chi, f32(1))), f32(1)).astype(f32)
    scaling = 1.0 / (H * W)
    sums = np.zeros(30)
    k = 0
    for i in range(6):
        for j in range(i, 6):
            sums[k] = ((J[:, i] * J[:, j] * lam).astype(f32).astype(np.float64) * scaling).sum()
            k += 1
    for i in range(6):
        sums[21 + i] = ((J[:, i] * e * lam).astype(f32).astype(np.float64) * scaling).sum()
    sums[27], sums[28], sums[29] = chi.astype(np.float64).sum(), len(win), int((~out_l).sum())
    return sums


def exp_so3(w):
    th2 = float(w @ w)
    th = np.sqrt(th2)
    Wm = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]], dtype=np.float64)
    if th2 < 1e-8:
        return np.eye(3) + Wm
    Km = Wm / th
    return np.eye(3) + np.sin(th) * Km + 2.0 * np.sin(th / 2.0) ** 2 * Km @ Km


def update_pose(T, sums):
    """photometric_cuda.cu:369-392: solve, X = X * dX (in float32)."""
    if sums[28] == 0 or sums[29] == 0:
        return T
    Hm = np.zeros((6, 6))
    Hm[np.triu_indices(6)] = sums[:21]
    Hm = Hm + np.triu(Hm, 1).T
    dx = np.linalg.solve(Hm, -sums[21:27])
    dX = np.eye(4, dtype=f32)
    dX[:3, :3] = exp_so3(dx[3:]).astype(f32)
    dX[:3, 3] = dx[:3].astype(f32)
    return (T.astype(f32) @ dX).astype(f32)


def align(ref_depth, query_depth, K, T0, iterations, min_depth=0.5, max_depth=80.0, omega_depth=5.0,
          d_threshold=0.8, e_threshold=1.0):
    ref = prepare_reference(ref_depth)
    cloud = query_cloud(query_depth, K, min_depth, max_depth)
    T = T0.astype(f32)
    sums = np.zeros(30)
    for _ in range(iterations):
        sums = linearize(ref, cloud, T, K, min_depth, max_depth, omega_depth, d_threshold, e_threshold)
        T = update_pose(T, sums)
    HW = ref_depth.size
    if sums[28] == 0 or sums[29] == 0:
        stats = (f32(sums[27]), f32(sums[29]) / f32(HW))
    else:
        stats = (f32(sums[27]) / f32(sums[28]), f32(sums[29]) / f32(HW))
    return T, stats, sums
