"""Developer tool (CPU): how many sub-tile visits a tighter per-surfel pixel box would save.

For every 4x4 sub-tile of a frame the C oracle counts the traversed list entries, those whose
box overlaps the sub-tile ("visited" by the compositing kernels) and those with at least one
passing pixel ("true").  Compared: the product's box (bounding sphere of the cut-off ellipse ->
spherical cap, dssr_preprocess.cu:cull_box) and a box from the ellipse's own extent (exact
azimuth tangents of its xy-projection; elevation from its z- and radial ranges).

    python tests/dev_mask_study.py [C1|C2]
"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import scenes  # noqa: E402
from oracle import oracle  # noqa: E402


def cut_of(opa):
    return (2.0 * np.log(255.0 * np.maximum(opa, 1e-12))) * 1.002 + 0.01


def sphere_boxes(T, opa, xy, fx, fy, cy):
    """numpy restatement of cull_box (float64)."""
    tu, tv, tw = T[:, 0:3], T[:, 4:7], T[:, 8:11]
    cut = cut_of(opa)
    a, c, b = (tu * tu).sum(1), (tv * tv).sum(1), (tu * tv).sum(1)
    lam = 0.5 * (a + c) + np.sqrt((0.5 * (a - c)) ** 2 + b * b)
    rs = np.sqrt(cut * lam) * 1.001
    dist = np.linalg.norm(tw, axis=1)
    r2 = np.sqrt(0.5 * cut) + 0.01
    st = rs / dist
    everything = ~(st < 0.999)
    st = np.minimum(st, 0.999)
    theta = np.arcsin(st) * 1.0005 + 1e-5
    rxy = np.hypot(tw[:, 0], tw[:, 1])
    phi = np.arctan2(tw[:, 2], rxy)
    e0, e1 = (phi - theta) * fy + cy, (phi + theta) * fy + cy
    ymin = np.minimum(xy[:, 1] - r2, np.minimum(e0, e1))
    ymax = np.maximum(xy[:, 1] + r2, np.maximum(e0, e1))
    pole = np.abs(phi) + theta >= 1.5607
    sd = st / np.cos(phi)
    dl = np.where(sd < 0.9999, np.arcsin(np.minimum(sd, 0.9999)) * 1.0005 + 1e-5, 3.2)
    hw = dl * abs(fx)
    xmin = np.where(pole, -1e30, np.minimum(xy[:, 0] - r2, xy[:, 0] - hw))
    xmax = np.where(pole, 1e30, np.maximum(xy[:, 0] + r2, xy[:, 0] + hw))
    M = 0.25
    box = np.stack([xmin - M, xmax + M, ymin - M, ymax + M], 1)
    box[everything] = [-1e30, 1e30, -1e30, 1e30]
    return box


def ellipse_boxes(T, opa, xy, fx, fy, cy):
    """Box from the ellipse {Tw + u Tu + v Tv : u^2 + v^2 <= cut} itself."""
    tu, tv, tw = T[:, 0:3], T[:, 4:7], T[:, 8:11]
    cut = cut_of(opa)
    r2 = np.sqrt(0.5 * cut) + 0.01
    # azimuth: tangents from the origin to the xy-projection.  With n = (-sin a, cos a):
    # (n.c)^2 = cut ((n.tu)^2 + (n.tv)^2)  ->  n^T Q n = 0, Q = c c^T - cut (tu tu^T + tv tv^T)
    cx_, cy_ = tw[:, 0], tw[:, 1]
    q11 = cx_ * cx_ - cut * (tu[:, 0] ** 2 + tv[:, 0] ** 2)
    q22 = cy_ * cy_ - cut * (tu[:, 1] ** 2 + tv[:, 1] ** 2)
    q12 = cx_ * cy_ - cut * (tu[:, 0] * tu[:, 1] + tv[:, 0] * tv[:, 1])
    # half-width of the azimuth interval around the centre's azimuth: rotate so the centre lies
    # on +x: then tangent directions are +-dl with tan(dl)^2 = -Q'22 / Q'11 ... use the invariant
    # form: sin^2(dl) = lambda_perp / |c|^2 where lambda_perp = cut * ((t.tu)^2 + (t.tv)^2) / ...
    # simpler and still exact: support of the projected ellipse along the direction t
    # perpendicular to c, and along c:
    rc = np.hypot(cx_, cy_)
    ux, uy = cx_ / rc, cy_ / rc                       # radial unit vector
    px, py = -uy, ux                                  # tangential unit vector
    s_perp = np.sqrt(cut * ((tu[:, 0] * px + tu[:, 1] * py) ** 2 + (tv[:, 0] * px + tv[:, 1] * py) ** 2))
    s_rad = np.sqrt(cut * ((tu[:, 0] * ux + tu[:, 1] * uy) ** 2 + (tv[:, 0] * ux + tv[:, 1] * uy) ** 2))
    rmin = rc - s_rad
    polar = rmin <= 1e-3 * rc               # the projection may contain the vertical axis
    rmin = np.maximum(rmin, 1e-9)
    # every point has |tangential offset| <= s_perp and radial coordinate >= rmin
    dl = np.arctan2(s_perp, rmin) * 1.0005 + 1e-5
    hw = dl * abs(fx)
    xmin = np.where(polar, -1e30, np.minimum(xy[:, 0] - r2, xy[:, 0] - hw))
    xmax = np.where(polar, 1e30, np.maximum(xy[:, 0] + r2, xy[:, 0] + hw))
    # elevation: z in [zc - sz, zc + sz]; horizontal distance in [rmin', rc + s_rad] where the
    # tangential spread can only increase the distance
    sz = np.sqrt(cut * (tu[:, 2] ** 2 + tv[:, 2] ** 2))
    zlo, zhi = tw[:, 2] - sz, tw[:, 2] + sz
    rmax = np.hypot(rc + s_rad, s_perp)
    el_hi = np.where(zhi > 0, np.arctan2(zhi, rmin), np.arctan2(zhi, rmax)) + 1e-5
    el_lo = np.where(zlo < 0, np.arctan2(zlo, rmin), np.arctan2(zlo, rmax)) - 1e-5
    e0, e1 = el_lo * fy + cy, el_hi * fy + cy
    ymin = np.where(polar, -1e30, np.minimum(xy[:, 1] - r2, np.minimum(e0, e1)))
    ymax = np.where(polar, 1e30, np.maximum(xy[:, 1] + r2, np.maximum(e0, e1)))
    M = 0.25
    _ = (q11, q22, q12)
    return np.stack([xmin - M, xmax + M, ymin - M, ymax + M], 1)


def main(cfg):
    if cfg == "edge":
        s = scenes.edge_scene()
    elif cfg == "ref":
        s = scenes.reference_test_scene()
    elif ":" in cfg:                       # e.g. C1:3:7 = config, keyframe, seed
        name, kf, seed = cfg.split(":")
        s = scenes.make_scene(name, seed=int(seed), keyframe=int(kf))
    else:
        s = scenes.make_scene(cfg)
    f = oracle.rasterize_forward(s["means3D"], s["scales"], s["rotations"], s["opacities"],
                                 s["viewmatrix"], s["projmatrix"], s["H"], s["W"])
    L = oracle._load()
    H, W = s["H"], s["W"]
    Pm = np.ascontiguousarray(s["projmatrix"], np.float32)
    fx, fy, cy = float(Pm.flat[0]), float(Pm.flat[5]), float(Pm.flat[9])
    T = np.ascontiguousarray(f["transMat"], np.float32).reshape(-1, 12)
    no = np.ascontiguousarray(f["normal_opacity"], np.float32).reshape(-1, 4)
    xy = np.ascontiguousarray(f["means2D"], np.float32).reshape(-1, 2)
    nsub = ((W + 3) // 4) * ((H + 3) // 4)
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)  # noqa: E731
    res = {}
    for name, fn in (("sphere", sphere_boxes), ("ellipse", ellipse_boxes)):
        with np.errstate(all="ignore"):
            boxes = np.ascontiguousarray(fn(T.astype(np.float64), no[:, 3].astype(np.float64),
                                            xy.astype(np.float64), fx, fy, cy), np.float32)
        boxes[~np.isfinite(boxes).all(1)] = [-1e30, 1e30, -1e30, 1e30]
        out = np.zeros(nsub * 7, np.uint32)
        ranges = np.ascontiguousarray(f["ranges"], np.uint32)
        pl = np.ascontiguousarray(f["point_list"], np.uint32)
        nc = np.ascontiguousarray(f["n_contrib"], np.uint32)
        L.orc_subtile_stats(W, H, p(ranges), p(pl), p(xy), p(T), p(no), p(Pm), p(nc), p(boxes), p(out))
        o = out.reshape(-1, 7).astype(np.int64)
        res[name] = o
        print(f"{name:8s}: two-at-a-time steps {o[:, 3].sum()} for {o[:, 1].sum()} visits "
              f"(ideal {(o[:, 1].sum() + 1) // 2}); 32-entry steps with a visit {o[:, 4].sum()} of {((o[:, 0] + 31) // 32).sum()}")
        print(f"{name:8s}: (pixel, visit) pairs {16 * o[:, 1].sum()}: pixel inside the box {o[:, 5].sum()} "
              f"({100 * o[:, 5].sum() / (16 * o[:, 1].sum()):.0f} %), pass the alpha test {o[:, 6].sum()} "
              f"({100 * o[:, 6].sum() / (16 * o[:, 1].sum()):.0f} %)")
        print(f"{name:8s}: traversed {o[:, 0].sum():>10d}  visited {o[:, 1].sum():>10d}  true {o[:, 2].sum():>10d}"
              f"  (per sub-tile {o[:, 0].mean():.0f} / {o[:, 1].mean():.0f} / {o[:, 2].mean():.0f})")
    a, b = res["sphere"], res["ellipse"]
    assert (b[:, 2] == a[:, 2]).all(), "the tighter box LOST true hits: not conservative"
    print(f"visits saved: {100 * (1 - b[:, 1].sum() / a[:, 1].sum()):.1f} %; false visits "
          f"{a[:, 1].sum() - a[:, 2].sum()} -> {b[:, 1].sum() - b[:, 2].sum()}")


if __name__ == "__main__":
    for c in (sys.argv[1:] or ["C2"]):
        print("==", c)
        main(c)
