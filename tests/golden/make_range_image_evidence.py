"""Reads what the reference tree holds about its range-image step (which it delegates to the
absent `pyprojections` package) and writes tests/golden/range_image_evidence.json.

Run in the build container only (needs /root/reference):
    python tests/golden/make_range_image_evidence.py

Sources, all under submodules/gsaligner/tests/data/ of the reference:
  K_ref.npy        camera matrix produced by pyp.calculate_spherical_intrinsics
  depth_query.npy  range image produced by pyp.Camera.project + preprocessing.py:56-58
  cloud_query.ply  depth_to_points(depth_query) (utils/graphic_utils.py:27-66, slam/tracker.py:
                   160-168): pixel centres, NOT the scan - so it carries no sub-pixel evidence
"""
import json
import os

import numpy as np

D = "/root/reference/submodules/gsaligner/tests/data/"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "range_image_evidence.json")


def read_ply_xyz(path):
    b = open(path, "rb").read()
    i = b.index(b"end_header\n") + len(b"end_header\n")
    n = int([l for l in b[:i].decode().split("\n") if l.startswith("element vertex")][0].split()[-1])
    return np.frombuffer(b[i:i + n * 48], dtype="<f8").reshape(n, 6)[:, :3]


def main():
    K = np.load(D + "K_ref.npy")
    dq = np.load(D + "depth_query.npy")
    H, W = dq.shape
    valid = dq > 0
    pts = read_ply_xyz(D + "cloud_query.ply")
    # the back-projected cloud lies on pixel (col, row) <-> image point (col - 0.5, row - 0.5);
    # a least-squares line through (azimuth, col - 0.5) recovers the K the query frame used
    az = np.arctan2(pts[:, 1], pts[:, 0])
    el = np.arctan2(pts[:, 2], np.hypot(pts[:, 0], pts[:, 1]))
    r = np.linalg.norm(pts, axis=1)
    u, v = K[0, 0] * az + K[0, 2], K[1, 1] * el + K[1, 2]
    col, row = np.floor(u).astype(int) + 1, np.floor(v).astype(int) + 1
    match = np.abs(dq[row.clip(0, H - 1), col.clip(0, W - 1)] - r) < 1e-5
    fx, cx = np.polyfit(az[match], col[match] - 0.5, 1)
    fy, cy = np.polyfit(el[match], row[match] - 0.5, 1)
    ev = {
        "source": "submodules/gsaligner/tests/data/{K_ref.npy,depth_query.npy,cloud_query.ply}",
        "H": H, "W": W,
        "K_ref": K.tolist(),
        "hfov_from_fx": float(-(W - 1) / K[0, 0]),
        "vfov_from_fy_deg": float(np.degrees(-(H - 1) / K[1, 1])),
        "K_query_fit": [float(fx), float(fy), float(cx), float(cy)],
        "query_valid": int(valid.sum()),
        "query_valid_per_row": valid.sum(1).tolist(),
        "query_valid_first_cols": valid.sum(0)[:4].tolist(),
        "query_valid_last_cols": valid.sum(0)[-4:].tolist(),
        "query_depth_min": float(dq[valid].min()), "query_depth_max": float(dq.max()),
        "cloud_query_matches_pixel_centres": float(match.mean()),
    }
    # round trip: the reference's back-projection of its own range image, projected again with
    # the camera matrix the reference's test pairs with it, must give the range image back
    sel = np.random.default_rng(0).choice(len(pts), 8192, replace=False)
    np.savez_compressed(os.path.join(os.path.dirname(OUT), "range_image_roundtrip.npz"),
                        cloud=pts[sel].astype(np.float32), K=K.astype(np.float32),
                        row=row[sel].astype(np.int16), col=col[sel].astype(np.int16),
                        depth=dq[row[sel], col[sel]].astype(np.float32))
    assert match.all()
    with open(OUT, "w") as f:
        json.dump(ev, f, indent=1)
    print(json.dumps(ev, indent=1)[:1500])


if __name__ == "__main__":
    main()
