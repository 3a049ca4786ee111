"""Generate tests/golden/photo_*.npz: inputs and outputs of the UNMODIFIED photometric
registration of the reference's gsaligner submodule (oracle/_ref/libref_photo.so, built by
oracle/Makefile from /root/reference).  Needs a GPU:

    python tests/golden/make_golden_photo.py [outdir]

Cases: the reference's own fixture (submodules/gsaligner/tests/data/depth_{ref,query}.npy,
K_ref.npy - read from /root/reference when present, else from an earlier photo_ref.npz) and a
small synthetic room.  Stored per case: both range images, K, the initial guess, the prepared
reference image (depth, derivatives, mask), the 30 sums of the first linearisation, and the
transform + stats after 1 and 10 iterations.

The initial guess is a small GENERIC motion, not the identity: at the identity every query
point re-projects onto (numerically almost exactly) an integer pixel, so the truncation in
getSubPixel and the +-0.5 rounding of the z-buffer pixel are decided by the last bit of the
sin / cos / atan2 chain - the sums then move by several per cent under a 1e-5 pixel nudge and
are no basis for a comparison.  The transform after 10 iterations FROM the identity (what the
reference's own test does) is stored as well; it is compared with a looser tolerance."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "tests")]
import reflib  # noqa: E402
import scenes  # noqa: E402


def run(name, ref, query, K, T0, out):
    H, W = ref.shape
    r = reflib.RefPhotometric(H, W)
    r.set_reference(torch.from_numpy(ref).cuda(), K)
    r.set_query(torch.from_numpy(query).cuda())
    res = {"ref": ref, "query": query, "K": K, "T0": T0,
           "ref_image": r.reference_image().numpy(),
           "sums0": np.array(r.linearize(T0))}
    for it in (1, 10):
        T, fit, err = r.align(T0, it)
        res[f"T{it}"] = T.numpy()
        res[f"stats{it}"] = np.array([err, fit], dtype=np.float32)     # residual, inlier ratio
    T, fit, err = r.align(np.eye(4, dtype=np.float32), 10)
    res["T10_identity"] = T.numpy()
    np.savez_compressed(os.path.join(out, f"photo_{name}.npz"), **res)
    print(name, res["sums0"][27:], res["stats10"], "\n", res["T10"])


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else HERE
    data = "/root/reference/submodules/gsaligner/tests/data"
    if os.path.isdir(data):
        ref, query = np.load(f"{data}/depth_ref.npy"), np.load(f"{data}/depth_query.npy")
        K = np.load(f"{data}/K_ref.npy")
    else:
        g = np.load(os.path.join(HERE, "photo_ref.npz"))
        ref, query, K = g["ref"], g["query"], g["K"]
    run("ref", ref.astype(np.float32), query.astype(np.float32), K.astype(np.float32),
        scenes.small_motion(), out)
    pose = scenes.pose(1)
    pose[:3, 3] *= 0.3
    a, K = scenes.room_range_image(32, 256, 40.0)
    b, _ = scenes.room_range_image(32, 256, 40.0, pose=pose, seed=1)
    run("room", a, b, K, scenes.small_motion(), out)


if __name__ == "__main__":
    main()
