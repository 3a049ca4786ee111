"""Developer tool: per-CTA phase timestamps of tile_sort_kernel (build with -DSORT_TRACE)."""
import ctypes, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "splat-loam_b200"), os.path.join(ROOT, "tests")]
import scenes
from diff_surfel_spherical_rasterization import _C as C
import splat_b200_native as n
s = scenes.make_scene(sys.argv[1] if len(sys.argv) > 1 else "C2")
s = {k: (torch.from_numpy(v).cuda() if isinstance(v, np.ndarray) else v) for k, v in s.items()}
empty = torch.empty(0, device="cuda")
for _ in range(3):
    C.rasterize_gaussians(s["means3D"], s["opacities"], s["scales"], s["rotations"], 1.0, empty,
                          s["viewmatrix"], s["projmatrix"], s["H"], s["W"], False, False)
torch.cuda.synchronize()
buf = np.zeros((1024, 16), dtype=np.uint64)
assert n.load().splat_debug_sort_trace(buf.ctypes.data_as(ctypes.c_void_p)) == 0
t = buf.astype(np.int64)
t = t[t[:, 12] > 0]
t0 = t[:, 0].min()
names = ["load", "p0 rank", "p0 scan", "p0 scatter", "p1 rank", "p1 scan", "p1 scatter", "p2 rank",
         "p2 scan", "p2 scatter", "write-out"]
print(f"{len(t)} CTAs; span {(t[:, 11].max() - t0) / 1000:.1f} us; list lengths min/median/max "
      f"{t[:, 12].min()} / {int(np.median(t[:, 12]))} / {t[:, 12].max()}")
print("second wave starts at %.1f us" % ((np.sort(t[:, 0])[min(148, len(t) - 1)] - t0) / 1000))
d = np.diff(t[:, :12], axis=1) / 1000.0
print("phase medians (us): " + ", ".join(f"{names[j]} {np.median(d[:, j]):.2f}" for j in range(11)))
print("per CTA total median %.1f us, max %.1f us" % (np.median(t[:, 11] - t[:, 0]) / 1000, (t[:, 11] - t[:, 0]).max() / 1000))
w = t[:, [10, 13, 14, 15, 11]]
dw = np.diff(w, axis=1) / 1000.0
print("write-out parts (median us): point_list %.2f, count+warp scan %.2f, block scan %.2f, list stores %.2f" % tuple(np.median(dw, axis=0)))
