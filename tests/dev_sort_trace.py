"""Developer tool: per-CTA phase timestamps of the last radix pass.  Build the library with
-DSORT_TRACE first:  python -c "import sys; sys.path.insert(0,'splat-loam_b200'); import build; build.NVCC_FLAGS.append('-DSORT_TRACE'); build.build(force=True)" """
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
t = t[t[:, 0] > 0]
rel = (t - t[:, 0].min()) / 1000.0
names = ["start", "ranked", "published", "scanned", "staged", "lookback", "synced", "end"]
print(f"{len(t)} CTAs; kernel span {rel[:, 7].max():.1f} us; second wave starts at {np.sort(rel[:, 0])[min(148, len(t) - 1)]:.1f}")
d = np.diff(rel[:, :8], axis=1)
print("all warps loaded after (median us):", np.median(rel[:, 8] - rel[:, 0]))
print("phase durations (median us): " + ", ".join(f"{names[j + 1]} {np.median(d[:, j]):.2f}" for j in range(7)))
