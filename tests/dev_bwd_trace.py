"""Developer tool: per-warp start / end times of the backward compositing kernel
(build with -DRENDER_TRACE)."""
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
    R, out, radii, geom, binning, img = C.rasterize_gaussians(
        s["means3D"], s["opacities"], s["scales"], s["rotations"], 1.0, empty,
        s["viewmatrix"], s["projmatrix"], s["H"], s["W"], False, False)
    C.rasterize_gaussians_backward(s["means3D"], radii, s["scales"], s["rotations"], 1.0, empty,
                                   s["viewmatrix"], s["projmatrix"], s["dL_dout"], geom, R, binning, img, False)
torch.cuda.synchronize()
buf = np.zeros((8192, 4), dtype=np.uint64)
assert n.load().splat_debug_bwd_trace(buf.ctypes.data_as(ctypes.c_void_p)) == 0
t = buf.astype(np.int64)
t = t[(t[:, 0] > 0) & (t[:, 1] > 0)]
t0 = t[:, 0].min()
start, end = (t[:, 0] - t0) / 1e3, (t[:, 1] - t0) / 1e3
dur, cost = end - start, t[:, 2]
span = end.max()
print(f"{len(t)} warps; kernel span {span:.1f} us; sum of warp time {dur.sum() / 1e3:.2f} ms = "
      f"{dur.sum() / span / 148:.1f} warps busy per SM on average")
print("warp duration us: median %.1f  p90 %.1f  p99 %.1f  max %.1f" % tuple(np.percentile(dur, [50, 90, 99, 100])))
print("start time us:    median %.1f  p75 %.1f p90 %.1f  max %.1f" % tuple(np.percentile(start, [50, 75, 90, 100])))
print("corr(duration, traversed entries) = %.3f" % np.corrcoef(dur, cost)[0, 1])
print("us per traversed entry: %.4f" % (dur.sum() / max(cost.sum(), 1)))
for q in range(100, int(span) + 100, 100):
    print(f"  running at {q} us: {((start <= q) & (end > q)).sum()}", end=";")
print()
# what an exact LPT order on 148*16 slots would give (list scheduling on measured durations)
import heapq
slots = [0.0] * (148 * 16)
heapq.heapify(slots)
for d in sorted(dur, reverse=True):
    heapq.heappush(slots, heapq.heappop(slots) + d)
print("LPT list-scheduling makespan with measured durations: %.1f us" % max(slots))

sb = np.zeros((8192, 5), dtype=np.uint64)
if hasattr(n.load(), "splat_debug_bwd_stages") and n.load().splat_debug_bwd_stages(sb.ctypes.data_as(ctypes.c_void_p)) == 0:
    sb = sb.astype(np.float64)
    sb = sb[sb.sum(axis=1) > 0]
    sh = sb / sb.sum(axis=1, keepdims=True)
    print("stage share of a warp's cycles (mean over warps): gather %.2f P1 %.2f P2 %.2f P3 %.2f reduce %.2f" % tuple(sh.mean(axis=0)))
