"""Developer tool: per-warp (= per 4x4 sub-tile) start / end times of the forward compositing
kernel.  Build with -DRENDER_TRACE first (see tests/dev_sort_trace.py for the recipe)."""
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
buf = np.zeros((8192, 4), dtype=np.uint64)
assert n.load().splat_debug_render_trace(buf.ctypes.data_as(ctypes.c_void_p)) == 0
t = buf.astype(np.int64)
t = t[t[:, 0] > 0]
t0 = t[:, 0].min()
start, end = (t[:, 0] - t0) / 1e3, (t[:, 1] - t0) / 1e3
dur = end - start
steps, pairs, length = t[:, 2] >> 32, t[:, 2] & 0xffffffff, t[:, 3]
print(f"{len(t)} warps; kernel span {end.max():.1f} us; sum of warp time {dur.sum() / 1e3:.2f} ms "
      f"= {dur.sum() / end.max() / 148:.1f} warps busy per SM on average")
print("warp duration us: median %.1f  p90 %.1f  p99 %.1f  max %.1f" % tuple(np.percentile(dur, [50, 90, 99, 100])))
print("start time us:    median %.1f  p90 %.1f  p99 %.1f  max %.1f" % tuple(np.percentile(start, [50, 90, 99, 100])))
i = np.argsort(-dur)[:8]
print("longest warps: dur", np.round(dur[i], 1), "start", np.round(start[i], 1), "steps", steps[i], "pairs", pairs[i], "list", length[i])
print("us per step (all): %.3f ; us per pair-step: %.3f" % (dur.sum() / steps.sum(), dur.sum() / max(pairs.sum(), 1)))
for q in (50, 100, 150, 200, 250):
    print(f"  warps still running at {q} us: {(end > q).sum()}", end=";")
print()

# where a warp's time goes (clock64 between the stage boundaries, -DRENDER_TRACE build)
sb = np.zeros((8192, 4), dtype=np.uint64)
if hasattr(n.load(), "splat_debug_render_stages") and n.load().splat_debug_render_stages(sb.ctypes.data_as(ctypes.c_void_p)) == 0:
    sb = sb.astype(np.float64)[: len(t)]
    tot = sb.sum(axis=1, keepdims=True) + 1e-9
    sh = (sb / tot)
    print("stage share of a warp's cycles (mean over warps): gather %.2f evaluate %.2f blend %.2f write-down %.2f" % tuple(sh.mean(axis=0)))
    print("cycles per step: gather %.0f evaluate %.0f blend %.0f write-down %.0f ; steps %.0f pairs/step %.1f" % (
        *(sb.sum(axis=0) / steps.sum()), steps.mean(), pairs.sum() / steps.sum()))
