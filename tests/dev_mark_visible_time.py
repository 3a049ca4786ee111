"""Developer tool (GPU): mark_visible time at C2 (CUDA events)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "splat-loam_b200"), os.path.join(ROOT, "tests")]
import scenes
from diff_surfel_spherical_rasterization import _C
s = scenes.make_scene("C2")
m, v, p = (torch.from_numpy(s[k]).cuda() for k in ("means3D", "viewmatrix", "projmatrix"))
for _ in range(5):
    _C.mark_visible(m, v, p)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(100):
    _C.mark_visible(m, v, p)
b.record(); b.synchronize()
print(f"mark_visible, P={m.shape[0]}: {a.elapsed_time(b) * 10:.2f} us per call (incl. the torch.zeros of the result)")
