"""Developer tool (GPU): run a few fwd+bwd frames of one config for ncu / timing."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "splat-loam_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import scenes  # noqa: E402
from diff_surfel_spherical_rasterization import _C as C  # noqa: E402

cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
what = sys.argv[3] if len(sys.argv) > 3 else "raster"
if what == "knn":
    import simple_knn._C as K
    pts = torch.from_numpy(scenes.knn_cloud(int(cfg))).cuda()
    for _ in range(iters):
        K.distCUDA2(pts)
    torch.cuda.synchronize()
    sys.exit(0)
s = scenes.make_scene(cfg)
s = {k: (torch.from_numpy(v).cuda() if isinstance(v, np.ndarray) else v) for k, v in s.items()}
H, W = s["H"], s["W"]
empty = torch.empty(0, device="cuda")
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
for it in range(iters):
    ev[0].record()
    R, out, radii, geom, binning, img = C.rasterize_gaussians(
        s["means3D"], s["opacities"], s["scales"], s["rotations"], 1.0, empty, s["viewmatrix"],
        s["projmatrix"], H, W, False, False)
    ev[1].record()
    C.rasterize_gaussians_backward(s["means3D"], radii, s["scales"], s["rotations"], 1.0, empty,
                                   s["viewmatrix"], s["projmatrix"], s["dL_dout"], geom, R,
                                   binning, img, False)
    ev[2].record()
    torch.cuda.synchronize()
    print(f"iter {it}: fwd {ev[0].elapsed_time(ev[1]):.3f} ms  bwd {ev[1].elapsed_time(ev[2]):.3f} ms  R={R}")

# host-side cost of one forward call (the GPU work is asynchronous in hint mode)
import time
torch.cuda.synchronize()
for rep in range(3):
    t0 = time.perf_counter()
    r = C.rasterize_gaussians(s["means3D"], s["opacities"], s["scales"], s["rotations"], 1.0,
                              empty, s["viewmatrix"], s["projmatrix"], H, W, False, False)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"host: call {1e3*(t1-t0):.3f} ms, +sync {1e3*(t2-t1):.3f} ms")
