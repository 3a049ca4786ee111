"""Developer tool (GPU): distCUDA2 time at a given size (CUDA events, after warm-up)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "splat-loam_b200"), os.path.join(ROOT, "tests")]
import scenes
import simple_knn._C as K
for P in [int(a) for a in sys.argv[1:]] or [4000000]:
    pts = torch.from_numpy(scenes.knn_cloud(P)).cuda()
    for _ in range(3):
        K.distCUDA2(pts)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        K.distCUDA2(pts)
    b.record(); b.synchronize()
    print(f"P={P}: {a.elapsed_time(b) / 5:.3f} ms per call")
