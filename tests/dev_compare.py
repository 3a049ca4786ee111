"""Developer tool (GPU): product vs the reference CUDA build on identical inputs.

    python tests/dev_compare.py C1 [C2 ...]      (run under gpurun)
"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "splat-loam_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

import reflib  # noqa: E402
import scenes  # noqa: E402
from diff_surfel_spherical_rasterization import _C  # noqa: E402


def to_cuda(s):
    return {k: (torch.from_numpy(v).cuda() if isinstance(v, np.ndarray) else v) for k, v in s.items()}


def rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / (b.norm() + 1e-30))


def cosine(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a @ b) / (a.norm() * b.norm() + 1e-30))


def compare(name, s):
    s = to_cuda(s)
    H, W = s["H"], s["W"]
    P = s["means3D"].shape[0]
    opa = s["opacities"].reshape(-1).contiguous()
    empty = torch.empty(0, device="cuda")
    ref = reflib.RefRasterizer()
    R_ref, out_ref, radii_ref = ref.forward(s["means3D"], opa, s["scales"], s["rotations"],
                                            s["viewmatrix"], s["projmatrix"], H, W)
    import diff_surfel_spherical_rasterization._C as C
    import splat_b200_native as n
    # debug bit1 keeps the unsorted keys
    os.environ["X"] = "1"
    R, out, radii, geom, binning, img = C.rasterize_gaussians(
        s["means3D"], s["opacities"], s["scales"], s["rotations"], 1.0, empty, s["viewmatrix"],
        s["projmatrix"], H, W, False, False)
    torch.cuda.synchronize()
    print(f"== {name}: P={P} HxW={H}x{W} R_ref={R_ref} R={R}")
    print("  radii equal:", bool((radii.cpu() == radii_ref.cpu()).all()),
          " visible:", int((radii_ref > 0).sum()))
    vis = (radii_ref > 0).cpu()
    for nm in ("depths", "means2D", "tiles_touched", "point_offsets", "transMat",
               "normal_opacity"):
        a = C.debug_state(geom, binning, img, nm)
        b = ref.state(nm)
        if nm in ("depths", "tiles_touched", "point_offsets"):
            m = vis if nm == "depths" else torch.ones_like(vis)
            eq = (a.view(torch.int32)[m] == b.view(torch.int32)[m])
        else:
            k = a.numel() // P
            eq = (a.view(torch.int32).view(P, k)[vis] == b.view(torch.int32).view(P, k)[vis])
        print(f"  {nm}: bit-equal {int(eq.sum())}/{eq.numel()}")
    if R == R_ref:
        for nm in ("keys", "point_list", "ranges"):
            a = C.debug_state(geom, binning, img, nm)
            b = ref.state(nm)
            eq = a == b
            print(f"  {nm}: equal {int(eq.sum())}/{eq.numel()}")
    for nm in ("final_T", "n_contrib"):
        a = C.debug_state(geom, binning, img, nm)
        b = ref.state(nm)
        if nm == "final_T":
            eq = a.view(torch.int32) == b.view(torch.int32)
        else:
            eq = a == b
        print(f"  {nm}: bit-equal {int(eq.sum())}/{eq.numel()}")
    oc, orc = out.cpu(), out_ref.cpu()
    for c, cn in enumerate(["depth", "alpha", "nx", "ny", "nz", "median", "distortion"]):
        eq = (oc[c].view(torch.int32) == orc[c].view(torch.int32)) | ((oc[c] == 0) & (orc[c] == 0))
        print(f"  map[{cn}]: bit-equal {int(eq.sum())}/{eq.numel()}  rel={rel(oc[c], orc[c]):.3e} "
              f"maxabs={float((oc[c]-orc[c]).abs().max()):.3e}")
    # backward
    g_ref = ref.backward(s["dL_dout"])
    g = C.rasterize_gaussians_backward(s["means3D"], radii, s["scales"], s["rotations"], 1.0,
                                       empty, s["viewmatrix"], s["projmatrix"], s["dL_dout"],
                                       geom, R, binning, img, False)
    torch.cuda.synchronize()
    names = ["means2D", "opacity", "means3D", "transMat", "scales", "rotations"]
    for nm, t in zip(names, g):
        r = g_ref[nm]
        print(f"  grad[{nm}]: rel={rel(t.cpu(), r.cpu()):.3e} cos={cosine(t.cpu(), r.cpu()):.8f} "
              f"max|ref|={float(r.abs().max()):.3e}")
    # reference's own run-to-run noise (float atomics)
    g_ref2 = ref.backward(s["dL_dout"])
    for nm in ("means3D", "scales"):
        print(f"  ref-vs-ref grad[{nm}]: rel={rel(g_ref2[nm].cpu(), g_ref[nm].cpu()):.3e}")

    def timeit(fn, n=5):
        torch.cuda.synchronize()
        ts = []
        for _ in range(n):
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        return min(ts) * 1e3

    def mine():
        R, out, radii, geom, binning, img = C.rasterize_gaussians(
            s["means3D"], s["opacities"], s["scales"], s["rotations"], 1.0, empty,
            s["viewmatrix"], s["projmatrix"], H, W, False, False)
        C.rasterize_gaussians_backward(s["means3D"], radii, s["scales"], s["rotations"], 1.0,
                                       empty, s["viewmatrix"], s["projmatrix"], s["dL_dout"],
                                       geom, R, binning, img, False)

    def mine_fwd():
        C.rasterize_gaussians(s["means3D"], s["opacities"], s["scales"], s["rotations"], 1.0,
                              empty, s["viewmatrix"], s["projmatrix"], H, W, False, False)

    def theirs():
        ref.forward(s["means3D"], opa, s["scales"], s["rotations"], s["viewmatrix"],
                    s["projmatrix"], H, W)
        ref.backward(s["dL_dout"])

    def theirs_fwd():
        ref.forward(s["means3D"], opa, s["scales"], s["rotations"], s["viewmatrix"],
                    s["projmatrix"], H, W)

    print(f"  wall fwd: mine {timeit(mine_fwd):.3f} ms  ref {timeit(theirs_fwd):.3f} ms")
    print(f"  wall fwd+bwd: mine {timeit(mine):.3f} ms  ref {timeit(theirs):.3f} ms")
    nc = C.debug_state(geom, binning, img, "n_contrib")[: H * W]
    rg = C.debug_state(geom, binning, img, "ranges").view(-1, 2)
    print(f"  n_contrib: mean {float(nc.float().mean()):.1f} max {int(nc.max())}; "
          f"tile len mean {float((rg[:,1]-rg[:,0]).float().mean()):.1f} max {int((rg[:,1]-rg[:,0]).max())}")


def compare_knn(P):
    import simple_knn._C as K
    pts = torch.from_numpy(scenes.knn_cloud(P)).cuda()
    a = K.distCUDA2(pts)
    b = reflib.dist2(pts)
    torch.cuda.synchronize()
    eq = a.view(torch.int32) == b.view(torch.int32)
    print(f"== knn P={P}: bit-equal {int(eq.sum())}/{P} rel={rel(a.cpu(), b.cpu()):.3e} "
          f"maxrel={float(((a-b).abs()/(b.abs()+1e-30)).max()):.3e}")
    torch.cuda.synchronize()
    for fn, nm in ((lambda: K.distCUDA2(pts), "mine"), (lambda: reflib.dist2(pts), "ref")):
        t0 = time.perf_counter(); fn(); torch.cuda.synchronize()
        print(f"  {nm}: {(time.perf_counter()-t0)*1e3:.3f} ms")


if __name__ == "__main__":
    torch.cuda.init()
    for a in sys.argv[1:]:
        if a == "G0":
            compare("G0", scenes.reference_test_scene())
        elif a.startswith("knn"):
            compare_knn(int(a[3:]))
        else:
            compare(a, scenes.make_scene(a))
