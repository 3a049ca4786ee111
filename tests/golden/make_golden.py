"""Generate tests/golden/*.npz with the UNMODIFIED reference CUDA build (oracle/_ref),
run on a GPU box:

    gpurun -- 'python tests/golden/make_golden.py gpurun_out/golden'   # then copy to tests/golden/

The reference has no CPU implementation and ships no golden vectors, so
these files are what pins both the CPU oracle and the CUDA product to the reference.
Inputs are regenerated from tests/scenes.py (seeded numpy); their digest is stored.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import reflib  # noqa: E402
import scenes  # noqa: E402


def run(s, precomp=None):
    c = {k: (torch.from_numpy(v).cuda() if isinstance(v, np.ndarray) else v) for k, v in s.items()}
    ref = reflib.RefRasterizer()
    opa = c["opacities"].reshape(-1).contiguous()
    pre = torch.from_numpy(precomp).cuda() if precomp is not None else None
    R, out, radii = ref.forward(c["means3D"], opa, None if pre is not None else c["scales"],
                                None if pre is not None else c["rotations"], c["viewmatrix"],
                                c["projmatrix"], s["H"], s["W"], precomp=pre)
    g = {"num_rendered": np.int64(R), "out": out.cpu().numpy(), "radii": radii.cpu().numpy()}
    for nm in ("depths", "means2D", "transMat", "normal_opacity", "tiles_touched", "keys",
               "point_list", "ranges", "final_T", "n_contrib"):
        g[nm] = ref.state(nm).numpy()
    if pre is None:   # the reference's backward dereferences scales/rotations
        gr = ref.backward(c["dL_dout"])
        for k, v in gr.items():
            g["grad_" + k] = v.cpu().numpy()
    g["digest"] = np.frombuffer(scenes.input_digest(s).encode(), dtype=np.uint8)
    return g


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    s = scenes.make_scene("C0")
    g = run(s)
    np.savez_compressed(os.path.join(outdir, "c0.npz"), **g)
    # transMat_precomp path: feed the reference's own T back in
    sp = scenes.precomp_scene()
    gp0 = run(sp)
    gp = run(sp, precomp=gp0["transMat"].reshape(-1, 12))
    gp["precomp"] = gp0["transMat"].reshape(-1, 12)
    np.savez_compressed(os.path.join(outdir, "c0_precomp.npz"), **gp)
    np.savez_compressed(os.path.join(outdir, "edge.npz"), **run(scenes.edge_scene()))
    # the reference's own test scene (2048x512): keep every 8th pixel + channel sums
    s = scenes.reference_test_scene()
    g = run(s)
    full = g.pop("out")
    g["out_sub"] = full[:, ::8, ::8].copy()
    g["out_sum"] = full.astype(np.float64).sum(axis=(1, 2))
    ft = g.pop("final_T"); nc = g.pop("n_contrib")
    H, W = s["H"], s["W"]
    g["n_contrib_sub"] = nc.reshape(2, H, W)[:, ::8, ::8].copy()
    g["final_T_sub"] = ft.reshape(3, H, W)[:, ::8, ::8].copy()
    np.savez_compressed(os.path.join(outdir, "g0.npz"), **g)
    # kNN
    pts = scenes.knn_cloud(5000, seed=3)
    d = reflib.dist2(torch.from_numpy(pts).cuda()).cpu().numpy()
    small = [reflib.dist2(torch.from_numpy(pts[:n].copy()).cuda()).cpu().numpy() for n in (1, 2, 3, 4, 5)]
    np.savez_compressed(os.path.join(outdir, "knn.npz"), dist=d,
                        **{f"small{n}": v for n, v in zip((1, 2, 3, 4, 5), small)})
    for f in sorted(os.listdir(outdir)):
        print(f, os.path.getsize(os.path.join(outdir, f)))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/golden")
