"""Shared helpers for the -m gpu tests: run the CUDA product through its reference-facing
Python surface (which calls the C-ABI) and fetch every intermediate as numpy."""
import numpy as np
import torch


def to_cuda(s):
    return {k: (torch.from_numpy(v).cuda() if isinstance(v, np.ndarray) else v) for k, v in s.items()}


STATE = ("depths", "means2D", "transMat", "normal_opacity", "tiles_touched", "point_offsets",
         "keys", "point_list", "ranges", "final_T", "n_contrib")


def run_product(s, precomp=None, backward=True, state=STATE):
    from diff_surfel_spherical_rasterization import _C
    c = to_cuda(s)
    empty = torch.empty(0, device="cuda")
    pre = torch.from_numpy(np.ascontiguousarray(precomp)).cuda() if precomp is not None else empty
    scales = empty if precomp is not None else c["scales"]
    rots = empty if precomp is not None else c["rotations"]
    R, out, radii, geom, binning, img = _C.rasterize_gaussians(
        c["means3D"], c["opacities"], scales, rots, 1.0, pre, c["viewmatrix"], c["projmatrix"],
        s["H"], s["W"], False, False)
    res = {"num_rendered": R, "out": out.cpu().numpy(), "radii": radii.cpu().numpy()}
    for nm in state:
        res[nm] = _C.debug_state(geom, binning, img, nm).numpy()
    if backward:
        g = _C.rasterize_gaussians_backward(c["means3D"], radii, scales, rots, 1.0, pre,
                                            c["viewmatrix"], c["projmatrix"], c["dL_dout"], geom,
                                            R, binning, img, False)
        for nm, t in zip(("means2D", "opacity", "means3D", "transMat", "scales", "rotations"), g):
            res["grad_" + nm] = t.cpu().numpy()
    torch.cuda.synchronize()
    return res


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)
