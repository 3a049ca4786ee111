"""Developer tool (GPU): where the end-to-end step of bench.py spends its time - the same
pipeline with parts switched off, and a timeline of one step's events."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "splat-loam_b200"), os.path.join(ROOT, "tests")]
import bench  # noqa: E402
from diff_surfel_spherical_rasterization import _C  # noqa: E402

dev = torch.device("cuda", 0)
t = bench.scene_tensors("C2", 0, dev)
hp = bench.scene_tensors("C2", 0, dev, pinned=True)
H, W, P = t["H"], t["W"], t["means3D"].shape[0]
empty = torch.empty(0, device=dev)
flush = torch.empty(192 << 20, dtype=torch.uint8, device=dev)
KEYS = bench.PARAM_KEYS + ("dL_dout",)
d_in = []
for b in range(2):
    d = {k: torch.empty_like(hp[k], device=dev) for k in KEYS}
    d["viewmatrix"], d["projmatrix"] = t["viewmatrix"], t["projmatrix"]
    d_in.append(d)
gfull = [torch.empty((P, 10), device=dev) for _ in range(2)]


def frame(x, g):
    R, allmap, radii, geom, binning, img = _C.rasterize_gaussians(
        x["means3D"], x["opacities"], x["scales"], x["rotations"], 1.0, empty, x["viewmatrix"],
        x["projmatrix"], H, W, False, False)
    _C.rasterize_gaussians_backward(x["means3D"], radii, x["scales"], x["rotations"], 1.0, empty,
                                    x["viewmatrix"], x["projmatrix"], x["dL_dout"], geom, R,
                                    binning, img, False, packed=g)
    return allmap


def run(steps, do_in=True, do_out=True, do_flush=True, timeline=False):
    main = torch.cuda.current_stream(dev)
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    h_out = [[torch.empty(sh).pin_memory() for sh in ((7, H, W), (P, 10))] for _ in range(2)]
    ev_in = [torch.cuda.Event() for _ in range(2)]
    ev_used = [torch.cuda.Event() for _ in range(2)]
    ev_res = [torch.cuda.Event() for _ in range(2)]
    ev_out = [torch.cuda.Event() for _ in range(2)]
    marks = []

    def mark(stream, name, i):
        if timeline:
            e = torch.cuda.Event(enable_timing=True)
            e.record(stream)
            marks.append((name, i, e))

    def upload(i):
        b = i & 1
        with torch.cuda.stream(s_in):
            if i >= 2:
                s_in.wait_event(ev_used[b])
            mark(s_in, "in0", i)
            for k in KEYS:
                d_in[b][k].copy_(hp[k], non_blocking=True)
            mark(s_in, "in1", i)
            ev_in[b].record(s_in)

    def loop(n):
        # the order bench.pipelined_e2e uses for plain copies: step i + 1's upload ahead of
        # step i's frame
        if do_in:
            upload(0)
        for i in range(n):
            b = i & 1
            if do_in:
                if i + 1 < n:
                    upload(i + 1)
                main.wait_event(ev_in[b])
            if i >= 2 and do_out:
                main.wait_event(ev_out[b])
            if do_flush:
                flush.zero_()
            mark(main, "f0", i)
            allmap = frame(d_in[b], gfull[b])
            mark(main, "f1", i)
            ev_used[b].record(main)
            ev_res[b].record(main)
            if do_out:
                with torch.cuda.stream(s_out):
                    s_out.wait_event(ev_res[b])
                    mark(s_out, "out0", i)
                    h_out[b][0].copy_(allmap, non_blocking=True)
                    h_out[b][1].copy_(gfull[b], non_blocking=True)
                    allmap.record_stream(s_out)
                    mark(s_out, "out1", i)
                    ev_out[b].record(s_out)
        main.wait_stream(s_out)
        main.wait_stream(s_in)

    timeline_save, timeline = timeline, False
    loop(4)
    timeline = timeline_save
    torch.cuda.synchronize()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(main)
    s_in.wait_event(a)
    loop(steps)
    z.record(main)
    z.synchronize()
    if timeline:
        for name, i, e in marks:
            if 3 <= i <= 5:
                print(f"  step {i} {name:5s} at {a.elapsed_time(e):8.3f} ms")
    return a.elapsed_time(z) / steps


for d in d_in:
    for k in KEYS:
        d[k].copy_(hp[k])
for label, kw in (("full", {}), ("no flush", dict(do_flush=False)), ("no D2H", dict(do_out=False)),
                  ("no H2D", dict(do_in=False)), ("compute only", dict(do_in=False, do_out=False))):
    print(f"{label:14s} {run(100, **kw):.3f} ms / step")
run(8, timeline=True)
