#!/usr/bin/env python
"""bench.py — surfel raster fwd+bwd frames/s (BASELINE.json metric) on N GPUs of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C2]

A "step" is one pass of the hot path over one synthetic frame per GPU: rasterizer forward +
backward through the reference-facing `_C` calls (which go through the C-ABI of
libsplat_b200.so).  N=1 runs BASELINE config C2 (64x1024 image, 500k surfels).  N>1 is the
keyframe-sharded window (weak scaling): rank r renders its own keyframe of the same surfel set
(equal work per GPU, see scene_tensors) and the per-surfel [P,10] gradient buffers are summed
over NVLink peer memory (push fused into the VJP kernel;
SPLAT_B200_XCHG=nccl selects one NCCL all-reduce instead); before anything is timed the
exchange's sum is checked against an NCCL all-reduce of the same buffers
(`exchange_check`).  At N>1 the line also carries `window`: the BASELINE window configs
(C3: 8 keyframes at 2/4/8 GPUs, C4: 32 keyframes at 8, BASELINE.md's poses) through
KeyframeWindow.gradients, with the same window timed on ONE GPU in the same run.

Rank 0 prints ONE JSON line.  See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "splat-loam_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import scenes  # noqa: E402

METRIC = "surfel raster fwd+bwd frames/s"
UNIT = "frames/s"


# ------------------------------------------------------------------ helpers ------------
def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """SM clocks and throttle reasons DURING the timed region.  NVML from a sampling thread
    (what nvidia-smi reads, without a process spawn or its driver-wide locks every few
    milliseconds); falls back to `nvidia-smi -lms` when pynvml is missing."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index, period_s=0.005):
        self.rows, self.proc, self.stop_flag, self.source = [], None, False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index
            h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            bits = [(0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"),
                    (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap")]

            def loop():
                while not self.stop_flag:
                    try:
                        sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                        rs = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h) \
                            if hasattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons") \
                            else pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                        self.rows.append([str(sm), str(mx)] +
                                         ["Active" if rs & b else "Not Active" for b, _ in bits])
                    except Exception:
                        pass
                    time.sleep(period_s)
            self.t = threading.Thread(target=loop, daemon=True)
            self.t.start()
            self.source = "nvml"
            return
        except Exception:
            pass
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "50"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
            self.source = "nvidia-smi"
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        elif self.source == "nvml":
            self.t.join(timeout=1)
        else:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        sm, mx, reasons = [], None, set()
        for r in list(self.rows):
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                for nm, v in zip(self.NAMES, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm), "source": self.source}


def scene_tensors(cfg, keyframe, device, pinned=False):
    # The weak-scaling line measures what the exchange costs on top of a frame, so every GPU
    # gets the SAME work: rank r's keyframe is the scene rotated by 22.5 degrees * r about the
    # sensor's vertical axis (ranges and footprints unchanged).  BASELINE.md's poses (0.5 m and
    # 2 degrees per keyframe) move the sensor through the synthetic cloud and change a frame's
    # cost by up to 40 % (its longest tile list: 5.3k entries at keyframe 0, 7.4k at keyframe 1,
    # CPU oracle) - a step would then time the slowest keyframe, not the exchange.  Those poses
    # are what the `window` entries run (run_window), where the keyframes are dealt round-robin.
    s = scenes.make_scene(cfg, keyframe=0)
    s["viewmatrix"] = scenes.yaw_viewmatrix(22.5 * keyframe)
    t = {}
    for k in ("means3D", "opacities", "scales", "rotations", "viewmatrix", "projmatrix", "dL_dout"):
        x = torch.from_numpy(s[k])
        t[k] = x.pin_memory() if pinned else x.to(device)
    t["H"], t["W"] = s["H"], s["W"]
    return t


def workload(cfg, H, W, P):
    """The same string for both arms (the implementation goes under config.implementation)."""
    return (f"{cfg}: {H}x{W} spherical range image, {P} surfels, fwd+bwd per frame, one keyframe "
            "per GPU (rank r: the scene yawed by 22.5 deg * r - equal work on every GPU)")


L2_NOTE = "192 MB buffer (1.5x the 126 MB L2) written between timed steps"


def timed_steps(step_fn, steps, flush, barrier):
    """The timed region of both arms: a barrier + synchronize on both sides, the L2 flush
    outside each step's event pair, CUDA events on the current stream.  Returns total ms."""
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(steps)]
    barrier()
    for a, b in ev:
        flush.zero_()                     # L2 flush between timed iterations (not timed)
        a.record()
        step_fn()
        b.record()
    barrier()
    return sum(a.elapsed_time(b) for a, b in ev)


def algorithmic_bytes(P, N, Nc, HW):
    """compulsory bytes per stage of one fwd+bwd frame."""
    b = {
        "preprocess": 132 * P,                 # 40 R + 84 W + scan 4 R + 4 W
        "emit": 20 * P + 12 * N,               # duplicateWithKeys
        "sort": (8 + 24 * 6 + 8) * N,          # histogram + 6 passes x (12 R + 12 W) + ranges
        "render_forward": 76 * Nc + 48 * HW,
        "render_backward": 180 * Nc + 48 * HW,  # 76 gather + 104 gradient RMW
        "backward_preprocess": 236 * P,        # 112 zero-init + 88 R + 36 W
    }
    b["frame"] = sum(b.values())
    return b


def consumed_instances(nc, ranges, H, W):
    """N_c = sum over tiles of max over the tile's pixels of n_contrib."""
    gx, gy = (W + 15) // 16, (H + 15) // 16
    last = nc[: H * W].reshape(H, W).astype(np.int64)
    pad = np.zeros((gy * 16, gx * 16), dtype=np.int64)
    pad[:H, :W] = last
    return int(pad.reshape(gy, 16, gx, 16).max(axis=(1, 3)).sum())


def pipelined_e2e(dev, steps, flush, copy_in, frame_fn, out_specs, prefetch=True):
    """End-to-end protocol shared by both arms: every step brings that step's inputs from
    pinned host memory into device slot b (`copy_in(b)`, enqueued on a side stream: H2D copies
    and, for a sharded upload, the all-gather that completes them), runs `frame_fn(b)` -> list
    of device result tensors, and copies the results back to pinned host memory, all inside
    the timed region.  Inputs, results and host buffers are double-buffered so the copies of
    neighbouring steps overlap the compute (copy engines on side streams); the L2 flush is also
    inside the timed region.  `prefetch`: step i + 1's upload is enqueued ahead of step i's
    frame (plain copies); without it, after the frame (an upload that ends in a collective:
    measured on 8 GPUs, the all-gather enqueued early costs more than its late start).
    Returns (ms per step, D2H bytes per step)."""
    main = torch.cuda.current_stream(dev)
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    h_out = [[torch.empty(sh, dtype=torch.float32).pin_memory() for sh in out_specs] for _ in range(2)]
    ev_in = [torch.cuda.Event() for _ in range(2)]
    ev_used = [torch.cuda.Event() for _ in range(2)]
    ev_res = [torch.cuda.Event() for _ in range(2)]
    ev_out = [torch.cuda.Event() for _ in range(2)]

    def upload(i):
        b = i & 1
        with torch.cuda.stream(s_in):
            if i >= 2:
                s_in.wait_event(ev_used[b])        # the frame two steps ago has read them
            copy_in(b)
            ev_in[b].record(s_in)

    def run(n):
        # The upload of step i + 1 is enqueued BEFORE the frame of step i: a frame's forward
        # blocks the host until its instance count has landed (both arms do), and an upload
        # enqueued after that wait would start late and leave the GPU idle at the next step.
        if prefetch:
            upload(0)
        for i in range(n):
            b = i & 1
            if prefetch and i + 1 < n:
                upload(i + 1)
            elif not prefetch:
                upload(i)
            main.wait_event(ev_in[b])
            if i >= 2:
                main.wait_event(ev_out[b])             # result slot b has been drained
            flush.zero_()
            res = frame_fn(b)
            ev_used[b].record(main)
            ev_res[b].record(main)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_res[b])
                for h, r in zip(h_out[b], res):
                    h.copy_(r, non_blocking=True)
                    r.record_stream(s_out)
                ev_out[b].record(s_out)
        main.wait_stream(s_out)
        main.wait_stream(s_in)

    run(4)
    torch.cuda.synchronize(dev)
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(main)
    s_in.wait_event(a)
    run(steps)
    z.record(main)
    z.synchronize()
    return a.elapsed_time(z) / steps, sum(h.numel() * 4 for h in h_out[0])


PARAM_KEYS = ("means3D", "opacities", "scales", "rotations")


# ------------------------------------------------------------------ our arm ------------
def run_ours(args, rank, world, local_rank):
    import torch.distributed as dist
    from diff_surfel_spherical_rasterization import _C
    import splat_b200_native as native
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    t = scene_tensors(args.config, rank, dev)          # rank r renders keyframe r
    H, W, P = t["H"], t["W"], t["means3D"].shape[0]
    empty = torch.empty(0, device=dev)
    gbuf = torch.empty((P, 10), dtype=torch.float32, device=dev)
    flush = torch.empty(192 << 20, dtype=torch.uint8, device=dev)   # > 126 MB of L2

    # keyframe-sharded window: the [P,10] gradients are summed over the ranks through NVLink
    # peer memory, the push half fused into the per-surfel VJP kernel (SPLAT_B200_XCHG=nccl:
    # one NCCL all-reduce instead)
    xchg = None
    if world > 1 and os.environ.get("SPLAT_B200_XCHG", "peer") != "nccl":
        from keyframe_window import PeerExchange
        try:
            xchg = PeerExchange(P, dev)
        except RuntimeError as e:    # all ranks agree on this outcome (see PeerExchange)
            if rank == 0:
                print(f"[bench] {e}; falling back to the NCCL all-reduce", file=sys.stderr)

    def fwd_bwd(x, packed=None, use_xchg=False):
        R, allmap, radii, geom, binning, img = _C.rasterize_gaussians(
            x["means3D"], x["opacities"], x["scales"], x["rotations"], 1.0, empty,
            x["viewmatrix"], x["projmatrix"], H, W, False, False)
        # the [P,10] packed gradients come straight out of the per-surfel VJP kernel
        g = _C.rasterize_gaussians_backward(
            x["means3D"], radii, x["scales"], x["rotations"], 1.0, empty, x["viewmatrix"],
            x["projmatrix"], x["dL_dout"], geom, R, binning, img, False,
            packed=packed, xchg=xchg.handle.value if use_xchg else None)
        return R, allmap, g, (geom, binning, img)

    def frame(x):
        if xchg:
            out = fwd_bwd(x, None, True)
            xchg.allreduce()
        else:
            out = fwd_bwd(x, gbuf)
            if world > 1:
                dist.all_reduce(gbuf)
        return out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        flush.zero_()
        R, allmap, g, bufs = frame(t)
    torch.cuda.synchronize()
    nc = _C.debug_state(*bufs, "n_contrib").numpy().view(np.uint32)
    rg = _C.debug_state(*bufs, "ranges").numpy()
    Nc = consumed_instances(nc, rg, H, W)

    # ---- N>1: the exchange's sum against an NCCL all-reduce of the same buffers -------
    exchange_check = None
    if world > 1:
        fwd_bwd(t, gbuf)                              # this rank's [P,10], no exchange
        want = gbuf.clone()
        dist.all_reduce(want)
        if xchg:
            fwd_bwd(t, None, True)
            got = xchg.allreduce().clone()
            xchg.check()
        else:
            got = want.clone()
        rel = float((got.double() - want.double()).norm() / want.double().norm())
        ref0 = got.clone()
        dist.broadcast(ref0, src=0)
        same = torch.tensor([1 if torch.equal(ref0, got) else 0], device=dev)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        relt = torch.tensor([rel], device=dev, dtype=torch.float64)
        dist.all_reduce(relt, op=dist.ReduceOp.MAX)
        exchange_check = {"rel_err": float(relt[0]), "bit_identical_across_ranks": bool(same[0]),
                          "against": "dist.all_reduce (NCCL) of the same per-rank [P,10] buffers",
                          "path": "peer memory" if xchg else "nccl (nothing to compare)"}
        assert exchange_check["rel_err"] < 1e-5 and exchange_check["bit_identical_across_ranks"], \
            exchange_check

    # ---- per-stage times (instrumented pass: every stage bracketed by events) ----
    native.profile_enable(True)
    barrier()
    nprof = min(args.steps, 20)
    for _ in range(nprof):
        flush.zero_()
        frame(t)
    torch.cuda.synchronize()
    stage_ms = {k: v / nprof for k, v in native.profile_read().items()}
    dom = max(stage_ms, key=stage_ms.get) if stage_ms else "render_backward"

    # ---- device-resident throughput (`value`): only the dominant kernel keeps its event
    # pair (the roofline's launch duration is measured inside this timed region) ----
    sampler = ClockSampler(local_rank) if rank == 0 else None
    native.profile_enable(True, only=dom)
    launches0 = native.launch_count()
    total_ms = timed_steps(lambda: frame(t), args.steps, flush, barrier)
    launches = native.launch_count() - launches0
    dom_ms = native.profile_read().get(dom, 0.0) / args.steps
    native.profile_enable(False)
    clocks = sampler.stop() if sampler else None

    # ---- the same through the reference's stock 14-argument backward (no packed=) ----
    stock_ms = None
    if world == 1:
        stock_ms = timed_steps(lambda: fwd_bwd(t), args.steps, flush, barrier) / args.steps

    # ---- end to end through the same calls with HOST buffers ----
    # N = 1: the frame's parameters, dL_dout in; maps and [P,10] gradients out.  N > 1: the
    # parameters are replicated, so each rank uploads 1/N of their rows and the ranks all-gather
    # them over NVLink; every rank uploads its own keyframe's dL_dout, downloads its own maps
    # and the slice of the summed gradients it owns in the exchange.  The byte counts are the
    # whole job's.
    hp = scene_tensors(args.config, rank, dev, pinned=True)
    rows = (P + world - 1) // world
    if xchg:
        rows = (rows + 255) // 256 * 256              # the exchange's slice (XCHG_SLICE_ALIGN)
    lo, hi = min(P, rank * rows), min(P, (rank + 1) * rows)
    shard = world > 1 and P % world == 0
    d_in = []
    for b in range(2):
        d = {k: torch.empty_like(hp[k], device=dev) for k in PARAM_KEYS + ("dL_dout",)}
        d["viewmatrix"], d["projmatrix"] = t["viewmatrix"], t["projmatrix"]
        d_in.append(d)
    if shard:
        # the rank's rows of the four parameter tensors travel as ONE packed [P/N,10] block:
        # one H2D copy, one all-gather, then four strided device copies into the tensors the
        # rasterizer reads (a collective call costs the host far more than a copy kernel)
        per = P // world
        cols = {"means3D": (0, 3), "opacities": (3, 4), "scales": (4, 6), "rotations": (6, 10)}
        h_slice = torch.cat([hp[k][rank * per:(rank + 1) * per] for k in PARAM_KEYS], dim=1).pin_memory()
        d_slice = [torch.empty_like(h_slice, device=dev) for _ in range(2)]
        d_full = [torch.empty((P, 10), dtype=torch.float32, device=dev) for _ in range(2)]

    def copy_in(b):
        d_in[b]["dL_dout"].copy_(hp["dL_dout"], non_blocking=True)
        if shard:
            d_slice[b].copy_(h_slice, non_blocking=True)
            dist.all_gather_into_tensor(d_full[b], d_slice[b])
            for k, (c0, c1) in cols.items():
                d_in[b][k].copy_(d_full[b][:, c0:c1])
        else:
            for k in PARAM_KEYS:
                d_in[b][k].copy_(hp[k], non_blocking=True)

    gslots = [torch.empty((hi - lo, 10), dtype=torch.float32, device=dev) for _ in range(2)]
    gfull = [gbuf, torch.empty_like(gbuf)]

    def e2e_frame(b):
        x = d_in[b]
        if xchg:       # the exchange owns ONE result array: park the owned slice in the slot
            _, allmap, _, _ = fwd_bwd(x, None, True)
            gslots[b].copy_(xchg.allreduce()[lo:hi])
            return [allmap, gslots[b]]
        _, allmap, _, _ = fwd_bwd(x, gfull[b])
        if world > 1:
            dist.all_reduce(gfull[b])
            gslots[b].copy_(gfull[b][lo:hi])
            return [allmap, gslots[b]]
        return [allmap, gfull[b]]

    barrier()
    e2e_ms_step, d2h = pipelined_e2e(dev, args.steps, flush, copy_in, e2e_frame,
                                     [(7, H, W), (hi - lo, 10)], prefetch=not shard)
    barrier()
    if xchg:
        xchg.check()
    e2e_ms = e2e_ms_step * args.steps
    h2d = 7 * H * W * 4 * world + P * 10 * 4 if shard else (7 * H * W + P * 10) * 4 * world
    d2h_job = 7 * H * W * 4 * world + P * 10 * 4 if world > 1 else d2h

    window = None
    if world > 1 and not args.no_window:
        window = [run_window("C3", 8, rank, world, dev, flush)]
        if world >= 8:
            window.append(run_window("C4", 32, rank, world, dev, flush))

    if world > 1:   # max over ranks, on the device timeline
        tt = torch.tensor([total_ms, e2e_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total_ms, e2e_ms = float(tt[0]), float(tt[1])
    if rank != 0:
        return None

    ms_per_step = total_ms / args.steps
    value = world * 1e3 / ms_per_step
    ab = algorithmic_bytes(P, R, Nc, H * W)
    peak, peak_src = measured_peak_gbs()
    stage_ms[dom] = dom_ms or stage_ms.get(dom)
    traffic = winst = None
    try:
        prof = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = prof.get(dom)
        winst = prof.get("warp_instructions", {}).get(dom) if args.config == "C2" else None
    except Exception:
        pass
    achieved = ab[dom] / (stage_ms[dom] * 1e-3) / 1e9 if stage_ms.get(dom) else None
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload(args.config, H, W, P),
                   "implementation": "libsplat_b200.so through the drop-in _C surface",
                   "exchange": (("[P,10] gradients summed over NVLink peer memory (push fused "
                                 "into the VJP kernel)" if xchg else
                                 "NCCL all-reduce of the [P,10] gradients")
                                if world > 1 else "none (one GPU)"),
                   "instances": R, "consumed_instances": Nc, "l2": L2_NOTE,
                   "parallelism": f"keyframe-sharded x{world}"},
        "e2e": {"value": world * 1e3 / (e2e_ms / args.steps), "unit": UNIT,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h_job,
                "note": ("whole-job bytes; parameters uploaded 1/N per rank and all-gathered over "
                         "NVLink, each rank downloads its maps and its owned gradient slice")
                if shard else "whole-job bytes"},
        # counted by the library (splat_launch_count) over the timed region, all of rank 0's
        # kernels: per frame 9 forward + 3 backward (+ the exchange's reduce kernel)
        "gpu_launches": launches,
        "stage_ms": stage_ms,
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak,
                     "peak_source": peak_src, "unit": "GB/s",
                     "frac": achieved / peak if achieved else None, "traffic": traffic,
                     "traffic_source": "profiles/traffic.json (ncu --set full of an earlier "
                                       "session, a constant - not measured by this run)",
                     "algorithmic_bytes": ab[dom]},
        "frame_roofline": {"algorithmic_bytes": ab["frame"],
                           "achieved": ab["frame"] / (ms_per_step * 1e-3) / 1e9, "unit": "GB/s",
                           "frac": ab["frame"] / (ms_per_step * 1e-3) / 1e9 / peak},
        "clocks": clocks,
    }
    if stock_ms:
        line["stock_signature"] = {
            "value": 1e3 / stock_ms, "unit": UNIT, "ms_per_step": stock_ms,
            "what": "the same frames through the reference's 14-argument "
                    "rasterize_gaussians_backward (no packed= extension)"}
    if exchange_check:
        line["exchange_check"] = exchange_check
    if window:
        line["window"] = window
    if winst and stage_ms.get(dom) and clocks and clocks.get("sm_mhz"):
        # the compositing kernels are bound by instruction issue, not by HBM: warp instructions
        # per launch (ncu, profiles/) / live launch duration, against 4 issue slots per SM and clock
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        ipeak = sms * 4 * clocks["sm_mhz"] * 1e6
        line["issue_roofline"] = {"kernel": dom, "warp_instructions": winst,
                                  "warp_instructions_source": "profiles/traffic.json (session "
                                                              "constant, ncu smsp__inst_executed)",
                                  "achieved": winst / (stage_ms[dom] * 1e-3), "peak": ipeak,
                                  "unit": "warp-inst/s", "frac": winst / (stage_ms[dom] * 1e-3) / ipeak}
    if world == 1:
        line["post"] = post_stage(allmap.detach(), t)
        line["optimizer"] = adam_stage(P, dev)
        line["photometric"] = photometric_stage(dev)
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args.config)
    return line


def run_window(cfg, K, rank, world, dev, flush, reps=10):
    """a window of K keyframes of `cfg` (BASELINE poses) through
    KeyframeWindow.gradients, sharded over the ranks, and the same window on ONE GPU (every
    rank runs it alone, at the same time, without communication) - all in this run.
    t = milliseconds per window, max over ranks; of_linear = t(1 GPU) / (G * t(G GPUs))."""
    import torch.distributed as dist
    from keyframe_window import KeyframeWindow
    s = scenes.make_scene(cfg)
    H, W = s["H"], s["W"]
    params = tuple(torch.from_numpy(s[k]).to(dev) for k in PARAM_KEYS)
    proj = torch.from_numpy(s["projmatrix"]).to(dev)
    cams = [(torch.from_numpy(scenes.viewmatrix(k)).to(dev), proj, H, W) for k in range(K)]
    dL = torch.from_numpy(s["dL_dout"]).to(dev)
    up = lambda k, allmap: dL  # noqa: E731
    win, solo = KeyframeWindow(), KeyframeWindow()

    def timed(fn):
        for _ in range(2):
            fn()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
              for _ in range(reps)]
        dist.barrier()
        torch.cuda.synchronize()
        for a, b in ev:
            flush.zero_()
            a.record()
            fn()
            b.record()
        dist.barrier()
        torch.cuda.synchronize()
        tt = torch.tensor([sum(a.elapsed_time(b) for a, b in ev) / reps], device=dev,
                          dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt[0])

    t_g = timed(lambda: win.gradients(params, cams, up, check=False))
    win.exchange_check()
    got = win.gradients(params, cams, up).clone()
    t_1 = timed(lambda: solo.accumulate(params, cams, up, local_only=True))
    want = solo.accumulate(params, cams, up, local_only=True)[0]
    rel = torch.tensor([float((got.double() - want.double()).norm() / want.double().norm())],
                       device=dev, dtype=torch.float64)
    dist.all_reduce(rel, op=dist.ReduceOp.MAX)
    del win, solo
    return {"config": f"{cfg}: {H}x{W}, {params[0].shape[0]} surfels, {K}-keyframe window "
                      f"sharded over {world} GPUs (keyframe k on rank k mod {world})",
            "keyframes": K, "ms_per_window": t_g, "keyframes_per_s": K * 1e3 / t_g,
            "ms_per_window_1gpu": t_1, "of_linear": t_1 / (world * t_g),
            "rel_err_vs_1gpu": float(rel[0])}


def photometric_stage(dev, iters=20):
    """SURVEY.md §8(f) rank 3, reported beside the headline: one photometric alignment (10
    Gauss-Newton iterations, 64x1024 range images of a synthetic room, generic initial guess)
    through gsaligner_photometric.PhotometricAligner, and through the unmodified reference
    build (oracle/_ref/libref_photo.so) when it is there.  Wall-clock per call: both end with
    the transform on the host."""
    from gsaligner_photometric import PhotometricAligner
    import reflib
    a, K = scenes.room_range_image(64, 1024, 33.2, seed=2)
    pose = scenes.pose(2)
    pose[:3, 3] *= 0.2
    b, _ = scenes.room_range_image(64, 1024, 33.2, pose=pose, seed=3)
    T0 = torch.from_numpy(scenes.small_motion())
    proj = torch.eye(4)
    proj[:3, :3] = torch.from_numpy(K).T
    da, db = torch.from_numpy(a)[None].to(dev), torch.from_numpy(b)[None].to(dev)
    al = PhotometricAligner(image_height=64, image_width=1024)
    al.set_reference(da, None, proj.to(dev))
    al.set_query(db, None, proj.to(dev))

    def clock(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(iters):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / iters * 1e3

    res = {"fused_ms": clock(lambda: al.align_photometric(T0)),
           "what": "photometric alignment, 64x1024, 10 iterations, transform back on the host"}
    if reflib.photo_available():
        r = reflib.RefPhotometric(64, 1024)
        r.set_reference(da[0], K)
        r.set_query(db[0])
        res["reference_ms"] = clock(lambda: r.align(T0.numpy(), 10))
        res["max_abs_diff_of_transforms"] = float(
            (al.align_photometric(T0)[0] - r.align(T0.numpy(), 10)[0]).abs().max())
    return res


def adam_stage(P, dev, iters=50):
    """SURVEY.md §8(f) rank 2, reported beside the headline: from the [P,10] gradients to
    updated parameters.  `torch_ms`: what the reference runs (activations, autograd through
    them and the scale regulariser, torch.optim.Adam.step); `fused_ms`: the same activations
    (for the next frame's rasterizer call) + ONE kernel (fused_adam.activation_adam_step)."""
    from fused_adam import activation_adam_step
    g = torch.Generator(device=dev).manual_seed(1)
    raw = {"xyz": torch.randn(P, 3, device=dev, generator=g),
           "opacity": torch.randn(P, 1, device=dev, generator=g),
           "scaling": torch.randn(P, 2, device=dev, generator=g) - 1.0,
           "rotation": torch.randn(P, 4, device=dev, generator=g)}
    G = torch.randn(P, 10, device=dev, generator=g)
    lrs = {"xyz": 1.6e-4, "opacity": 5e-2, "scaling": 5e-3, "rotation": 1e-3}

    def make():
        ps = {k: torch.nn.Parameter(v.clone()) for k, v in raw.items()}
        return ps, torch.optim.Adam([{"params": [ps[k]], "lr": lrs[k], "name": k} for k in ps],
                                    lr=0.0, eps=1e-15)

    def acts(ps):
        return (ps["xyz"], torch.sigmoid(ps["opacity"]), torch.exp(ps["scaling"]),
                torch.nn.functional.normalize(ps["rotation"]))

    pa, oa = make()
    pb, ob = make()

    def torch_step():
        oa.zero_grad(set_to_none=True)
        a = acts(pa)
        smax = a[2].max(dim=1).values
        reg = (0.2 * (smax[smax >= 0.5] - 0.5)).sum()
        torch.autograd.backward(list(a) + [reg], [G[:, 0:3], G[:, 3:4], G[:, 4:6], G[:, 6:10],
                                                  torch.ones_like(reg)])
        oa.step()

    def fused_step():
        with torch.no_grad():
            acts(pb)
        activation_adam_step(ob, G, 0.5, 0.2)

    res = {}
    for name, fn in (("fused_ms", fused_step), ("torch_ms", torch_step)):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        b.synchronize()
        res[name] = a.elapsed_time(b) / iters
    res["what"] = f"activations + their backward + scale regulariser + Adam, P={P}"
    return res


def post_stage(allmap, t, iters=50):
    """SURVEY.md §8(f) rank 1, reported beside the headline: render()'s map post-processing,
    forward + backward, as the fused kernel pair (gaussian_renderer.postprocess) and as the
    reference's own sequence of PyTorch ops (the oracle's torch restatement, which reproduces
    the reference's render() to 1e-7) on the same GPU.  CUDA events, after warm-up."""
    import gaussian_renderer as gr
    from oracle import render_post as rp
    V, Pm = t["viewmatrix"], t["projmatrix"]
    g = {k: torch.randn(s, device=allmap.device) for k, s in
         (("rend_alpha", allmap[1:2].shape), ("rend_normal", allmap[2:5].shape),
          ("rend_dist", allmap[6:7].shape), ("surf_depth", allmap[0:1].shape),
          ("surf_normal", allmap[2:5].shape))}

    def step(fn):
        a = allmap.clone().requires_grad_(True)
        out = fn(a, V, Pm, 0.0)
        torch.autograd.backward([out[k] for k in g], [g[k] for k in g])
        return a.grad

    res = {}
    for name, fn in (("fused_ms", gr.postprocess), ("torch_ops_ms", rp.post_forward)):
        for _ in range(5):
            step(fn)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            step(fn)
        b.record()
        b.synchronize()
        res[name] = a.elapsed_time(b) / iters
    res["what"] = "render() map post-processing fwd+bwd per frame (incl. clone + autograd glue)"

    # the same with the mapper's image losses on top (slam/mapper.py:158-195)
    import mapping_loss as ml

    class Cam:
        world_view_transform, projection_matrix = V, Pm
        image_depth = (allmap[5:6] + 0.1).contiguous()
        image_valid = (allmap[1:2] > 0).to(torch.uint8)

    def fused_loss():
        a = allmap.clone().requires_grad_(True)
        ml.mapping_image_losses(a, Cam, 0.0, 0.1, 0.1)["total"].backward()

    def torch_loss():
        a = allmap.clone().requires_grad_(True)
        pkg = rp.post_forward(a, V, Pm, 0.0)
        rp.mapper_image_losses(pkg, Cam.image_depth, Cam.image_valid, 0.1, 0.1)[3].backward()

    for name, fn in (("fused_with_losses_ms", fused_loss), ("torch_ops_with_losses_ms", torch_loss)):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        b.synchronize()
        res[name] = a.elapsed_time(b) / iters
    return res


# ------------------------------------------------------------------ baselines ----------
def cpu_baseline(cfg, full=True):
    """The plain-C oracle (a PORT of the reference algorithm; the reference itself has no CPU
    implementation) timed on the host cores for one whole fwd+bwd frame of the workload."""
    from oracle import oracle
    s = scenes.make_scene(cfg)
    oracle.build()
    t0 = time.perf_counter()
    f = oracle.rasterize_forward(s["means3D"], s["scales"], s["rotations"], s["opacities"],
                                 s["viewmatrix"], s["projmatrix"], s["H"], s["W"])
    oracle.rasterize_backward(f, s["dL_dout"])
    dt = time.perf_counter() - t0
    return {"value": 1.0 / dt, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
            "sample": f"1 full fwd+bwd frame of {cfg} ({dt:.2f} s, OpenMP on all cores; "
                      "preprocess/sort single-threaded)"}


def run_reference(args, rank, world, local_rank):
    """The UNMODIFIED reference CUDA rasterizer (oracle/_ref: the reference's .cu sources
    compiled where they lie + oracle/ref_shim.cu, a lean extern "C" wrapper) on the same frames.
    The reference has no CPU implementation of this path; its CUDA build is what BASELINE.json's
    target (>= 3x) is quoted against, so that is what this arm times, with the CPU oracle port
    reported beside it.  At N > 1 every rank runs the reference on its own GPU and keyframe (N
    independent replicas - the reference has no multi-GPU form and no gradient sum), value =
    N frames per slowest rank's step."""
    import reflib
    line = {"impl": "reference", "metric": METRIC, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic"}
    if not reflib.available():
        if rank != 0:
            return None
        cb = cpu_baseline(args.config)
        H, W, _, P, _ = scenes.CONFIGS[args.config]
        line.update(value=cb["value"], ms_per_step=1e3 / cb["value"], cpu_baseline=cb,
                    config={"workload": workload(args.config, H, W, P),
                            "implementation": "CPU oracle port (the reference CUDA build "
                                              "oracle/_ref or a GPU is not available)"},
                    e2e={"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0,
                         "d2h_bytes_per_step": 0})
        return line
    import torch.distributed as dist
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    t = scene_tensors(args.config, rank, dev)
    H, W, P = t["H"], t["W"], t["means3D"].shape[0]
    opa = t["opacities"].reshape(-1).contiguous()
    ref = reflib.RefRasterizer()
    flush = torch.empty(192 << 20, dtype=torch.uint8, device=dev)

    def frame(x, o):
        R, out, radii = ref.forward(x["means3D"], o, x["scales"], x["rotations"],
                                    x["viewmatrix"], x["projmatrix"], H, W, sync=False)
        g = ref.backward(x["dL_dout"], sync=False)
        return R, out, g

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        frame(t, opa)
    torch.cuda.synchronize()

    sampler = ClockSampler(local_rank) if rank == 0 else None
    total_ms = timed_steps(lambda: frame(t, opa), args.steps, flush, barrier)
    clocks = sampler.stop() if sampler else None
    hp = scene_tensors(args.config, rank, dev, pinned=True)
    keys = PARAM_KEYS + ("dL_dout",)
    d_in = []
    for b in range(2):
        d = {k: torch.empty_like(hp[k], device=dev) for k in keys}
        d["viewmatrix"], d["projmatrix"] = t["viewmatrix"], t["projmatrix"]
        d_in.append(d)

    def copy_in(b):
        for k in keys:
            d_in[b][k].copy_(hp[k], non_blocking=True)

    def e2e_frame(b):
        x = d_in[b]
        R, out, g = frame(x, x["opacities"].reshape(-1))
        return [out, g["means3D"], g["opacity"], g["scales"], g["rotations"]]

    barrier()
    ms_e2e, d2h = pipelined_e2e(dev, args.steps, flush, copy_in, e2e_frame,
                                [(7, H, W), (P, 3), (P, 1), (P, 2), (P, 4)])
    barrier()
    e2e_ms = ms_e2e * args.steps
    if world > 1:
        tt = torch.tensor([total_ms, e2e_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total_ms, e2e_ms = float(tt[0]), float(tt[1])
    if rank != 0:
        return None
    ms = total_ms / args.steps
    line.update(
        value=world * 1e3 / ms, ms_per_step=ms, clocks=clocks,
        config={"workload": workload(args.config, H, W, P),
                "implementation": "UNMODIFIED reference CUDA rasterizer: its .cu sources compiled "
                                  "where they lie + oracle/ref_shim.cu (oracle/_ref), not the "
                                  "torch binding of its setup.py",
                "exchange": "none (the reference is single-GPU: N independent replicas)"
                if world > 1 else "none (one GPU)",
                "l2": L2_NOTE, "parallelism": f"{world} independent replicas"},
        kind="reference",
        e2e={"value": world * 1e3 / (e2e_ms / args.steps), "unit": UNIT,
             "h2d_bytes_per_step": sum(hp[k].numel() * 4 for k in keys) * world,
             "d2h_bytes_per_step": d2h * world})
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args.config)
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C2", choices=list(scenes.CONFIGS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-window", action="store_true",
                    help="N>1: skip the C3 / C4 keyframe-window measurement")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "ours" and not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
    if world > 1 and torch.cuda.is_available():
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # keep stdout to the JSON line
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", rank=rank, world_size=world,
                                device_id=torch.device("cuda", local_rank))
    elif world > 1 and rank != 0:
        return
    line = (run_reference if args.impl == "reference" else run_ours)(args, rank, world, local_rank)
    if world > 1 and torch.cuda.is_available():
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
