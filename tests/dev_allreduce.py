"""Developer tool: where does the multi-GPU step time go?  (torchrun, one rank per GPU)"""
import os, sys, json
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "splat-loam_b200"), os.path.join(ROOT, "tests")]
import bench

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", lr); torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
from diff_surfel_spherical_rasterization import _C
t = bench.scene_tensors("C2", rank, dev)
H, W, P = t["H"], t["W"], t["means3D"].shape[0]
empty = torch.empty(0, device=dev)
gbuf = torch.empty((P, 10), dtype=torch.float32, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

from keyframe_window import PeerExchange
xchg = PeerExchange(P, dev)

def frame_peer():
    R, allmap, radii, geom, binning, img = _C.rasterize_gaussians(
        t["means3D"], t["opacities"], t["scales"], t["rotations"], 1.0, empty,
        t["viewmatrix"], t["projmatrix"], H, W, False, False)
    _C.rasterize_gaussians_backward(t["means3D"], radii, t["scales"], t["rotations"], 1.0, empty,
        t["viewmatrix"], t["projmatrix"], t["dL_dout"], geom, R, binning, img, False, packed=None,
        xchg=xchg.handle.value)
    xchg.allreduce()

def peer_only():
    xchg.push(gbuf); xchg.allreduce()

def frame(ar):
    R, allmap, radii, geom, binning, img = _C.rasterize_gaussians(
        t["means3D"], t["opacities"], t["scales"], t["rotations"], 1.0, empty,
        t["viewmatrix"], t["projmatrix"], H, W, False, False)
    _C.rasterize_gaussians_backward(t["means3D"], radii, t["scales"], t["rotations"], 1.0, empty,
        t["viewmatrix"], t["projmatrix"], t["dL_dout"], geom, R, binning, img, False, packed=gbuf)
    if ar: dist.all_reduce(gbuf)

def timed(fn, n=30, do_flush=True):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    dist.barrier(); torch.cuda.synchronize()
    for a, b in ev:
        if do_flush: flush.zero_()
        a.record(); fn(); b.record(); b.synchronize()
    return sum(a.elapsed_time(b) for a, b in ev) / n

for _ in range(5): frame(True)
res = {"rank": rank,
       "frame_only_ms": timed(lambda: frame(False)),
       "frame_ar_ms": timed(lambda: frame(True)),
       "frame_peer_ms": timed(frame_peer),
       "peer_only_sync_ms": timed(peer_only, do_flush=False),
       "ar_only_sync_ms": timed(lambda: dist.all_reduce(gbuf), do_flush=False)}
# back-to-back all-reduces without host sync
torch.cuda.synchronize(); dist.barrier()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(50): dist.all_reduce(gbuf)
b.record(); b.synchronize()
res["ar_back_to_back_ms"] = a.elapsed_time(b) / 50
torch.cuda.synchronize(); dist.barrier()
a.record()
for _ in range(50): peer_only()
b.record(); b.synchronize()
res["peer_back_to_back_ms"] = a.elapsed_time(b) / 50
# per-kernel split of the unfused exchange
import splat_b200_native as native
lib = native.load()
st = torch.cuda.current_stream().cuda_stream
evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(30)]
torch.cuda.synchronize(); dist.barrier()
for e in evs:
    e[0].record(); native.check(lib.splat_xchg_push(xchg.handle, gbuf.data_ptr(), st))
    e[1].record(); native.check(lib.splat_xchg_reduce(xchg.handle, st))
    e[2].record(); native.check(lib.splat_xchg_wait(xchg.handle, st))
    e[3].record()
torch.cuda.synchronize()
for i, nm in enumerate(("push_us", "reduce_us", "wait_us")):
    res[nm] = 1e3 * sum(e[i].elapsed_time(e[i + 1]) for e in evs[5:]) / len(evs[5:])
out = [None] * world
dist.all_gather_object(out, res)
if rank == 0:
    for r in out: print(json.dumps(r))
dist.destroy_process_group()
