"""Worker of tests/test_gpu_exchange_mp.py (one process per GPU, launched by torchrun):
the keyframe window's peer-memory exchange against the process-group all-reduce."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "splat-loam_b200"), os.path.join(ROOT, "tests")]
import scenes  # noqa: E402
from gpu_helpers import to_cuda  # noqa: E402
from keyframe_window import KeyframeWindow  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    nkf = int(sys.argv[1]) if len(sys.argv) > 1 else 2 * world + 1
    frames = [to_cuda(scenes.make_scene("C0", keyframe=k)) for k in range(nkf)]
    f0 = frames[0]
    params = (f0["means3D"], f0["opacities"], f0["scales"], f0["rotations"])
    cams = [(f["viewmatrix"], f["projmatrix"], f["H"], f["W"]) for f in frames]
    up = lambda k, allmap: frames[k]["dL_dout"]  # noqa: E731
    ref = KeyframeWindow(exchange="collective").gradients(params, cams, up).clone()
    win = KeyframeWindow(exchange="peer")
    for it in range(3):                     # the exchange is reused: epochs, flags, staging
        got = win.gradients(params, cams, up)
        err = float((got - ref).norm() / ref.norm())
        assert err < 1e-5, (rank, it, err)
        # bit-identical on every rank: compare with rank 0's copy
        mine = got.clone()
        dist.broadcast(mine, src=0)
        assert torch.equal(mine, got), (rank, it)
    # fewer keyframes than ranks: some ranks push zeros
    if world > 1:
        ref1 = KeyframeWindow(exchange="collective").gradients(params, cams[:1], up).clone()
        got1 = win.gradients(params, cams[:1], up)
        assert float((got1 - ref1).norm() / ref1.norm()) < 1e-5
    torch.cuda.synchronize()
    dist.barrier()
    if rank == 0:
        print(f"EXCHANGE_OK world={world} keyframes={nkf} rel_err={err:.2e}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
