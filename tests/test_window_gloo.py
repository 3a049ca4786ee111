"""CPU tests of the N>1 path: keyframe sharding + gradient all-reduce on a 2-rank gloo group.
The frame function is a deterministic CPU stand-in (the CUDA rasterizer needs a GPU); what is
tested is the host logic: partition, accumulation buffer, collective, equality with the
single-process sum."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _fake_frame(params, camera, dL_fn):
    means3D, opacities, scales, rotations = params
    view, proj, H, W = camera
    allmap = torch.zeros(7, H, W)
    dL = dL_fn(allmap)
    w = float(view[3, 0]) + 1.0 + float(dL.sum())
    g = torch.cat([means3D * w, opacities * w, scales * w, rotations * w], dim=1)
    return g, allmap


def _scene(P=257, K=5):
    g = torch.Generator().manual_seed(0)
    params = (torch.randn(P, 3, generator=g), torch.rand(P, 1, generator=g),
              torch.rand(P, 2, generator=g), torch.randn(P, 4, generator=g))
    cams = []
    for k in range(K):
        v = torch.eye(4)
        v[3, 0] = 0.5 * k
        cams.append((v, torch.eye(4), 4, 8))
    return params, cams


def _dL(k, allmap):
    return torch.full_like(allmap, 0.001 * (k + 1))


def _worker(rank, world, port, out):
    import sys
    sys.path.insert(0, os.path.join(ROOT, "splat-loam_b200"))
    from keyframe_window import KeyframeWindow, shard_keyframes
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    params, cams = _scene()
    win = KeyframeWindow(frame_fn=_fake_frame)
    local, mine = win.accumulate(params, cams, _dL)
    assert mine == shard_keyframes(len(cams), rank, world)
    total = win.gradients(params, cams, _dL)
    torch.save((mine, total), os.path.join(out, f"r{rank}.pt"))
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_shard_keyframes():
    import sys
    sys.path.insert(0, os.path.join(ROOT, "splat-loam_b200"))
    from keyframe_window import shard_keyframes
    for K in (0, 1, 5, 8, 32):
        for world in (1, 2, 3, 8):
            parts = [shard_keyframes(K, r, world) for r in range(world)]
            assert sorted(sum(parts, [])) == list(range(K))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


@pytest.mark.timeout(120)
def test_two_rank_window_equals_single_process_sum(tmp_path):
    import sys
    sys.path.insert(0, os.path.join(ROOT, "splat-loam_b200"))
    from keyframe_window import KeyframeWindow
    params, cams = _scene()
    want = KeyframeWindow(frame_fn=_fake_frame).gradients(params, cams, _dL).clone()
    single = sum(_fake_frame(params, cams[k], lambda m, k=k: _dL(k, m))[0] for k in range(len(cams)))
    assert torch.allclose(want, single, rtol=1e-6, atol=1e-6)
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    got = [torch.load(os.path.join(tmp_path, f"r{r}.pt")) for r in range(2)]
    assert got[0][0] == [0, 2, 4] and got[1][0] == [1, 3]
    assert torch.equal(got[0][1], got[1][1])                      # identical on every rank
    assert torch.allclose(got[0][1], want, rtol=1e-5, atol=1e-6)  # == sum over all keyframes
    m, o, s, r = KeyframeWindow.split(got[0][1])
    assert m.shape[1] == 3 and o.shape[1] == 1 and s.shape[1] == 2 and r.shape[1] == 4
