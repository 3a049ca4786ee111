"""`simple_knn._C`: distCUDA2 with the reference's signature
(submodules/simple-knn/ext.cpp:15-17, spatial.cu:15-26) over the C-ABI sknn_dist2."""
import torch

import splat_b200_native as _n

_lib = _n.load()
_workspaces = {}   # device index -> cached scratch tensor (the reference cudaMallocs per call)


def distCUDA2(points):
    if not points.is_cuda:
        raise RuntimeError("points must be a CUDA tensor")
    P = int(points.size(0))
    dev = points.device
    means = torch.zeros((P,), dtype=torch.float32, device=dev)
    if P == 0:
        return means
    with torch.cuda.device(dev):
        pts = points.contiguous()
        if pts.dtype != torch.float32:
            pts = pts.float()
        need = int(_lib.sknn_workspace_bytes(P))
        ws = _workspaces.get(dev.index)
        if ws is None or ws.numel() < need:
            ws = torch.empty(need, dtype=torch.uint8, device=dev)
            _workspaces[dev.index] = ws
        _n.check(_lib.sknn_dist2(P, pts.data_ptr(), means.data_ptr(), ws.data_ptr(), ws.numel(),
                                 torch.cuda.current_stream(dev).cuda_stream))
    return means
