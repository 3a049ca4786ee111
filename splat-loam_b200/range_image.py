"""Range-image front end on the GPU (C-ABI rimg_*, csrc/range_image.cu): what
`Preprocessor.__call__` of the reference (scene/preprocessing.py:24-83) does on the CPU with
numpy and `pyprojections` - intrinsics from the cloud, spherical projection with a z-buffer,
range / normal / valid images - as three kernel launches on a cloud that is already (or is
put) on the device.

    pre = RangeImagePreprocessor(image_height=64, image_width=1024, depth_min=0.5, depth_max=80.0)
    out = pre(cloud)          # cloud: [N,3] float32, numpy or CUDA tensor
    out["K"]                  # (3,3)   float32   -> Camera(K=...)
    out["image_depth"]        # (1,H,W) float32   range image, 0 where no return
    out["image_normal"]       # (3,H,W) float32   -p/|p| of the pixel's point
    out["image_valid"]        # (1,H,W) bool
    out["lut"]                # (H,W)   int64     index of the pixel's point, -1 where empty

The dictionary holds CUDA tensors with the shapes and meaning of the arrays the reference hands
to `Camera(...)` (scene/cameras.py:10-50; `Camera` copies them to the device right away).
PARITY UNPINNED - see csrc/range_image.cu: `pyprojections` is not part of the reference tree.
Only the default normal path (normals towards the sensor, preprocessing.py:112-113) exists;
Open3D's PCA normals and ground segmentation are outside this path.  No CPU path.
"""
import numpy as np
import torch

import splat_b200_native as _n

_lib = _n.load()


class RangeImagePreprocessor:
    def __init__(self, image_height, image_width, depth_min, depth_max, device="cuda"):
        self.H, self.W = int(image_height), int(image_width)
        self.depth_min, self.depth_max = float(depth_min), float(depth_max)
        self.device = torch.device(device)
        self._ws = None

    def __call__(self, cloud, K=None):
        """cloud: [N,3]; K: optional (3,3) camera matrix (default: computed from the cloud, as
        the reference does for every scan)."""
        if isinstance(cloud, np.ndarray):
            cloud = torch.from_numpy(np.ascontiguousarray(cloud, dtype=np.float32)).to(self.device)
        if not cloud.is_cuda:
            raise RuntimeError("cloud must be a numpy array or a CUDA tensor")
        if cloud.dim() != 2 or cloud.shape[1] != 3:
            raise RuntimeError("cloud must have shape (N, 3)")
        cloud = cloud.contiguous().float()
        dev, H, W, P = cloud.device, self.H, self.W, int(cloud.shape[0])
        if P == 0 and K is None:
            raise RuntimeError("an empty cloud has no intrinsics")
        with torch.cuda.device(dev):
            st = torch.cuda.current_stream(dev).cuda_stream
            if self._ws is None or self._ws.device != dev:
                self._ws = torch.empty(_lib.rimg_workspace_bytes(H, W), dtype=torch.uint8, device=dev)
            Kd = torch.empty((3, 3), dtype=torch.float32, device=dev)
            if K is None:
                _n.check(_lib.rimg_spherical_intrinsics(P, cloud.data_ptr(), H, W,
                                                        self._ws.data_ptr(), Kd.data_ptr(), st))
            else:
                Kd.copy_(torch.as_tensor(K, dtype=torch.float32))
            lut = torch.empty((H, W), dtype=torch.int64, device=dev)
            depth = torch.empty((1, H, W), dtype=torch.float32, device=dev)
            normal = torch.empty((3, H, W), dtype=torch.float32, device=dev)
            valid = torch.empty((1, H, W), dtype=torch.bool, device=dev)
            _n.check(_lib.rimg_project(P, cloud.data_ptr() if P else None, H, W,
                                       None if K is None else Kd.data_ptr(), self.depth_min,
                                       self.depth_max, self._ws.data_ptr(), lut.data_ptr(),
                                       depth.data_ptr(), normal.data_ptr(), valid.data_ptr(), st))
        return {"K": Kd, "image_depth": depth, "image_normal": normal, "image_valid": valid,
                "lut": lut}
