"""Photometric (range-image) alignment of a query scan to a reference range image - the
photometric half of the reference's `gsaligner.GSAligner`
(submodules/gsaligner/gsaligner/__init__.py:24-146), which Splat-LOAM's tracker calls once per
frame (slam/tracker.py:160-188) - over C-ABI gsa_photo_* (csrc/photometric.cu).

Same constructor keywords, method names, argument meaning and return tuples as the reference
class for this half:

    aligner = PhotometricAligner(image_height=H, image_width=W, photo_omega_depth=5.0, ...)
    aligner.set_reference(depth, cloud, projmat)        # rendered depth of the keyframe
    aligner.set_query(depth, cloud, projmat)            # the new scan
    T, fitness, err = aligner.align_photometric(init)   # (4,4) float32, inlier ratio, residual

The geometric half (MAD-tree ICP, CPU, double precision) is not part of this path; the keyword
arguments that configure it are accepted and ignored, `cloud` likewise.  The whole
Gauss-Newton loop runs on the GPU without a host round trip per iteration.  No CPU path.
"""
import ctypes

import torch

import splat_b200_native as _n

_lib = _n.load()


class PhotometricAligner:
    # GSAligner::GSAligner, aligner.h:99-105: fixed by the reference
    DEPTH_DERIVATIVE_THRESHOLD = 0.5
    MASK_GROW_RADIUS = 1

    def __init__(self, image_height, image_width, photo_omega_depth=5.0,
                 photo_depth_rejection_threshold=0.8, photo_rho_kernel=1.0, photo_min_depth=0.5,
                 photo_max_depth=80.0, photo_iterations=10, **geometric_parameters):
        self.H, self.W = int(image_height), int(image_width)
        self.omega_depth = float(photo_omega_depth)
        self.depth_rejection_threshold = float(photo_depth_rejection_threshold)
        self.rho_kernel = float(photo_rho_kernel)
        self.min_depth, self.max_depth = float(photo_min_depth), float(photo_max_depth)
        self.iterations = int(photo_iterations)
        self._ws = None
        self._K = None
        self._have = set()
        self.last_system = None

    def _workspace(self, device):
        if self._ws is None or self._ws.device != device:
            self._ws = torch.empty(_lib.gsa_photo_workspace_bytes(self.H, self.W),
                                   dtype=torch.uint8, device=device)
            self._have.clear()
        return self._ws

    def _depth(self, depth_image):
        if not depth_image.is_cuda:
            raise RuntimeError("depth_image must be a CUDA tensor")
        if depth_image.dim() != 3 or tuple(depth_image.shape[1:]) != (self.H, self.W):
            raise RuntimeError("depth_image must be a (1xHxW) tensor")
        return depth_image.contiguous().float()

    @staticmethod
    def _camera(projmat):
        # the (4,4) matrix arrives TRANSPOSED (projection_matrix, scene/cameras.py:48-50):
        # K = projmat[:3,:3]^T, aligner.cu:203-206
        p = projmat.detach().float().cpu()
        return float(p[0, 0]), float(p[1, 1]), float(p[2, 0]), float(p[2, 1])

    def set_reference(self, depth_image, cloud, projmat):
        d = self._depth(depth_image)
        ws = self._workspace(d.device)
        self._K = self._camera(projmat)
        with torch.cuda.device(d.device):
            _n.check(_lib.gsa_photo_set_reference(
                ws.data_ptr(), self.H, self.W, d.data_ptr(), self.DEPTH_DERIVATIVE_THRESHOLD,
                self.MASK_GROW_RADIUS, torch.cuda.current_stream(d.device).cuda_stream))
        self._have.add("reference")

    def set_query(self, depth_image, cloud, projmat):
        d = self._depth(depth_image)
        ws = self._workspace(d.device)
        if self._K is None:
            raise RuntimeError("set_reference must be called before set_query")
        fx, fy, cx, cy = self._K        # the query is back-projected with the REFERENCE sensor
        with torch.cuda.device(d.device):                               # (aligner.cu:246)
            _n.check(_lib.gsa_photo_set_query(
                ws.data_ptr(), self.H, self.W, d.data_ptr(), fx, fy, cx, cy, self.min_depth,
                self.max_depth, torch.cuda.current_stream(d.device).cuda_stream))
        self._have.add("query")

    def debug_reference_image(self):
        """Parity-test accessor: [4,H,W] = depth | d/dcol | d/drow | masked of the reference."""
        dev = self._ws.device
        out = torch.empty((4, self.H, self.W), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _n.check(_lib.gsa_photo_debug_reference(self._ws.data_ptr(), self.H, self.W,
                                                    out.data_ptr(),
                                                    torch.cuda.current_stream(dev).cuda_stream))
        return out

    def align_photometric(self, init_guess):
        if self._have != {"reference", "query"}:
            raise RuntimeError("set_reference and set_query must be called before aligning")
        dev = self._ws.device
        fx, fy, cx, cy = self._K
        ig = (ctypes.c_float * 16)(*init_guess.detach().float().cpu().reshape(-1).tolist())
        out, stats = (ctypes.c_float * 16)(), (ctypes.c_float * 2)()
        system = (ctypes.c_double * 30)()
        with torch.cuda.device(dev):
            p = _n.GsaPhotoParams(self.H, self.W, fx, fy, cx, cy, self.min_depth, self.max_depth,
                                  self.omega_depth, self.depth_rejection_threshold,
                                  self.rho_kernel, torch.cuda.current_stream(dev).cuda_stream)
            _n.check(_lib.gsa_photo_align(self._ws.data_ptr(), ctypes.byref(p), ig,
                                          self.iterations, out, stats, system))
        self.last_system = list(system)
        T = torch.tensor(list(out), dtype=torch.float32).reshape(4, 4)
        # alignPhotometric returns (transform, stats.second, stats.first), aligner.cu:341
        return T, float(stats[1]), float(stats[0])
