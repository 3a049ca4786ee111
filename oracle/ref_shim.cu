// oracle/ref_shim.cu — TEST INFRASTRUCTURE, not product code.
//
// A thin extern "C" shim that lets tests/ and bench.py's reference arm drive the
// UNMODIFIED reference CUDA rasterizer and simple-knn (compiled from the sources
// where they lie under /root/reference, see oracle/Makefile) without torch or
// pybind.  Nothing in the product path (splat-loam_b200/) links or loads this.
//
// It wraps the reference's own C++ entry points:
//   CudaRasterizer::Rasterizer::forward / backward / markVisible
//     (submodules/diff-surfel-spherical-rasterization/cuda_rasterizer/rasterizer.h:7-31)
//   SimpleKNN::knn (submodules/simple-knn/simple_knn.h:15-19)
// and re-derives the reference's internal state pointers with the reference's own
// GeometryState/BinningState/ImageState::fromChunk
//   (cuda_rasterizer/rasterizer_impl.h:18-65) so parity tests can read keys,
// sort order and tile ranges bit for bit.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <string>

#include <cuda_runtime.h>

#include "cuda_rasterizer/rasterizer.h"
#include "cuda_rasterizer/rasterizer_impl.h"
#include "simple_knn.h"

namespace {

struct Chunk {
  char *ptr = nullptr;
  size_t cap = 0;
  size_t used = 0;
  char *grow(size_t n) {
    if (n > cap) {
      if (ptr) cudaFree(ptr);
      cap = n + (n >> 2) + 256;
      if (cudaMalloc(&ptr, cap) != cudaSuccess) {
        ptr = nullptr;
        cap = 0;
      }
    }
    used = n;
    return ptr;
  }
  ~Chunk() {
    if (ptr) cudaFree(ptr);
  }
};

struct RefHandle {
  Chunk geom, binning, img;
  int P = 0, W = 0, H = 0, R = 0;
};

}  // namespace

extern "C" {

void *ref_dssr_create() { return new RefHandle(); }

void ref_dssr_destroy(void *h) { delete static_cast<RefHandle *>(h); }

// Returns num_rendered (>= 0) or -1 on CUDA error.
int ref_dssr_forward(void *h, int P, int W, int H, const float *means3D,
                     const float *opacities, const float *scales,
                     float scale_modifier, const float *rotations,
                     const float *transMat_precomp, const float *viewmatrix,
                     const float *projmatrix, int prefiltered, float *out_others,
                     int *radii) {
  RefHandle *s = static_cast<RefHandle *>(h);
  s->P = P;
  s->W = W;
  s->H = H;
  s->R = 0;
  // mirrors RasterizeGaussiansCUDA (rasterize_points.cu:59-61): outputs start at 0
  cudaMemset(out_others, 0, sizeof(float) * 7 * (size_t)W * H);
  cudaMemset(radii, 0, sizeof(int) * (size_t)P);
  if (P == 0) return 0;
  std::function<char *(size_t)> g = [s](size_t n) { return s->geom.grow(n); };
  std::function<char *(size_t)> b = [s](size_t n) { return s->binning.grow(n); };
  std::function<char *(size_t)> i = [s](size_t n) { return s->img.grow(n); };
  int R = CudaRasterizer::Rasterizer::forward(
      g, b, i, P, W, H, means3D, opacities, scales, scale_modifier, rotations,
      transMat_precomp, viewmatrix, projmatrix, prefiltered != 0, out_others,
      radii, false);
  s->R = R;
  if (cudaDeviceSynchronize() != cudaSuccess) return -1;
  return R;
}

int ref_dssr_backward(void *h, const float *means3D, const float *scales,
                      float scale_modifier, const float *rotations,
                      const float *transMat_precomp, const float *viewmatrix,
                      const float *projmatrix, const int *radii,
                      const float *dL_dout, float *dL_dmean2D, float *dL_dnormal,
                      float *dL_dopacity, float *dL_dmean3D, float *dL_dtransMat,
                      float *dL_dscale, float *dL_drot) {
  RefHandle *s = static_cast<RefHandle *>(h);
  const size_t P = s->P;
  // mirrors RasterizeGaussiansBackwardCUDA (rasterize_points.cu:116-122)
  cudaMemset(dL_dmean3D, 0, sizeof(float) * 3 * P);
  cudaMemset(dL_dmean2D, 0, sizeof(float) * 3 * P);
  cudaMemset(dL_dnormal, 0, sizeof(float) * 3 * P);
  cudaMemset(dL_dopacity, 0, sizeof(float) * 1 * P);
  cudaMemset(dL_dtransMat, 0, sizeof(float) * 12 * P);
  cudaMemset(dL_dscale, 0, sizeof(float) * 2 * P);
  cudaMemset(dL_drot, 0, sizeof(float) * 4 * P);
  if (P == 0) return 0;
  CudaRasterizer::Rasterizer::backward(
      s->P, s->R, s->W, s->H, means3D, scales, scale_modifier, rotations,
      transMat_precomp, viewmatrix, projmatrix, radii, s->geom.ptr,
      s->binning.ptr, s->img.ptr, dL_dout, dL_dmean2D, dL_dnormal, dL_dopacity,
      dL_dmean3D, dL_dtransMat, dL_dscale, dL_drot, false);
  return cudaDeviceSynchronize() == cudaSuccess ? 0 : -1;
}

// Launch-only variants (no trailing synchronise, no memsets of their own beyond
// what the reference binding does) for timing with CUDA events on stream 0.
int ref_dssr_num_rendered(void *h) { return static_cast<RefHandle *>(h)->R; }

// Internal state of the last forward, by name.  Returns 0 and fills (*ptr,
// *count) on success.  Element types: see rasterizer_impl.h:18-65.
int ref_dssr_state(void *h, const char *name, void **ptr, size_t *count) {
  RefHandle *s = static_cast<RefHandle *>(h);
  if (!s->geom.ptr) return -1;
  using namespace CudaRasterizer;
  char *gp = s->geom.ptr;
  GeometryState gs = GeometryState::fromChunk(gp, s->P);
  char *ip = s->img.ptr;
  ImageState is = ImageState::fromChunk(ip, (size_t)s->W * s->H);
  BinningState bs{};
  if (s->binning.ptr) {
    char *bp = s->binning.ptr;
    bs = BinningState::fromChunk(bp, s->R);
  }
  const std::string n(name);
  const size_t P = s->P, R = s->R, HW = (size_t)s->W * s->H;
  const size_t tiles = (size_t)((s->W + 15) / 16) * ((s->H + 15) / 16);
  if (n == "depths") { *ptr = gs.depths; *count = P; }
  else if (n == "means2D") { *ptr = gs.means2D; *count = 2 * P; }
  else if (n == "transMat") { *ptr = gs.transMat; *count = 12 * P; }
  else if (n == "normal_opacity") { *ptr = gs.normal_opacity; *count = 4 * P; }
  else if (n == "tiles_touched") { *ptr = gs.tiles_touched; *count = P; }
  else if (n == "point_offsets") { *ptr = gs.point_offsets; *count = P; }
  else if (n == "keys_unsorted") { *ptr = bs.point_list_keys_unsorted; *count = R; }
  else if (n == "keys") { *ptr = bs.point_list_keys; *count = R; }
  else if (n == "point_list_unsorted") { *ptr = bs.point_list_unsorted; *count = R; }
  else if (n == "point_list") { *ptr = bs.point_list; *count = R; }
  else if (n == "ranges") { *ptr = is.ranges; *count = 2 * tiles; }
  else if (n == "final_T") { *ptr = is.accum_alpha; *count = 3 * HW; }
  else if (n == "n_contrib") { *ptr = is.n_contrib; *count = 2 * HW; }
  else return -2;
  return 0;
}

int ref_dssr_mark_visible(int P, float *means3D, float *viewmatrix,
                          float *projmatrix, unsigned char *present) {
  cudaMemset(present, 0, (size_t)P);
  if (P == 0) return 0;
  CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix,
                                          reinterpret_cast<bool *>(present));
  return cudaDeviceSynchronize() == cudaSuccess ? 0 : -1;
}

// distCUDA2 (simple-knn/spatial.cu:15-26): means start at 0, then SimpleKNN::knn.
int ref_sknn_dist2(int P, const float *points, float *mean_dists) {
  cudaMemset(mean_dists, 0, sizeof(float) * (size_t)P);
  if (P == 0) return 0;
  SimpleKNN::knn(P, (float3 *)points, mean_dists);
  return cudaDeviceSynchronize() == cudaSuccess ? 0 : -1;
}

// Timing variants: enqueue only, the caller brackets with events on stream 0
// (legacy default stream, where the reference launches everything).
int ref_dssr_forward_nosync(void *h, int P, int W, int H, const float *means3D,
                            const float *opacities, const float *scales,
                            float scale_modifier, const float *rotations,
                            const float *transMat_precomp,
                            const float *viewmatrix, const float *projmatrix,
                            int prefiltered, float *out_others, int *radii) {
  RefHandle *s = static_cast<RefHandle *>(h);
  s->P = P; s->W = W; s->H = H; s->R = 0;
  cudaMemsetAsync(out_others, 0, sizeof(float) * 7 * (size_t)W * H, 0);
  cudaMemsetAsync(radii, 0, sizeof(int) * (size_t)P, 0);
  if (P == 0) return 0;
  std::function<char *(size_t)> g = [s](size_t n) { return s->geom.grow(n); };
  std::function<char *(size_t)> b = [s](size_t n) { return s->binning.grow(n); };
  std::function<char *(size_t)> i = [s](size_t n) { return s->img.grow(n); };
  s->R = CudaRasterizer::Rasterizer::forward(
      g, b, i, P, W, H, means3D, opacities, scales, scale_modifier, rotations,
      transMat_precomp, viewmatrix, projmatrix, prefiltered != 0, out_others,
      radii, false);
  return s->R;
}

int ref_dssr_backward_nosync(void *h, const float *means3D, const float *scales,
                             float scale_modifier, const float *rotations,
                             const float *transMat_precomp,
                             const float *viewmatrix, const float *projmatrix,
                             const int *radii, const float *dL_dout,
                             float *dL_dmean2D, float *dL_dnormal,
                             float *dL_dopacity, float *dL_dmean3D,
                             float *dL_dtransMat, float *dL_dscale,
                             float *dL_drot) {
  RefHandle *s = static_cast<RefHandle *>(h);
  const size_t P = s->P;
  cudaMemsetAsync(dL_dmean3D, 0, sizeof(float) * 3 * P, 0);
  cudaMemsetAsync(dL_dmean2D, 0, sizeof(float) * 3 * P, 0);
  cudaMemsetAsync(dL_dnormal, 0, sizeof(float) * 3 * P, 0);
  cudaMemsetAsync(dL_dopacity, 0, sizeof(float) * 1 * P, 0);
  cudaMemsetAsync(dL_dtransMat, 0, sizeof(float) * 12 * P, 0);
  cudaMemsetAsync(dL_dscale, 0, sizeof(float) * 2 * P, 0);
  cudaMemsetAsync(dL_drot, 0, sizeof(float) * 4 * P, 0);
  if (P == 0) return 0;
  CudaRasterizer::Rasterizer::backward(
      s->P, s->R, s->W, s->H, means3D, scales, scale_modifier, rotations,
      transMat_precomp, viewmatrix, projmatrix, radii, s->geom.ptr,
      s->binning.ptr, s->img.ptr, dL_dout, dL_dmean2D, dL_dnormal, dL_dopacity,
      dL_dmean3D, dL_dtransMat, dL_dscale, dL_drot, false);
  return 0;
}

int ref_sknn_dist2_nosync(int P, const float *points, float *mean_dists) {
  cudaMemsetAsync(mean_dists, 0, sizeof(float) * (size_t)P, 0);
  if (P == 0) return 0;
  SimpleKNN::knn(P, (float3 *)points, mean_dists);
  return 0;
}

}  // extern "C"
