// range_image.cu — the range-image front end: a LiDAR cloud becomes the spherical range /
// normal / valid images a Splat-LOAM Camera holds.
//
// Replaces Preprocessor.__call__ (scene/preprocessing.py:24-83) up to the Camera constructor:
//   pyp.calculate_spherical_intrinsics(cloud.T, H, W)     :41-43
//   pyp.Camera(...).project(cloud.T) -> lut, valid         :44-53
//   range_image = ranges[lut], 0 where invalid             :56-58
//   normals = -cloud / |cloud| for depth_min < r <= depth_max, normals[lut], 0 where invalid
//                                                           :59-64, :112-113 (default path)
// The reference does this on the CPU with numpy and `pyprojections` 0.0.3, a PyPI dependency
// (pixi.toml:39) that is NOT vendored under the reference tree and is not in this image, so its
// rounding, tie and intrinsics rules cannot be read: PARITY UNPINNED.  What the tree does hold
// fixes most of the convention (DESIGN.md section 7.6 has the measurements):
//   * submodules/gsaligner/tests/data/K_ref.npy, a camera matrix this step produced: fx, fy < 0,
//     fx = -(W-1)/hfov with hfov the cloud's own azimuth span, cx = W/2 - fx * mid-azimuth (the
//     extreme azimuths land on 0.5 and W - 0.5); the same form is used in y;
//   * submodules/gsaligner/tests/data/depth_query.npy, a range image this step produced: row 0
//     and column 0 are exactly empty and each of the other 63 rows holds ~1000 returns.  With
//     the extreme beams on 0.5 and H - 0.5 and the beams between them about one pixel apart,
//     floor() would fill row 0, and rounding to nearest would sit on the knife edge for every
//     beam (some rows hit twice, some left empty); pixel = floor(coordinate) + 1 gives exactly
//     this picture (0.5 -> 1, ..., H - 0.5 -> H, dropped).  It is also the inverse of the
//     reference's own back-projection, which places pixel (c, r) at (c - 0.5, r - 0.5)
//     (utils/graphic_utils.py:45-49);
//   * preprocessing.py:56-62 itself: range = numpy's float32 norm ((x*x + y*y) + z*z, no fused
//     multiply-add), a point takes part when depth_min < range <= depth_max.
// So:
//   * azimuth = atan2(y, x), elevation = atan2(z, hypot(x, y)); hfov / vfov = the spans of the
//     cloud's azimuths / elevations; fx = -(W-1)/hfov, fy = -(H-1)/vfov,
//     cx = W/2 - fx * (az_min + az_max)/2, cy = H/2 - fy * (el_min + el_max)/2;
//   * column = floor(fx * azimuth + cx) + 1, row = floor(fy * elevation + cy) + 1;
//   * the point with the smallest range wins a pixel, ties go to the smaller index (unknowable
//     without the package; gsaligner's own z-buffer keeps the nearest return too).
// Three launches: (az, el) extrema, z-buffer (atomicMin of range bits | index), resolve.
#include <cfloat>

#include "../../include/splat_b200.h"
#include "common.cuh"

namespace {

struct RimgState {                 // head of the workspace
  float lim[4];                    // az_min, az_max, el_min, el_max
  float K[4];                      // fx, fy, cx, cy
  unsigned int pad[8];
};

__device__ __forceinline__ void angles(float x, float y, float z, float &az, float &el) {
  az = atan2f(y, x);
  el = atan2f(z, sqrtf(fmaf(y, y, x * x)));
}

// float atomics by CAS on the bit pattern are not needed: the order of non-negative floats is
// the order of their bits, and of negative floats the reverse - map to an ordered unsigned key
__device__ __forceinline__ unsigned int ordered(float f) {
  const unsigned int u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float unordered(unsigned int k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

__global__ void __launch_bounds__(256)
extrema_kernel(int P, const float *__restrict__ cloud, unsigned int *keys /* [4] */) {
  __shared__ unsigned int s[4][8];
  unsigned int lo_a = 0xffffffffu, hi_a = 0u, lo_e = 0xffffffffu, hi_e = 0u;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
    const float x = cloud[3 * i], y = cloud[3 * i + 1], z = cloud[3 * i + 2];
    float az, el;
    angles(x, y, z, az, el);
    if (az == az && el == el) {    // (NaN points take no part)
      const unsigned int ka = ordered(az), ke = ordered(el);
      lo_a = min(lo_a, ka); hi_a = max(hi_a, ka);
      lo_e = min(lo_e, ke); hi_e = max(hi_e, ke);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo_a = min(lo_a, __shfl_xor_sync(0xffffffffu, lo_a, o));
    hi_a = max(hi_a, __shfl_xor_sync(0xffffffffu, hi_a, o));
    lo_e = min(lo_e, __shfl_xor_sync(0xffffffffu, lo_e, o));
    hi_e = max(hi_e, __shfl_xor_sync(0xffffffffu, hi_e, o));
  }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { s[0][wid] = lo_a; s[1][wid] = hi_a; s[2][wid] = lo_e; s[3][wid] = hi_e; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; w++) {
      lo_a = min(lo_a, s[0][w]); hi_a = max(hi_a, s[1][w]);
      lo_e = min(lo_e, s[2][w]); hi_e = max(hi_e, s[3][w]);
    }
    atomicMin(&keys[0], lo_a); atomicMax(&keys[1], hi_a);
    atomicMin(&keys[2], lo_e); atomicMax(&keys[3], hi_e);
  }
}

__global__ void intrinsics_kernel(const unsigned int *keys, int H, int W, RimgState *st,
                                  float *K_out /* [9] row-major, optional */) {
  if (threadIdx.x != 0) return;
  const float a0 = unordered(keys[0]), a1 = unordered(keys[1]);
  const float e0 = unordered(keys[2]), e1 = unordered(keys[3]);
  const float hfov = a1 - a0, vfov = e1 - e0;
  const float fx = -(float)(W - 1) / hfov, fy = -(float)(H - 1) / vfov;
  const float cx = 0.5f * (float)W - fx * (0.5f * (a0 + a1));
  const float cy = 0.5f * (float)H - fy * (0.5f * (e0 + e1));
  st->lim[0] = a0; st->lim[1] = a1; st->lim[2] = e0; st->lim[3] = e1;
  st->K[0] = fx; st->K[1] = fy; st->K[2] = cx; st->K[3] = cy;
  if (K_out) {
    K_out[0] = fx;  K_out[1] = 0.f; K_out[2] = cx;
    K_out[3] = 0.f; K_out[4] = fy;  K_out[5] = cy;
    K_out[6] = 0.f; K_out[7] = 0.f; K_out[8] = 1.f;
  }
}

__global__ void __launch_bounds__(256)
zbuffer_kernel(int P, const float *__restrict__ cloud, const float *__restrict__ K /* fx fy cx cy */,
               int H, int W, float depth_min, float depth_max, unsigned long long *zbuf) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const float x = cloud[3 * i], y = cloud[3 * i + 1], z = cloud[3 * i + 2];
  // numpy's float32 norm, preprocessing.py:56 (no contraction into fused multiply-adds)
  const float r = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z)));
  if (!(r > depth_min) || !(r <= depth_max)) return;           // preprocessing.py:60-61
  float az, el;
  angles(x, y, z, az, el);
  const float u = fmaf(K[0], az, K[2]), v = fmaf(K[1], el, K[3]);
  if (!(fabsf(u) < 1e9f) || !(fabsf(v) < 1e9f)) return;        // NaN / overflow of the casts
  const int col = (int)floorf(u) + 1, row = (int)floorf(v) + 1;
  if ((unsigned)col >= (unsigned)W || (unsigned)row >= (unsigned)H) return;
  atomicMin(&zbuf[row * W + col], ((unsigned long long)__float_as_uint(r) << 32) | (unsigned)i);
}

__global__ void __launch_bounds__(256)
resolve_kernel(int HW, const float *__restrict__ cloud, const unsigned long long *__restrict__ zbuf,
               long long *lut, float *range_image, float *normal_image, unsigned char *valid) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= HW) return;
  const unsigned long long key = zbuf[p];
  long long idx = -1;
  float r = 0.f, nx = 0.f, ny = 0.f, nz = 0.f;
  if (key != ~0ull) {
    idx = (long long)(unsigned)(key & 0xffffffffull);
    r = __uint_as_float((unsigned)(key >> 32));
    const float x = cloud[3 * idx], y = cloud[3 * idx + 1], z = cloud[3 * idx + 2];
    nx = -x / r; ny = -y / r; nz = -z / r;          // preprocessing.py:113
  }
  if (lut) lut[p] = idx;
  range_image[p] = r;
  normal_image[p] = nx; normal_image[HW + p] = ny; normal_image[2 * HW + p] = nz;
  valid[p] = idx >= 0;
}

}  // namespace

extern "C" {

size_t rimg_workspace_bytes(int H, int W) {
  if (H <= 0 || W <= 0) return 0;
  return align_up(sizeof(RimgState) + 64, 256) + (size_t)H * W * 8 + 512;
}

int rimg_spherical_intrinsics(int P, const float *cloud, int H, int W, void *workspace,
                              float *K_out, void *stream) {
  if (P <= 0 || !cloud || !workspace || H <= 0 || W <= 0)
    return splat_fail("rimg_spherical_intrinsics: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  char *b = align_ptr((const char *)workspace);
  RimgState *state = (RimgState *)b;
  unsigned int *keys = (unsigned int *)(b + sizeof(RimgState));
  static const unsigned int init[4] = {0xffffffffu, 0u, 0xffffffffu, 0u};
  cudaError_t e = cudaMemcpyAsync(keys, init, sizeof init, cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return splat_fail("rimg_spherical_intrinsics: %s", cudaGetErrorString(e));
  int blocks = (P + 255) / 256;
  const int cap = splat_sm_count() * 8;
  if (blocks > cap) blocks = cap;
  extrema_kernel<<<blocks, 256, 0, LS(st)>>>(P, cloud, keys);
  intrinsics_kernel<<<1, 32, 0, LS(st)>>>(keys, H, W, state, K_out);
  e = cudaGetLastError();
  return e == cudaSuccess ? 0 : splat_fail("rimg_spherical_intrinsics: %s", cudaGetErrorString(e));
}

int rimg_project(int P, const float *cloud, int H, int W, const float *K, float depth_min,
                 float depth_max, void *workspace, long long *lut, float *range_image,
                 float *normal_image, unsigned char *valid, void *stream) {
  if (P < 0 || H <= 0 || W <= 0 || !workspace || !range_image || !normal_image || !valid ||
      (P > 0 && !cloud))
    return splat_fail("rimg_project: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  char *b = align_ptr((const char *)workspace);
  RimgState *state = (RimgState *)b;
  unsigned long long *zbuf = (unsigned long long *)(b + align_up(sizeof(RimgState) + 64, 256));
  const int HW = H * W;
  cudaError_t e = cudaMemsetAsync(zbuf, 0xff, (size_t)HW * 8, st);
  if (e != cudaSuccess) return splat_fail("rimg_project: %s", cudaGetErrorString(e));
  const float *k4 = state->K;      // the intrinsics rimg_spherical_intrinsics left in the workspace
  if (K) {                         // or the caller's row-major 3x3 (device)
    float *dst = state->K;
    e = cudaMemcpy2DAsync(dst, sizeof(float), K, 4 * sizeof(float), sizeof(float), 2,
                          cudaMemcpyDeviceToDevice, st);              // K[0], K[4] -> fx, fy
    if (e == cudaSuccess)
      e = cudaMemcpy2DAsync(dst + 2, sizeof(float), K + 2, 3 * sizeof(float), sizeof(float), 2,
                            cudaMemcpyDeviceToDevice, st);            // K[2], K[5] -> cx, cy
    if (e != cudaSuccess) return splat_fail("rimg_project: %s", cudaGetErrorString(e));
  }
  if (P > 0)
    zbuffer_kernel<<<(P + 255) / 256, 256, 0, LS(st)>>>(P, cloud, k4, H, W, depth_min, depth_max, zbuf);
  resolve_kernel<<<(HW + 255) / 256, 256, 0, LS(st)>>>(HW, cloud, zbuf, lut, range_image,
                                                       normal_image, valid);
  e = cudaGetLastError();
  return e == cudaSuccess ? 0 : splat_fail("rimg_project: %s", cudaGetErrorString(e));
}

}  // extern "C"
