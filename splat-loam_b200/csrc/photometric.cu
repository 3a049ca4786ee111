// photometric.cu — photometric (range-image) registration of a query scan against a
// reference range image: the tracking stage that runs downstream of the rasterizer.
//
// Replaces, for the configuration the reference uses it in (spherical camera, depth channel
// only, identity sensor offset - GSAligner::GSAligner, submodules/gsaligner/gsaligner/
// aligner.h:94-121; the binding fills nothing but depth, aligner_utils.cu:3-16):
//   ImageSet::growMaskDevice / updateDerivativesDevice   tools/image_set_impl.cu:254-376
//   ImageSet::toCloudDevice                              tools/image_set_impl.cu:20-120
//   PhotometricFac::computeProjections (project_kernel)  registration/photometric_cuda.cu:8-97
//   PhotometricFac::linearize (linearize_kernel,
//     sum_reduce, host-side final sum, LDLT, pose update) registration/photometric_cuda.cu:99-395
//   PhotometricFac::compute                              registration/photometric.cpp:55-77
// driven from slam/tracker.py:160-188 through GSAligner::alignPhotometric (aligner.cu:329-342).
//
// The reference runs, per iteration: a workspace fill, two projection kernels (z-buffer, then
// a 160-byte workspace entry per pixel), a fill of a 240-byte linear-system entry per pixel, the
// linearisation, a block reduction of those entries, a cudaMalloc / cudaFree pair, a D2H copy
// and the final sum + 6x6 solve on the host - with a cudaDeviceSynchronize after every launch.
// Here an iteration is three stream-ordered operations and no host round trip:
//   1. clear the 8-byte z-buffer;
//   2. project_kernel: every query point -> atomicMin of (depth bits << 32 | point index);
//   3. linearize_kernel: every pixel recomputes its winner's projection (the same device
//      function, so the same floats), interpolates the reference image, forms the 1x6 Jacobian
//      row, and the 21 + 6 + 3 sums are reduced in double by warp shuffles -> shared memory ->
//      one atomic per CTA; the LAST CTA solves the 6x6 system and updates the pose in device
//      memory, where the next iteration's kernels read it.
// With only the depth channel populated the reference's 5-vector error has one non-zero
// component and its 5x6 Jacobian one non-zero row; the zero rows contribute exact zeros to
// every product, so they are not computed.
#include <cfloat>
#include <cmath>
#include <cstring>

#include "../../include/splat_b200.h"
#include "common.cuh"

namespace {

struct PhotoDevice {        // the part of the workspace the kernels read and write
  float R[9], t[3];         // ref_T_query, row-major rotation + translation
  double accum[30];         // upper H (21) | b (6) | chi | good | inliers
  double last[30];          // the sums of the last linearisation (kept for the caller)
  float stats[2];           // residual, inlier ratio of the last linearisation
  unsigned int ticket;
  unsigned int pad;
};

struct PhotoView {
  PhotoDevice *dev;
  float *ref_depth, *ref_dc, *ref_dr;     // reference image: depth, d/dcol, d/drow
  unsigned char *ref_mask, *ref_mask0;    // masked (after / before growing)
  float4 *cloud;                          // query points (sensor frame), zeros where invalid
  unsigned long long *zbuf;
};

__host__ PhotoView make_photo_view(void *ws, int H, int W) {
  const size_t n = (size_t)H * W;
  char *b = align_ptr((const char *)ws);
  PhotoView v;
  size_t o = 0;
  v.dev = (PhotoDevice *)(b + o);        o = align_up(o + sizeof(PhotoDevice), 256);
  v.ref_depth = (float *)(b + o);        o = align_up(o + n * 4, 256);
  v.ref_dc = (float *)(b + o);           o = align_up(o + n * 4, 256);
  v.ref_dr = (float *)(b + o);           o = align_up(o + n * 4, 256);
  v.ref_mask = (unsigned char *)(b + o); o = align_up(o + n, 256);
  v.ref_mask0 = (unsigned char *)(b + o); o = align_up(o + n, 256);
  v.cloud = (float4 *)(b + o);           o = align_up(o + n * 16, 256);
  v.zbuf = (unsigned long long *)(b + o);
  return v;
}

size_t photo_bytes(int H, int W) {
  const size_t n = (size_t)H * W;
  return align_up(sizeof(PhotoDevice), 256) + 3 * align_up(n * 4, 256) + 2 * align_up(n, 256) +
         align_up(n * 16, 256) + align_up(n * 8, 256) + 256;
}

// aligner_utils.cu:3-16: depth > 1e-5 unmasks a pixel, the depth is stored either way
__global__ void ref_init_kernel(int n, const float *__restrict__ depth, PhotoView v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float d = depth[i];
  v.ref_depth[i] = d;
  v.ref_mask0[i] = !(d > 1e-5);
}

// __grow_mask_kernel (image_set_impl.cu:318-334): every masked pixel masks the pixels at the
// FLATTENED offsets r * cols + c of the disc r^2 + c^2 <= radius^2 (no row wrap check), as
// long as the target index is inside the image.  Written as a gather (the disc is symmetric).
__global__ void ref_grow_kernel(int H, int W, int radius, PhotoView v) {
  const int n = H * W;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  bool m = v.ref_mask0[i] != 0;
  for (int r = -radius; r <= radius && !m; r++)
    for (int c = -radius; c <= radius; c++) {
      if (r * r + c * c > radius * radius) continue;
      const int src = i - (r * W + c);
      if (src >= 0 && src < n && v.ref_mask0[src]) { m = true; break; }
    }
  v.ref_mask[i] = m;
}

// updateDerivatives_kernel (image_set_impl.cu:269-300) for the depth channel, policy Clamp
__global__ void ref_deriv_kernel(int H, int W, float thr2, PhotoView v) {
  const int col = blockIdx.x * blockDim.x + threadIdx.x, row = blockIdx.y * blockDim.y + threadIdx.y;
  if (row >= H || col >= W) return;
  const int i = row * W + col;
  float dc = 0.f, dr = 0.f;
  if (row >= 1 && row < H - 1 && col >= 1 && col < W - 1) {
    dr = 0.5f * v.ref_depth[i + W] - 0.5f * v.ref_depth[i - W];
    dc = 0.5f * v.ref_depth[i + 1] - 0.5f * v.ref_depth[i - 1];
    if (!v.ref_mask[i]) {
      const float n = dc * dc + dr * dr;
      if (!(n < thr2)) {
        const float s = sqrtf(thr2 / n);
        dc *= s;
        dr *= s;
      }
    }
  }
  v.ref_dc[i] = dc;
  v.ref_dr[i] = dr;
}

// toCloud_kernel (image_set_impl.cu:20-77), spherical: masked or out-of-range pixels stay at
// the zero point the reference fills the cloud with (and are projected like any other point,
// photometric_cuda.cu:19-27 does not look at the status)
__global__ void query_cloud_kernel(int H, int W, const float *__restrict__ depth, float ifx,
                                   float ify, float cx, float cy, float min_depth,
                                   float max_depth, PhotoView v) {
  const int col = blockIdx.x * blockDim.x + threadIdx.x, row = blockIdx.y * blockDim.y + threadIdx.y;
  if (row >= H || col >= W) return;
  const float w = depth[row * W + col];
  float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
  if (w > 1e-5 && !(w < min_depth) && !(w > max_depth)) {
    const float az = ifx * ((float)col - cx), el = ify * ((float)row - cy);
    const float s0 = sinf(az), c0 = cosf(az), s1 = sinf(el), c1 = cosf(el);
    p = make_float4(c0 * c1 * w, s0 * c1 * w, s1 * w, 1.0f);
  }
  v.cloud[row * W + col] = p;
}

// parity-test accessor: depth | d/dcol | d/drow | masked of the prepared reference image
__global__ void ref_export_kernel(int n, PhotoView v, float *out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[i] = v.ref_depth[i];
  out[n + i] = v.ref_dc[i];
  out[2 * n + i] = v.ref_dr[i];
  out[3 * n + i] = v.ref_mask[i] ? 1.0f : 0.0f;
}

struct Projected {
  float px, py, pz;      // the query point
  float tx, ty, tz;      // transformed into the reference frame
  float depth, u, v;     // range and image point (column, row)
  int irow, icol;
  bool good;
};

// SX * point and utils_photo.cuh:54-82 (spherical) + the rounding of photometric_cuda.cu:36-38.
// One function for both kernels: the z-buffer winner is re-derived bit for bit.
__device__ __forceinline__ Projected project_point(const PhotoDevice *__restrict__ d, float4 p,
                                                   float fx, float fy, float cx, float cy,
                                                   float min_depth, float max_depth) {
  Projected r;
  r.px = p.x; r.py = p.y; r.pz = p.z;
  r.tx = fmaf(d->R[2], p.z, fmaf(d->R[1], p.y, d->R[0] * p.x)) + d->t[0];
  r.ty = fmaf(d->R[5], p.z, fmaf(d->R[4], p.y, d->R[3] * p.x)) + d->t[1];
  r.tz = fmaf(d->R[8], p.z, fmaf(d->R[7], p.y, d->R[6] * p.x)) + d->t[2];
  const float rxy2 = fmaf(r.ty, r.ty, r.tx * r.tx);
  r.depth = sqrtf(fmaf(r.tz, r.tz, rxy2));
  r.good = !(r.depth < min_depth) && !(r.depth > max_depth);
  r.u = fx * atan2f(r.ty, r.tx) + cx;
  r.v = fy * atan2f(r.tz, sqrtf(rxy2)) + cy;
  r.irow = (int)(r.v + (r.v >= 0 ? 0.5f : -0.5f));
  r.icol = (int)(r.u + (r.u >= 0 ? 0.5f : -0.5f));
  return r;
}

__global__ void __launch_bounds__(256)
project_kernel(int H, int W, float fx, float fy, float cx, float cy, float min_depth,
               float max_depth, PhotoView v) {
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= H * W) return;
  const Projected r = project_point(v.dev, v.cloud[tid], fx, fy, cx, cy, min_depth, max_depth);
  if (!r.good) return;
  if ((unsigned)r.irow >= (unsigned)H || (unsigned)r.icol >= (unsigned)W) return;
  const int pix = r.irow * W + r.icol;
  if (v.ref_mask[pix]) return;
  atomicMin(&v.zbuf[pix],
            ((unsigned long long)__float_as_uint(r.depth) << 32) | (unsigned)tid);
}

struct PhotoParams {
  int H, W, iterations;
  float fx, fy, cx, cy, min_depth, max_depth;
  float weight_depth, d_threshold, e_threshold;
  double scaling;
};

// 6x6 symmetric positive definite solve (LDL^T without pivoting), then the pose update of
// photometric_cuda.cu:383-392: dX = (expMapSO3(dx.tail(3)), dx.head(3)), X = X * dX in float.
__device__ void solve_and_update(PhotoDevice *d, const double *acc, int HW) {
  const double chi = acc[27], good = acc[28], inliers = acc[29];
  for (int k = 0; k < 30; k++) d->last[k] = acc[k];
  if (good == 0.0 || inliers == 0.0) {
    d->stats[0] = (float)chi;
    d->stats[1] = (float)inliers / (float)HW;
    return;
  }
  d->stats[0] = (float)chi / (float)good;
  d->stats[1] = (float)inliers / (float)HW;
  double A[6][6], b[6], x[6];
  {
    int k = 0;
    for (int i = 0; i < 6; i++)
      for (int j = i; j < 6; j++) A[i][j] = A[j][i] = acc[k++];
  }
  for (int i = 0; i < 6; i++) b[i] = -acc[21 + i];
  // A = L D L^T in place (L below the diagonal, D on it)
  for (int j = 0; j < 6; j++) {
    double dj = A[j][j];
    for (int k = 0; k < j; k++) dj -= A[j][k] * A[j][k] * A[k][k];
    A[j][j] = dj;
    for (int i = j + 1; i < 6; i++) {
      double s = A[i][j];
      for (int k = 0; k < j; k++) s -= A[i][k] * A[j][k] * A[k][k];
      A[i][j] = s / dj;
    }
  }
  for (int i = 0; i < 6; i++) {
    double s = b[i];
    for (int k = 0; k < i; k++) s -= A[i][k] * x[k];
    x[i] = s;
  }
  for (int i = 0; i < 6; i++) x[i] /= A[i][i];
  for (int i = 5; i >= 0; i--) {
    double s = x[i];
    for (int k = i + 1; k < 6; k++) s -= A[k][i] * x[k];
    x[i] = s;
  }
  // lie::expMapSO3 (tools/lie_algebra.h:41-53)
  const double wx = x[3], wy = x[4], wz = x[5];
  const double th2 = wx * wx + wy * wy + wz * wz, th = sqrt(th2);
  double Wm[3][3] = {{0, -wz, wy}, {wz, 0, -wx}, {-wy, wx, 0}}, dR[3][3];
  if (th2 < 1e-8) {
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) dR[i][j] = (i == j ? 1.0 : 0.0) + Wm[i][j];
  } else {
    const double s = sin(th), omc = 2.0 * sin(th / 2.0) * sin(th / 2.0);
    double K[3][3];
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) K[i][j] = Wm[i][j] / th;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) {
        double kk = 0.0;
        for (int k = 0; k < 3; k++) kk += K[i][k] * K[k][j];
        dR[i][j] = (i == j ? 1.0 : 0.0) + s * K[i][j] + omc * kk;
      }
  }
  float Rn[9], tn[3];
  for (int i = 0; i < 3; i++) {
    for (int j = 0; j < 3; j++) {
      float s = 0.f;
      for (int k = 0; k < 3; k++) s += d->R[3 * i + k] * (float)dR[k][j];
      Rn[3 * i + j] = s;
    }
    float s = 0.f;
    for (int k = 0; k < 3; k++) s += d->R[3 * i + k] * (float)x[k];
    tn[i] = s + d->t[i];
  }
  for (int i = 0; i < 9; i++) d->R[i] = Rn[i];
  for (int i = 0; i < 3; i++) d->t[i] = tn[i];
}

__global__ void __launch_bounds__(256)
linearize_kernel(PhotoParams q, PhotoView v) {
  __shared__ double s_part[8][30];
  __shared__ bool s_last;
  const int H = q.H, W = q.W, n = H * W;
  const int pix = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  double acc[30];
#pragma unroll
  for (int k = 0; k < 30; k++) acc[k] = 0.0;
  bool good = false;
  if (pix < n) {
    const unsigned long long key = v.zbuf[pix];
    if (key != ~0ull) {
      const int t = (int)(unsigned)(key & 0xffffffffull);
      const Projected r = project_point(v.dev, v.cloud[t], q.fx, q.fy, q.cx, q.cy, q.min_depth,
                                        q.max_depth);
      // getSubPixel (tools/image_set.cuh:75-107) at the un-rounded image point
      const int r0 = (int)r.v, c0 = (int)r.u;
      if ((unsigned)r0 < (unsigned)H && (unsigned)c0 < (unsigned)W && (unsigned)(r0 + 1) < (unsigned)H &&
          (unsigned)(c0 + 1) < (unsigned)W) {
        const int i00 = r0 * W + c0, i01 = i00 + 1, i10 = i00 + W, i11 = i10 + 1;
        if (!(v.ref_mask[i00] | v.ref_mask[i01] | v.ref_mask[i10] | v.ref_mask[i11])) {
          const float dr = r.v - (float)r0, dc = r.u - (float)c0, dr1 = 1.f - dr, dc1 = 1.f - dc;
          const float meas = (v.ref_depth[i00] * dc1 + v.ref_depth[i01] * dc) * dr1 +
                             (v.ref_depth[i10] * dc1 + v.ref_depth[i11] * dc) * dr;
          const float gc = (v.ref_dc[i00] * dc1 + v.ref_dc[i01] * dc) * dr1 +
                           (v.ref_dc[i10] * dc1 + v.ref_dc[i11] * dc) * dr;
          const float gr = (v.ref_dr[i00] * dc1 + v.ref_dr[i01] * dc) * dr1 +
                           (v.ref_dr[i10] * dc1 + v.ref_dr[i11] * dc) * dr;
          const float e = (r.depth - meas) * q.weight_depth;
          if (!(e * e > q.d_threshold)) {
            good = true;
            // errorAndJacobian (photometric_cuda.cu:99-235), spherical camera, depth row
            const float ir = 1.f / r.depth, ir2 = ir * ir;
            const float rxy2 = r.tx * r.tx + r.ty * r.ty;
            const float irxy2 = (float)(1.0 / (double)rxy2);
            const float rxy = sqrtf(rxy2);
            const float irxy = (float)(1.0 / (double)rxy);
            // J_proj (2x3)
            const float p00 = -q.fx * r.ty * irxy2, p01 = q.fx * r.tx * irxy2;
            const float p10 = -q.fy * r.tx * r.tz * irxy * ir2, p11 = -q.fy * r.ty * r.tz * irxy * ir2,
                        p12 = q.fy * rxy * ir2;
            // d(depth error) / d(transformed point): J_depth - derivatives * J_proj
            const float a0 = r.tx * ir - (gc * p00 + gr * p10);
            const float a1 = r.ty * ir - (gc * p01 + gr * p11);
            const float a2 = r.tz * ir - gr * p12;
            // J_icp = [R | -2 R skew(point)]
            const float *R = v.dev->R;
            float J[6];
#pragma unroll
            for (int c = 0; c < 3; c++) J[c] = a0 * R[c] + a1 * R[3 + c] + a2 * R[6 + c];
            {
              const float S[3][3] = {{0.f, -r.pz, r.py}, {r.pz, 0.f, -r.px}, {-r.py, r.px, 0.f}};
#pragma unroll
              for (int c = 0; c < 3; c++) {
                float m0 = 0.f, m1 = 0.f, m2 = 0.f;   // column c of R * skew(point)
#pragma unroll
                for (int k = 0; k < 3; k++) {
                  m0 += R[k] * S[k][c];
                  m1 += R[3 + k] * S[k][c];
                  m2 += R[6 + k] * S[k][c];
                }
                J[3 + c] = -2.f * (a0 * m0 + a1 * m1 + a2 * m2);
              }
            }
#pragma unroll
            for (int c = 0; c < 6; c++) J[c] *= q.weight_depth;
            const float chi = e * e;
            float lambda = 1.f;
            bool inlier = true;
            if (chi > q.e_threshold) {
              lambda = sqrtf(q.e_threshold / chi);
              inlier = false;
            }
            int k = 0;
#pragma unroll
            for (int i = 0; i < 6; i++)
#pragma unroll
              for (int j = i; j < 6; j++) acc[k++] = (double)(J[i] * J[j] * lambda) * q.scaling;
#pragma unroll
            for (int i = 0; i < 6; i++) acc[21 + i] = (double)(J[i] * e * lambda) * q.scaling;
            acc[27] = (double)chi;
            acc[28] = 1.0;
            acc[29] = inlier ? 1.0 : 0.0;
          }
        }
      }
    }
  }
  // warp -> CTA -> global reduction of the 30 sums, skipped by warps with nothing to add
  const bool any = __any_sync(0xffffffffu, good);
  if (any) {
#pragma unroll
    for (int k = 0; k < 30; k++) {
      double x = acc[k];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
      acc[k] = x;
    }
  }
  if (lane < 30) s_part[wid][lane] = 0.0;
  __syncwarp();
  if (any && lane == 0) {
#pragma unroll
    for (int k = 0; k < 30; k++) s_part[wid][k] = acc[k];
  }
  __syncthreads();
  if (threadIdx.x < 30) {
    double x = 0.0;
    for (int w = 0; w < 8; w++) x += s_part[w][threadIdx.x];
    if (x != 0.0) atomicAdd(&v.dev->accum[threadIdx.x], x);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(&v.dev->ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (s_last && threadIdx.x == 0) {
    __threadfence();
    double tot[30];
    for (int k = 0; k < 30; k++) tot[k] = atomicAdd(&v.dev->accum[k], 0.0);
    solve_and_update(v.dev, tot, n);
    for (int k = 0; k < 30; k++) v.dev->accum[k] = 0.0;
    v.dev->ticket = 0u;
    __threadfence();
  }
}

}  // namespace

extern "C" {

size_t gsa_photo_workspace_bytes(int H, int W) {
  if (H <= 0 || W <= 0) return 0;
  return photo_bytes(H, W);
}

int gsa_photo_set_reference(void *workspace, int H, int W, const float *depth,
                            float depth_derivative_threshold, int mask_grow_radius, void *stream) {
  if (!workspace || !depth || H <= 0 || W <= 0 || mask_grow_radius < 0)
    return splat_fail("gsa_photo_set_reference: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  const PhotoView v = make_photo_view(workspace, H, W);
  const int n = H * W;
  const dim3 blk(32, 8), grd((W + 31) / 32, (H + 7) / 8);
  ref_init_kernel<<<(n + 255) / 256, 256, 0, LS(st)>>>(n, depth, v);
  ref_grow_kernel<<<(n + 255) / 256, 256, 0, LS(st)>>>(H, W, mask_grow_radius, v);
  ref_deriv_kernel<<<grd, blk, 0, LS(st)>>>(
      H, W, depth_derivative_threshold * depth_derivative_threshold, v);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : splat_fail("gsa_photo_set_reference: %s", cudaGetErrorString(e));
}

int gsa_photo_debug_reference(void *workspace, int H, int W, float *out, void *stream) {
  if (!workspace || !out || H <= 0 || W <= 0) return splat_fail("gsa_photo_debug_reference: bad arguments");
  const PhotoView v = make_photo_view(workspace, H, W);
  const int n = H * W;
  ref_export_kernel<<<(n + 255) / 256, 256, 0, LS((cudaStream_t)stream)>>>(n, v, out);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : splat_fail("gsa_photo_debug_reference: %s", cudaGetErrorString(e));
}

int gsa_photo_set_query(void *workspace, int H, int W, const float *depth, float fx, float fy,
                        float cx, float cy, float min_depth, float max_depth, void *stream) {
  if (!workspace || !depth || H <= 0 || W <= 0)
    return splat_fail("gsa_photo_set_query: bad arguments");
  const PhotoView v = make_photo_view(workspace, H, W);
  const dim3 blk(32, 8), grd((W + 31) / 32, (H + 7) / 8);
  query_cloud_kernel<<<grd, blk, 0, LS((cudaStream_t)stream)>>>(H, W, depth, 1.f / fx, 1.f / fy, cx, cy,
                                                                min_depth, max_depth, v);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : splat_fail("gsa_photo_set_query: %s", cudaGetErrorString(e));
}

int gsa_photo_align(void *workspace, const gsa_photo_params *p, const float *initial_guess,
                    int iterations, float *transform, float *stats, double *system) {
  if (!workspace || !p || !initial_guess || !transform || !stats)
    return splat_fail("gsa_photo_align: NULL argument");
  if (p->H <= 0 || p->W <= 0 || iterations < 0) return splat_fail("gsa_photo_align: bad sizes");
  cudaStream_t st = (cudaStream_t)p->stream;
  const PhotoView v = make_photo_view(workspace, p->H, p->W);
  const int n = p->H * p->W;
  PhotoDevice init;
  memset(&init, 0, sizeof init);
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 3; c++) init.R[3 * r + c] = initial_guess[4 * r + c];
    init.t[r] = initial_guess[4 * r + 3];
  }
  cudaError_t e = cudaMemcpyAsync(v.dev, &init, sizeof init, cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return splat_fail("gsa_photo_align: %s", cudaGetErrorString(e));
  PhotoParams q;
  q.H = p->H; q.W = p->W; q.iterations = iterations;
  q.fx = p->fx; q.fy = p->fy; q.cx = p->cx; q.cy = p->cy;
  q.min_depth = p->min_depth; q.max_depth = p->max_depth;
  q.weight_depth = sqrtf(p->omega_depth);
  q.d_threshold = p->depth_rejection_threshold;
  q.e_threshold = p->kernel_chi_threshold;
  q.scaling = 1.0 / (double)n;
  const int blocks = (n + 255) / 256;
  for (int it = 0; it < iterations; it++) {
    cudaMemsetAsync(v.zbuf, 0xff, (size_t)n * 8, st);
    project_kernel<<<blocks, 256, 0, LS(st)>>>(p->H, p->W, q.fx, q.fy, q.cx, q.cy, q.min_depth,
                                               q.max_depth, v);
    linearize_kernel<<<blocks, 256, 0, LS(st)>>>(q, v);
  }
  PhotoDevice out;
  e = cudaMemcpyAsync(&out, v.dev, sizeof out, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) return splat_fail("gsa_photo_align: %s", cudaGetErrorString(e));
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 3; c++) transform[4 * r + c] = out.R[3 * r + c];
    transform[4 * r + 3] = out.t[r];
  }
  transform[12] = transform[13] = transform[14] = 0.f;
  transform[15] = 1.f;
  stats[0] = out.stats[0];
  stats[1] = out.stats[1];
  if (system)
    for (int k = 0; k < 30; k++) system[k] = out.last[k];
  return 0;
}

}  // extern "C"
