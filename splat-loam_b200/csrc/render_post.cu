// render_post.cu — the map post-processing of the reference's `render()` as ONE kernel per
// direction.
//
// Replaces, for a 7xHxW rasterizer output, gaussian_renderer/__init__.py:51-92 and
// utils/graphic_utils.py:26-88 (depth_to_points, depth_to_normal): about 25 PyTorch kernels
// forward (two torch.linalg.inv, a meshgrid, sin/cos, matmuls, masked in-place divisions, a
// cross product, normalize, ...) and about twice that through autograd backward.
//
//   rend_normal = (W[:3,:3] n_view) / alpha          where alpha > 0
//   surf_depth  = (D / alpha) (1 - r) + D_median r    (D / alpha only where alpha > 0)
//   surf_normal = alpha * normalize((P[y+1,x] - P[y-1,x]) x (P[y,x+1] - P[y,x-1])),  0 on the border
//   P = surf_depth * ray(x, y) + origin,  ray = R_sw [cos az cos el, sin az cos el, sin el],
//   (az, el, .) = K^-1 (x - 0.5, y - 0.5, 1)
//
// The origin cancels in the differences; it is never added here (the reference adds and
// subtracts it in fp32, which only costs it precision).  A CTA handles a 32x8 pixel tile;
// points (and, backward, the cross-product cotangents) of the tile plus halo live in
// shared memory, so every input is read once.
#include <cmath>

#include "../../include/splat_b200.h"
#include "common.cuh"

namespace {

constexpr int TX = 32, TY = 8;

struct PostCam {
  float Kinv[9];   // row-major inverse of K = projection_matrix[:3,:3]^T
  float Rsw[9];    // sensor -> world rotation = inv(world_view_transform^T)[:3,:3]
  float M[9];      // world_view_transform[:3,:3] as stored (applied to view normals)
};

// 3x3 inverse by the adjugate, in double (thread 0 of each CTA; ~60 flops)
__device__ void inverse3(const double a[9], double out[9]) {
  const double c0 = a[4] * a[8] - a[5] * a[7], c1 = a[5] * a[6] - a[3] * a[8],
               c2 = a[3] * a[7] - a[4] * a[6];
  const double inv_det = 1.0 / (a[0] * c0 + a[1] * c1 + a[2] * c2);
  out[0] = c0 * inv_det;
  out[1] = (a[2] * a[7] - a[1] * a[8]) * inv_det;
  out[2] = (a[1] * a[5] - a[2] * a[4]) * inv_det;
  out[3] = c1 * inv_det;
  out[4] = (a[0] * a[8] - a[2] * a[6]) * inv_det;
  out[5] = (a[2] * a[3] - a[0] * a[5]) * inv_det;
  out[6] = c2 * inv_det;
  out[7] = (a[1] * a[6] - a[0] * a[7]) * inv_det;
  out[8] = (a[0] * a[4] - a[1] * a[3]) * inv_det;
}

// world_view_transform is the TRANSPOSED world->sensor matrix (scene/cameras.py:43-46), so
// V = wvt^T.  The upper-left 3x3 of inv(V) is inv(V[:3,:3]) whenever V's last row is
// (0,0,0,1), which is what the reference's cameras produce; the translation only enters
// through the ray origin, which cancels.
__device__ void load_camera(const float *__restrict__ wvt, const float *__restrict__ proj,
                            PostCam &c) {
  double V[9], K[9], o[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      V[3 * i + j] = (double)wvt[4 * j + i];
      K[3 * i + j] = (double)proj[4 * j + i];
      c.M[3 * i + j] = wvt[4 * i + j];
    }
  inverse3(V, o);
  for (int i = 0; i < 9; i++) c.Rsw[i] = (float)o[i];
  inverse3(K, o);
  for (int i = 0; i < 9; i++) c.Kinv[i] = (float)o[i];
}

__device__ __forceinline__ float3 pixel_ray(const PostCam &c, int x, int y) {
  const float px = (float)x - 0.5f, py = (float)y - 0.5f;
  const float az = c.Kinv[0] * px + c.Kinv[1] * py + c.Kinv[2];
  const float el = c.Kinv[3] * px + c.Kinv[4] * py + c.Kinv[5];
  float sa, ca, se, ce;
  sincosf(az, &sa, &ca);
  sincosf(el, &se, &ce);
  const float lx = ca * ce, ly = sa * ce, lz = se;
  return make_float3(c.Rsw[0] * lx + c.Rsw[1] * ly + c.Rsw[2] * lz,
                     c.Rsw[3] * lx + c.Rsw[4] * ly + c.Rsw[5] * lz,
                     c.Rsw[6] * lx + c.Rsw[7] * ly + c.Rsw[8] * lz);
}

__device__ __forceinline__ float surf_depth_at(const float *__restrict__ allmap, size_t HW,
                                               size_t pix, float ratio) {
  const float alpha = allmap[HW + pix];
  float d = allmap[pix];
  if (alpha > 0.0f) d = d / alpha;
  return d * (1.0f - ratio) + allmap[5 * HW + pix] * ratio;
}

__device__ __forceinline__ float3 sub3(float3 a, float3 b) {
  return make_float3(a.x - b.x, a.y - b.y, a.z - b.z);
}
__device__ __forceinline__ float3 cross3(float3 a, float3 b) {
  return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ float dot3(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// Fills s_pts[(TY+2*HALO)][(TX+2*HALO)] with surf_depth * ray for the tile plus halo
// (zeros outside the image).
template <int HALO>
__device__ __forceinline__ void stage_points(const PostCam &cam, const float *__restrict__ allmap,
                                             int W, int H, float ratio, int x0, int y0,
                                             float3 (*s_pts)[TX + 2 * HALO]) {
  constexpr int SW = TX + 2 * HALO, SH = TY + 2 * HALO;
  const size_t HW = (size_t)W * H;
  for (int i = threadIdx.y * TX + threadIdx.x; i < SW * SH; i += TX * TY) {
    const int sx = i % SW, sy = i / SW;
    const int x = x0 + sx - HALO, y = y0 + sy - HALO;
    float3 p = make_float3(0.f, 0.f, 0.f);
    if (x >= 0 && x < W && y >= 0 && y < H) {
      const float d = surf_depth_at(allmap, HW, (size_t)y * W + x, ratio);
      const float3 r = pixel_ray(cam, x, y);
      p = make_float3(d * r.x, d * r.y, d * r.z);
    }
    s_pts[sy][sx] = p;
  }
}

// The mapper's per-keyframe image losses (slam/mapper.py:158-190), evaluated on the maps
// while they are still in registers:
//   geom_l1 = mean_{H*W} | valid (surf_depth - gt_depth) |
//   normal  = lambda_n * mean_{valid} (1 - <rend_normal, surf_normal>)
//   alpha   = lambda_a * mean_{valid} BCE(rend_alpha, 1)      (valid <=> gt_alpha == 1)
struct LossTerms {
  const float *gt_depth;            // [H,W]
  const unsigned char *gt_valid;    // [H,W]
  float lambda_normal, lambda_alpha;
};

struct LossWorkspace {   // 64 bytes, zeroed once by the caller; every launch leaves it zeroed
  double sum[4];         // |depth error|, 1 - <n, n'>, -log(alpha), #valid
  unsigned int ticket;
};

__device__ __forceinline__ float3 rend_normal_at(const PostCam &cam, const float *__restrict__ allmap,
                                                 size_t HW, size_t pix, float alpha) {
  const float nx = allmap[2 * HW + pix], ny = allmap[3 * HW + pix], nz = allmap[4 * HW + pix];
  float3 v = make_float3(cam.M[0] * nx + cam.M[1] * ny + cam.M[2] * nz,
                         cam.M[3] * nx + cam.M[4] * ny + cam.M[5] * nz,
                         cam.M[6] * nx + cam.M[7] * ny + cam.M[8] * nz);
  if (alpha > 0.0f) v = make_float3(v.x / alpha, v.y / alpha, v.z / alpha);
  return v;
}

template <bool LOSS>
__global__ void __launch_bounds__(TX *TY)
post_forward_kernel(int W, int H, float ratio, const float *__restrict__ allmap,
                    const float *__restrict__ wvt, const float *__restrict__ proj,
                    float *__restrict__ rend_normal, float *__restrict__ surf_depth,
                    float *__restrict__ surf_normal, LossTerms lt, LossWorkspace *ws,
                    float *__restrict__ losses) {
  __shared__ PostCam cam;
  __shared__ float3 s_pts[TY + 2][TX + 2];
  __shared__ float s_red[TX * TY / 32][4];
  if (threadIdx.x == 0 && threadIdx.y == 0) load_camera(wvt, proj, cam);
  __syncthreads();
  const int x0 = blockIdx.x * TX, y0 = blockIdx.y * TY;
  stage_points<1>(cam, allmap, W, H, ratio, x0, y0, s_pts);
  __syncthreads();
  const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
  const size_t HW = (size_t)W * H, pix = (size_t)y * W + x;
  float part[4] = {0.f, 0.f, 0.f, 0.f};
  if (x < W && y < H) {
    const float alpha = allmap[HW + pix];
    const float3 rn = rend_normal_at(cam, allmap, HW, pix, alpha);
    const float sd = surf_depth_at(allmap, HW, pix, ratio);
    float3 n = make_float3(0.f, 0.f, 0.f);
    if (x > 0 && x < W - 1 && y > 0 && y < H - 1) {
      const int sx = threadIdx.x + 1, sy = threadIdx.y + 1;
      const float3 dv = sub3(s_pts[sy + 1][sx], s_pts[sy - 1][sx]);
      const float3 dh = sub3(s_pts[sy][sx + 1], s_pts[sy][sx - 1]);
      const float3 c = cross3(dv, dh);
      const float len = fmaxf(sqrtf(dot3(c, c)), 1e-12f);   // F.normalize's eps
      n = make_float3(c.x / len * alpha, c.y / len * alpha, c.z / len * alpha);
    }
    if (rend_normal) {
      rend_normal[pix] = rn.x;
      rend_normal[HW + pix] = rn.y;
      rend_normal[2 * HW + pix] = rn.z;
    }
    if (surf_depth) surf_depth[pix] = sd;
    if (surf_normal) {
      surf_normal[pix] = n.x;
      surf_normal[HW + pix] = n.y;
      surf_normal[2 * HW + pix] = n.z;
    }
    if (LOSS) {
      const bool valid = lt.gt_valid[pix] == 1;
      part[0] = valid ? fabsf(sd - lt.gt_depth[pix]) : 0.0f;
      if (valid) {
        part[1] = 1.0f - dot3(rn, n);
        part[2] = -fmaxf(logf(alpha), -100.0f);     // BCE against target 1, torch's log clamp
        part[3] = 1.0f;
      }
    }
  }
  if (!LOSS) return;
  // CTA reduction, one double atomic per term per CTA; the last CTA normalises and re-arms
  const int tid = threadIdx.y * TX + threadIdx.x, lane = tid & 31, wid = tid >> 5;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    float v = part[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_red[wid][k] = v;
  }
  __syncthreads();
  if (tid == 0) {
    for (int k = 0; k < 4; k++) {
      double v = 0.0;
      for (int w = 0; w < TX * TY / 32; w++) v += (double)s_red[w][k];
      atomicAdd(&ws->sum[k], v);
    }
    __threadfence();
    if (atomicAdd(&ws->ticket, 1u) == gridDim.x * gridDim.y - 1u) {
      __threadfence();
      const double geom = *(volatile double *)&ws->sum[0], nrm = *(volatile double *)&ws->sum[1],
                   bce = *(volatile double *)&ws->sum[2], cnt = *(volatile double *)&ws->sum[3];
      const float l0 = (float)(geom / (double)HW);
      const float l1 = lt.lambda_normal * (float)(nrm / cnt);   // 0/0 = NaN, as torch's empty mean
      const float l2 = lt.lambda_alpha * (float)(bce / cnt);
      losses[0] = l0;
      losses[1] = l1;
      losses[2] = l2;
      losses[3] = l0 + l2 + l1;      // mapper.py:193-195: geom_l1 + alpha_loss + normal_loss
      losses[4] = (float)cnt;
      ws->sum[0] = ws->sum[1] = ws->sum[2] = ws->sum[3] = 0.0;
      ws->ticket = 0u;
    }
  }
}

// Backward.  With c = dv x dh, n = c / max(|c|, eps), surf_normal = alpha n:
//   dL/dc  = alpha (g - n (n.g)) / |c|          (alpha g / eps when |c| <= eps)
//   dL/ddv = dh x dL/dc ,  dL/ddh = dL/dc x dv
//   dL/dP[y+-1, x] = +-dL/ddv ,  dL/dP[y, x+-1] = +-dL/ddh ,  dL/dsurf_depth = ray . dL/dP
// LOSS: the upstream gradients are not read from memory but derived on the spot from the
// loss definition (weights w = dL/d(geom, normal, alpha) incl. the share of `total`):
//   g_surf_depth  = w_g sign(valid (sd - gt)) valid / (H W)
//   g_rend_normal = -w_n lambda_n valid surf_normal / n_valid
//   g_surf_normal = -w_n lambda_n valid rend_normal / n_valid
//   g_alpha      += w_a lambda_a valid (alpha - 1) / max((1 - alpha) alpha, 1e-12) / n_valid
template <bool LOSS>
__global__ void __launch_bounds__(TX *TY)
post_backward_kernel(int W, int H, float ratio, const float *__restrict__ allmap,
                     const float *__restrict__ wvt, const float *__restrict__ proj,
                     const float *__restrict__ g_rend_normal, const float *__restrict__ g_surf_depth,
                     const float *__restrict__ g_surf_normal, float *__restrict__ dL_dallmap,
                     LossTerms lt, const float *__restrict__ losses,
                     const float *__restrict__ g_losses) {
  float w_geom = 0.f, w_norm = 0.f, w_alpha = 0.f;
  if (LOSS) {
    const float gt = g_losses ? g_losses[3] : 1.0f;
    const float nv = losses[4];
    w_geom = ((g_losses ? g_losses[0] : 0.0f) + gt) / (float)((size_t)W * H);
    w_norm = ((g_losses ? g_losses[1] : 0.0f) + gt) * lt.lambda_normal / nv;
    w_alpha = ((g_losses ? g_losses[2] : 0.0f) + gt) * lt.lambda_alpha / nv;
  }
  const bool have_sn = LOSS || g_surf_normal != nullptr;
  __shared__ PostCam cam;
  __shared__ float3 s_pts[TY + 4][TX + 4];
  __shared__ float3 s_dv[TY + 2][TX + 2], s_dh[TY + 2][TX + 2];   // dL/ddv, dL/ddh per pixel
  if (threadIdx.x == 0 && threadIdx.y == 0) load_camera(wvt, proj, cam);
  __syncthreads();
  const int x0 = blockIdx.x * TX, y0 = blockIdx.y * TY;
  const size_t HW = (size_t)W * H;
  stage_points<2>(cam, allmap, W, H, ratio, x0, y0, s_pts);
  __syncthreads();
  for (int i = threadIdx.y * TX + threadIdx.x; i < (TX + 2) * (TY + 2); i += TX * TY) {
    const int sx = i % (TX + 2), sy = i / (TX + 2);
    const int x = x0 + sx - 1, y = y0 + sy - 1;
    float3 ddv = make_float3(0.f, 0.f, 0.f), ddh = ddv;
    if (have_sn && x > 0 && x < W - 1 && y > 0 && y < H - 1) {
      const size_t pix = (size_t)y * W + x;
      const float alpha = allmap[HW + pix];
      float3 g;
      if (LOSS) {
        const float k = lt.gt_valid[pix] == 1 ? -w_norm * alpha : 0.0f;
        const float3 rn = rend_normal_at(cam, allmap, HW, pix, alpha);
        g = make_float3(k * rn.x, k * rn.y, k * rn.z);
      } else {
        g = make_float3(g_surf_normal[pix] * alpha, g_surf_normal[HW + pix] * alpha,
                        g_surf_normal[2 * HW + pix] * alpha);
      }
      const float3 dv = sub3(s_pts[sy + 2][sx + 1], s_pts[sy][sx + 1]);
      const float3 dh = sub3(s_pts[sy + 1][sx + 2], s_pts[sy + 1][sx]);
      const float3 c = cross3(dv, dh);
      const float len = sqrtf(dot3(c, c));
      float3 dc;
      if (len > 1e-12f) {
        const float3 n = make_float3(c.x / len, c.y / len, c.z / len);
        const float ng = dot3(n, g);
        dc = make_float3((g.x - n.x * ng) / len, (g.y - n.y * ng) / len, (g.z - n.z * ng) / len);
      } else {
        dc = make_float3(g.x / 1e-12f, g.y / 1e-12f, g.z / 1e-12f);
      }
      ddv = cross3(dh, dc);
      ddh = cross3(dc, dv);
    }
    s_dv[sy][sx] = ddv;
    s_dh[sy][sx] = ddh;
  }
  __syncthreads();
  const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
  if (x >= W || y >= H) return;
  const size_t pix = (size_t)y * W + x;
  const int sx = threadIdx.x + 1, sy = threadIdx.y + 1;
  // pixel (x,y) is P[y+1] of the pixel above it, P[y-1] of the one below, and likewise in x
  const float3 up = s_dv[sy - 1][sx], down = s_dv[sy + 1][sx];
  const float3 left = s_dh[sy][sx - 1], right = s_dh[sy][sx + 1];
  const float3 dP = make_float3(up.x - down.x + left.x - right.x, up.y - down.y + left.y - right.y,
                                up.z - down.z + left.z - right.z);
  const float3 ray = pixel_ray(cam, x, y);
  float G_sd = dot3(ray, dP);
  const float alpha = allmap[HW + pix];
  const bool hit = alpha > 0.0f;
  const bool valid = LOSS && lt.gt_valid[pix] == 1;
  if (LOSS) {
    if (valid) {
      const float e = surf_depth_at(allmap, HW, pix, ratio) - lt.gt_depth[pix];
      G_sd += w_geom * (e > 0.0f ? 1.0f : (e < 0.0f ? -1.0f : 0.0f));
    }
  } else if (g_surf_depth) {
    G_sd += g_surf_depth[pix];
  }

  float d_alpha = 0.0f;
  float3 g_rn = make_float3(0.f, 0.f, 0.f);   // upstream gradient of rend_normal at this pixel
  if (!LOSS && g_rend_normal)
    g_rn = make_float3(g_rend_normal[pix], g_rend_normal[HW + pix], g_rend_normal[2 * HW + pix]);
  // surf_normal = alpha n  ->  dL/dalpha += g . n  (n recomputed from the staged points)
  if (have_sn && x > 0 && x < W - 1 && y > 0 && y < H - 1) {
    const float3 dv = sub3(s_pts[sy + 2][sx + 1], s_pts[sy][sx + 1]);
    const float3 dh = sub3(s_pts[sy + 1][sx + 2], s_pts[sy + 1][sx]);
    const float3 c = cross3(dv, dh);
    const float len = fmaxf(sqrtf(dot3(c, c)), 1e-12f);
    const float3 n = make_float3(c.x / len, c.y / len, c.z / len);
    if (LOSS) {
      if (valid) {
        const float3 rn = rend_normal_at(cam, allmap, HW, pix, alpha);
        d_alpha += -w_norm * dot3(rn, n);
        g_rn = make_float3(-w_norm * n.x * alpha, -w_norm * n.y * alpha, -w_norm * n.z * alpha);
      }
    } else {
      d_alpha += g_surf_normal[pix] * n.x + g_surf_normal[HW + pix] * n.y +
                 g_surf_normal[2 * HW + pix] * n.z;
    }
  }
  if (valid) d_alpha += w_alpha * (alpha - 1.0f) / fmaxf((1.0f - alpha) * alpha, 1e-12f);
  // surf_depth = (D / alpha)(1 - r) + D_med r
  const float gd = G_sd * (1.0f - ratio);
  const float D = allmap[pix];
  dL_dallmap[pix] = hit ? gd / alpha : gd;
  if (hit) d_alpha -= gd * D / (alpha * alpha);
  dL_dallmap[5 * HW + pix] = G_sd * ratio;
  dL_dallmap[6 * HW + pix] = 0.0f;
  // rend_normal_i = sum_j M[i][j] n_j / alpha
  float gn[3] = {0.f, 0.f, 0.f};
  if (LOSS || g_rend_normal) {
    const float g0 = g_rn.x, g1 = g_rn.y, g2 = g_rn.z;
#pragma unroll
    for (int j = 0; j < 3; j++) gn[j] = cam.M[j] * g0 + cam.M[3 + j] * g1 + cam.M[6 + j] * g2;
    if (hit) {
      const float nx = allmap[2 * HW + pix], ny = allmap[3 * HW + pix], nz = allmap[4 * HW + pix];
      d_alpha -= (gn[0] * nx + gn[1] * ny + gn[2] * nz) / (alpha * alpha);
#pragma unroll
      for (int j = 0; j < 3; j++) gn[j] /= alpha;
    }
  }
  dL_dallmap[2 * HW + pix] = gn[0];
  dL_dallmap[3 * HW + pix] = gn[1];
  dL_dallmap[4 * HW + pix] = gn[2];
  dL_dallmap[HW + pix] = d_alpha;
}

}  // namespace

extern "C" {

int dssr_post_forward(const dssr_post_forward_args *a) {
  if (!a) return splat_fail("dssr_post_forward: NULL argument");
  if (a->W <= 0 || a->H <= 0) return splat_fail("dssr_post_forward: bad sizes W=%d H=%d", a->W, a->H);
  if (!a->allmap || !a->viewmatrix || !a->projmatrix || !a->rend_normal || !a->surf_depth ||
      !a->surf_normal)
    return splat_fail("dssr_post_forward: NULL pointer");
  dim3 grid((a->W + TX - 1) / TX, (a->H + TY - 1) / TY), block(TX, TY);
  post_forward_kernel<false><<<grid, block, 0, LS((cudaStream_t)a->stream)>>>(
      a->W, a->H, a->depth_ratio, a->allmap, a->viewmatrix, a->projmatrix, a->rend_normal,
      a->surf_depth, a->surf_normal, LossTerms(), nullptr, nullptr);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : splat_fail("dssr_post_forward: %s", cudaGetErrorString(e));
}

int dssr_post_backward(const dssr_post_backward_args *a) {
  if (!a) return splat_fail("dssr_post_backward: NULL argument");
  if (a->W <= 0 || a->H <= 0) return splat_fail("dssr_post_backward: bad sizes W=%d H=%d", a->W, a->H);
  if (!a->allmap || !a->viewmatrix || !a->projmatrix || !a->dL_dallmap)
    return splat_fail("dssr_post_backward: NULL pointer");
  dim3 grid((a->W + TX - 1) / TX, (a->H + TY - 1) / TY), block(TX, TY);
  post_backward_kernel<false><<<grid, block, 0, LS((cudaStream_t)a->stream)>>>(
      a->W, a->H, a->depth_ratio, a->allmap, a->viewmatrix, a->projmatrix, a->dL_drend_normal,
      a->dL_dsurf_depth, a->dL_dsurf_normal, a->dL_dallmap, LossTerms(), nullptr, nullptr);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : splat_fail("dssr_post_backward: %s", cudaGetErrorString(e));
}

int dssr_post_loss_forward(const dssr_post_loss_forward_args *a) {
  if (!a) return splat_fail("dssr_post_loss_forward: NULL argument");
  if (a->W <= 0 || a->H <= 0) return splat_fail("dssr_post_loss_forward: bad sizes W=%d H=%d", a->W, a->H);
  if (!a->allmap || !a->viewmatrix || !a->projmatrix || !a->gt_depth || !a->gt_valid ||
      !a->losses || !a->workspace)
    return splat_fail("dssr_post_loss_forward: NULL pointer");
  if ((uintptr_t)a->workspace & 7) return splat_fail("dssr_post_loss_forward: workspace must be 8-byte aligned");
  static_assert(sizeof(LossWorkspace) <= DSSR_POST_LOSS_WORKSPACE_BYTES, "workspace size");
  LossTerms lt{a->gt_depth, a->gt_valid, a->lambda_normal, a->lambda_alpha};
  dim3 grid((a->W + TX - 1) / TX, (a->H + TY - 1) / TY), block(TX, TY);
  post_forward_kernel<true><<<grid, block, 0, LS((cudaStream_t)a->stream)>>>(
      a->W, a->H, a->depth_ratio, a->allmap, a->viewmatrix, a->projmatrix, a->rend_normal,
      a->surf_depth, a->surf_normal, lt, (LossWorkspace *)a->workspace, a->losses);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : splat_fail("dssr_post_loss_forward: %s", cudaGetErrorString(e));
}

int dssr_post_loss_backward(const dssr_post_loss_backward_args *a) {
  if (!a) return splat_fail("dssr_post_loss_backward: NULL argument");
  if (a->W <= 0 || a->H <= 0) return splat_fail("dssr_post_loss_backward: bad sizes W=%d H=%d", a->W, a->H);
  if (!a->allmap || !a->viewmatrix || !a->projmatrix || !a->gt_depth || !a->gt_valid ||
      !a->losses || !a->dL_dallmap)
    return splat_fail("dssr_post_loss_backward: NULL pointer");
  LossTerms lt{a->gt_depth, a->gt_valid, a->lambda_normal, a->lambda_alpha};
  dim3 grid((a->W + TX - 1) / TX, (a->H + TY - 1) / TY), block(TX, TY);
  post_backward_kernel<true><<<grid, block, 0, LS((cudaStream_t)a->stream)>>>(
      a->W, a->H, a->depth_ratio, a->allmap, a->viewmatrix, a->projmatrix, nullptr, nullptr,
      nullptr, a->dL_dallmap, lt, a->losses, a->dL_dlosses);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : splat_fail("dssr_post_loss_backward: %s", cudaGetErrorString(e));
}

}  // extern "C"
