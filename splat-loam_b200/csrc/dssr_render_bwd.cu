// dssr_render_bwd.cu — reverse-order tile compositing backward and the per-surfel VJP.
//
// Replaces (submodules/diff-surfel-spherical-rasterization/cuda_rasterizer/):
//   renderCUDA<3> (backward)    backward.cu:19-301   (launch :413-425)
//   preprocessCUDA<3> (backward) backward.cu:377-396 with compute_transmat_aabb :303-375
//   quat_to_rotmat_vjp          auxiliary.h:300-339
//
// The reference issues 13 global float atomicAdds per (pixel, surfel) pair, all 256
// threads of a tile hitting the same 13 addresses.  Here a warp owns a 4x4 pixel
// sub-tile, walks its tile's list back to front from the furthest entry any of its
// pixels blended, skips entries whose cull box misses the sub-tile (exact, see
// dssr_render_fwd.cu), evaluates the survivors two at a time (one per half warp) and
// reduces each surfel's 13 per-pixel contributions inside its half warp with a
// 15-shuffle transposed butterfly, after which 13 lanes issue ONE atomic instruction
// (one RED per component per sub-tile instead of one per pixel).  The reference's backward is not the adjoint of its forward (SURVEY.md §8 a10,
// quirks Q1-Q4, Q7); it is reproduced as written.
#include "peer_exchange.cuh"
#include "raster_math.cuh"

#ifdef RENDER_TRACE
__device__ unsigned long long g_btrace[8192][4];
extern "C" __attribute__((visibility("default"))) int splat_debug_bwd_trace(unsigned long long *out) {
  return (int)cudaMemcpyFromSymbol(out, g_btrace, sizeof(g_btrace));
}
__device__ __forceinline__ unsigned long long btime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#endif

namespace {

constexpr int SUB = 4;                                            // 4x4 pixels per warp
constexpr int SUBS_PER_TILE = (SPLAT_TILE / SUB) * (SPLAT_TILE / SUB);   // 16

// Per-pixel state of the reverse traversal (backward.cu:62-98) and its constants.
struct BackState {
  float T, last_dL_dT, accum_depth_rec, accum_alpha_rec, accum_n0, accum_n1, accum_n2;
  float last_alpha, last_depth, last_n0, last_n1, last_n2;
};
struct PixelConst {
  float gD, gA, gN0, gN1, gN2, gMed, gReg, final_D, final_D2, final_A;
  uint32_t last_contributor, median_contributor;
};
struct StepOut {
  float T, w, dL_dalpha, dLd_depth;
};

// backward.cu:183-237 for one surviving (pixel, surfel) pair, as written (quirks kept)
__device__ __forceinline__ StepOut recur(BackState &b, const PixelConst &c, float alpha,
                                         float depth, float nx, float ny, float nz, uint32_t pos) {
  StepOut o;
  b.T = __fdiv_rn(b.T, 1.0f - alpha);
  o.T = b.T;
  o.w = alpha * b.T;
  const float dL_dweight = (c.final_D2 + depth * depth * c.final_A - 2.0f * depth * c.final_D) * c.gReg;
  float dL_dalpha = dL_dweight - b.last_dL_dT;
  b.last_dL_dT = dL_dweight * alpha + (1.0f - alpha) * b.last_dL_dT;
  float dLd_depth = 2.0f * (b.T * alpha) * (depth * c.final_A - c.final_D) * c.gReg;
  if (pos == c.median_contributor - 1u) dLd_depth += c.gMed;
  dLd_depth += o.w * c.gD;
  b.accum_depth_rec = b.last_alpha * b.last_depth + (1.0f - b.last_alpha) * b.accum_depth_rec;
  b.last_depth = depth;
  dL_dalpha += (depth - b.accum_depth_rec) * c.gD;
  b.accum_n0 = b.last_alpha * b.last_n0 + (1.0f - b.last_alpha) * b.accum_n0;
  b.accum_n1 = b.last_alpha * b.last_n1 + (1.0f - b.last_alpha) * b.accum_n1;
  b.accum_n2 = b.last_alpha * b.last_n2 + (1.0f - b.last_alpha) * b.accum_n2;
  b.last_n0 = nx; b.last_n1 = ny; b.last_n2 = nz;
  dL_dalpha += (nx - b.accum_n0) * c.gN0;
  dL_dalpha += (ny - b.accum_n1) * c.gN1;
  dL_dalpha += (nz - b.accum_n2) * c.gN2;
  b.accum_alpha_rec = b.last_alpha + (1.0f - b.last_alpha) * b.accum_alpha_rec;
  dL_dalpha += (1.0f - b.accum_alpha_rec) * c.gA;
  dL_dalpha *= b.T;
  b.last_alpha = alpha;
  o.dL_dalpha = dL_dalpha;
  o.dLd_depth = dLd_depth;
  return o;
}

// One WARP per CTA = one 4x4 sub-tile, CTAs in order of decreasing tile list length.  Same
// traversal machinery as the forward (sub-tile masks, double-buffered staging), walked back
// to front; lanes 0-15 handle the later surfel of a pair, lanes 16-31 the earlier one.
__global__ void __launch_bounds__(32)
render_backward_kernel(const uint2 *__restrict__ ranges, const uint32_t *__restrict__ bwd_order,
                       const uint32_t *__restrict__ point_list,
                       const uint16_t *__restrict__ sub_mask_base, const BinHeader *__restrict__ bhdr,
                       const float4 *__restrict__ rec, const float *__restrict__ projmatrix, int W,
                       int H, int tiles_x, uint32_t P, const float *__restrict__ final_T,
                       const uint32_t *__restrict__ n_contrib, const float *__restrict__ dL_dout,
                       float *__restrict__ dL_dtransMat, float *__restrict__ dL_dnormal,
                       float *__restrict__ dL_dopacity) {
  __shared__ float4 s_rec[2][32][4];
  __shared__ float s_cut[2][32];
  __shared__ uint32_t s_pos[2][32];
  __shared__ uint32_t s_id[2][32];
  const int lane = threadIdx.x, half = lane >> 4, p16 = lane & 15;
  const uint32_t work = blockIdx.x;
#ifdef RENDER_TRACE
  if (lane == 0 && work < 8192) {   // written at once: nothing extra stays live in registers
    g_btrace[work][0] = btime();
    g_btrace[work][1] = 0;
  }
#endif
  const uint32_t unit = bwd_order[work];          // sub-tiles by decreasing walk length
  const uint32_t tile = unit / SUBS_PER_TILE;
  const int sub = unit % SUBS_PER_TILE;
  const int x0 = (tile % tiles_x) * SPLAT_TILE + (sub & 3) * SUB;
  const int y0 = (tile / tiles_x) * SPLAT_TILE + (sub >> 2) * SUB;
  if (x0 >= W || y0 >= H) return;
  const int px = x0 + (p16 & 3), py = y0 + (p16 >> 2);
  const bool inside = px < W && py < H;
  const size_t HW = (size_t)W * H;
  const size_t pix = (size_t)py * W + px;

  PixelConst pc;
  pc.last_contributor = inside ? n_contrib[pix] : 0u;
  uint32_t max_last = pc.last_contributor;
#pragma unroll
  for (int o = 8; o > 0; o >>= 1)
    max_last = max(max_last, __shfl_xor_sync(0xffffffffu, max_last, o));
#ifdef RENDER_TRACE
  if (lane == 0 && work < 8192) g_btrace[work][2] = max_last;
#endif
  if (max_last == 0) return;
  // the forward stored the offset of the mask array in the buffer's header (the capacity it
  // was laid out with is not known to the backward's caller)
  const uint16_t *sub_mask = reinterpret_cast<const uint16_t *>(
      reinterpret_cast<const char *>(sub_mask_base) + bhdr->off_sub_mask);

  const float fx = projmatrix[0], fy = projmatrix[5], cx = projmatrix[8], cy = projmatrix[9];
  const PixelRay ray = make_ray((float)px, (float)py, fx, fy, cx, cy);
  const uint2 range = ranges[tile];
  const uint32_t lt = (1u << lane) - 1u;

  BackState bs;
  const float T_final = inside ? final_T[pix] : 0.0f;
  bs.T = T_final;
  bs.last_dL_dT = bs.accum_depth_rec = bs.accum_alpha_rec = 0.f;
  bs.accum_n0 = bs.accum_n1 = bs.accum_n2 = 0.f;
  bs.last_alpha = bs.last_depth = bs.last_n0 = bs.last_n1 = bs.last_n2 = 0.f;
  pc.final_D = inside ? final_T[pix + HW] : 0.0f;
  pc.final_D2 = inside ? final_T[pix + 2 * HW] : 0.0f;
  pc.final_A = 1.0f - T_final;
  pc.median_contributor = inside ? n_contrib[pix + HW] : 0u;
  pc.gD = pc.gA = pc.gN0 = pc.gN1 = pc.gN2 = pc.gMed = pc.gReg = 0.f;
  if (inside) {
    pc.gD = dL_dout[pix];
    pc.gA = dL_dout[pix + HW];
    pc.gN0 = dL_dout[pix + 2 * HW];
    pc.gN1 = dL_dout[pix + 3 * HW];
    pc.gN2 = dL_dout[pix + 4 * HW];
    pc.gMed = dL_dout[pix + 5 * HW];
    pc.gReg = dL_dout[pix + 6 * HW];
  }

  // ---- prologue: stage the top step, fetch masks / indices of the one below ----
  int chunk = (int)((max_last - 1) >> 5);
  uint32_t m;
  {
    const uint32_t q = ((uint32_t)chunk << 5) + lane;
    const bool hit = (q < max_last) && ((sub_mask[range.x + q] >> sub) & 1u);
    m = __ballot_sync(0xffffffffu, hit);
    if (hit) {
      const int slot = __popc(m & lt);
      const uint32_t id = min(point_list[range.x + q], P - 1u);
      const float4 *r = rec + 4 * (size_t)id;
      const float4 a = __ldg(r), b = __ldg(r + 1), c = __ldg(r + 2), d = __ldg(r + 3);
      s_rec[0][slot][0] = a; s_rec[0][slot][1] = b; s_rec[0][slot][2] = c; s_rec[0][slot][3] = d;
      s_cut[0][slot] = rho_cutoff(d.x);
      s_pos[0][slot] = q;
      s_id[0][slot] = id;
    }
  }
  uint32_t nmask = 0u, nid = 0u;
  if (chunk >= 1) {
    const uint32_t q = ((uint32_t)(chunk - 1) << 5) + lane;
    nmask = sub_mask[range.x + q];
    nid = point_list[range.x + q];
  }
  __syncwarp();
  const uint32_t hmask = half ? 0xffff0000u : 0x0000ffffu;
  int cur = 0;
  for (; chunk >= 0; chunk--, cur ^= 1) {
    // ---- issue the loads of the step below ----
    const bool nhit = (chunk >= 1) && ((nmask >> sub) & 1u);
    const uint32_t nm = __ballot_sync(0xffffffffu, nhit);
    const uint32_t id1 = min(nid, P - 1u);
    if (chunk >= 2) {
      const uint32_t q = ((uint32_t)(chunk - 2) << 5) + lane;
      nmask = sub_mask[range.x + q];
      nid = point_list[range.x + q];
    }
    float4 na, nb, nc, nd;
    if (nhit) {
      const float4 *r = rec + 4 * (size_t)id1;
      na = __ldg(r); nb = __ldg(r + 1); nc = __ldg(r + 2); nd = __ldg(r + 3);
    }
    // ---- this step's survivors, back to front, two at a time ----
    const int nsurv = __popc(m);
    for (int top = nsurv - 1; top >= 0; top -= 2) {
      // lanes 0-15 take survivor `top` (later in the list), lanes 16-31 survivor `top - 1`
      const int slot = top - half;
      const bool mine = slot >= 0;
      const int sl = mine ? slot : top;
      const SurfelRec s = unpack_rec(s_rec[cur][sl][0], s_rec[cur][sl][1], s_rec[cur][sl][2],
                                     s_rec[cur][sl][3]);
      const uint32_t pos = s_pos[cur][sl];   // == the reference's `contributor` after decrement
      Hit h;
      const bool maybe = mine && pos < pc.last_contributor && intersect_begin(ray, s, s_cut[cur][sl], h);
      if (!__any_sync(0xffffffffu, maybe)) continue;
      const bool ok = maybe && intersect_finish(ray, s, h);
      const bool ok_o = __shfl_xor_sync(0xffffffffu, (int)ok, 16) != 0;
      const float alpha_o = __shfl_xor_sync(0xffffffffu, h.alpha, 16);
      const float depth_o = __shfl_xor_sync(0xffffffffu, h.depth, 16);
      const bool ok1 = half ? ok_o : ok, ok2 = half ? ok : ok_o;
      StepOut o1, o2;
      o1.T = o1.w = o1.dL_dalpha = o1.dLd_depth = 0.f;
      o2 = o1;
      if (ok1) {
        const float4 c = s_rec[cur][top][2];
        o1 = recur(bs, pc, half ? alpha_o : h.alpha, half ? depth_o : h.depth, c.y, c.z, c.w,
                   s_pos[cur][top]);
      }
      if (ok2) {
        const float4 c = s_rec[cur][top - 1][2];
        o2 = recur(bs, pc, half ? h.alpha : alpha_o, half ? h.depth : depth_o, c.y, c.z, c.w,
                   s_pos[cur][top - 1]);
      }
      // each half finishes the gradient of ITS surfel
      if ((__ballot_sync(0xffffffffu, ok) & hmask) == 0u) continue;
      const StepOut o = half ? o2 : o1;
      float v[16];
#pragma unroll
      for (int i = 0; i < 16; i++) v[i] = 0.0f;
      if (ok) {
        const float alpha = h.alpha, depth = h.depth, T = o.T, dL_dalpha = o.dL_dalpha,
                    dLd_depth = o.dLd_depth;
        v[9] = o.w * pc.gN0;
        v[10] = o.w * pc.gN1;
        v[11] = o.w * pc.gN2;
        v[12] = h.G * dL_dalpha;   // Q1: the 0.99 clamp is ignored here
        if (h.rho3d <= h.rho2d) {
          const float inv_depth = 1.0f / depth;
          const float bTu = s.tux * h.bx + s.tuy * h.by + s.tuz * h.bz;
          const float bTv = s.tvx * h.bx + s.tvy * h.by + s.tvz * h.bz;
          const float adl = -alpha * dL_dalpha;
          const float q2 = alpha * T * dLd_depth * inv_depth;   // Q2: alpha*T applied twice
          const float dsx = adl * h.sx + bTu * q2;
          const float dsy = adl * h.sy + bTv * q2;
          const float ipz = 1.0f / h.pz;
          const float dpx = dsx * ipz, dpy = dsy * ipz;
          const float dpz = -(dsx * h.px + dsy * h.py) * (ipz * ipz);
          // k1 = dL_dp x k, k2 = dL_dp x l
          const float k1x = dpy * h.kz - dpz * h.ky, k1y = dpz * h.kx - dpx * h.kz,
                      k1z = dpx * h.ky - dpy * h.kx;
          const float k2x = dpy * h.lz - dpz * h.ly, k2y = dpz * h.lx - dpx * h.lz,
                      k2z = dpx * h.ly - dpy * h.lx;
          const float cu = dLd_depth * h.sx * inv_depth, cv = dLd_depth * h.sy * inv_depth,
                      cw = dLd_depth * inv_depth;
          v[0] = k1x * ray.ex - k2x * ray.ax + cu * h.bx;
          v[1] = k1x * ray.ey - k2x * ray.ay + cu * h.by;
          v[2] = k1x * ray.ez + cu * h.bz;
          v[3] = k1y * ray.ex - k2y * ray.ax + cv * h.bx;
          v[4] = k1y * ray.ey - k2y * ray.ay + cv * h.by;
          v[5] = k1y * ray.ez + cv * h.bz;
          v[6] = k1z * ray.ex - k2z * ray.ax + cw * h.bx;
          v[7] = k1z * ray.ey - k2z * ray.ay + cw * h.by;
          v[8] = k1z * ray.ez + cw * h.bz;
        } else {
          // 2D-filter branch (backward.cu:278-298), Q3 and Q7 kept
          const float c2 = -2.0f * alpha * dL_dalpha * 2.0f;
          const float ddx = c2 * h.dx, ddy = c2 * h.dy;
          const float ss = s.twx * s.twx + s.twy * s.twy;
          const float sn = sqrtf(ss);
          const float inv_ss = 1.0f / ss, inv_snd = 1.0f / (sn * s.d2d), inv_d2d = 1.0f / s.d2d;
          const float gx = fx * ddx, gy = fy * ddy;
          const float cw = dLd_depth * inv_d2d;
          v[6] = gx * (-s.twy * inv_ss) + gy * (-s.twx * s.twz * inv_snd) + cw * s.twx;
          v[7] = gx * (s.twx * inv_ss) + gy * (-s.twy * s.twz * inv_snd) + cw * s.twy;
          v[8] = gy * sn * inv_d2d + cw * s.twz;
        }
      }
      // transposed butterfly inside the half warp: 16 values x 16 lanes -> lane p holds v[p]
      {
        const bool hi = p16 & 8;
#pragma unroll
        for (int i = 0; i < 8; i++) {
          float send = hi ? v[i] : v[i + 8], keep = hi ? v[i + 8] : v[i];
          v[i] = keep + __shfl_xor_sync(hmask, send, 8);
        }
      }
      {
        const bool hi = p16 & 4;
#pragma unroll
        for (int i = 0; i < 4; i++) {
          float send = hi ? v[i] : v[i + 4], keep = hi ? v[i + 4] : v[i];
          v[i] = keep + __shfl_xor_sync(hmask, send, 4);
        }
      }
      {
        const bool hi = p16 & 2;
#pragma unroll
        for (int i = 0; i < 2; i++) {
          float send = hi ? v[i] : v[i + 2], keep = hi ? v[i + 2] : v[i];
          v[i] = keep + __shfl_xor_sync(hmask, send, 2);
        }
      }
      {
        const bool hi = p16 & 1;
        float send = hi ? v[0] : v[1], keep = hi ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(hmask, send, 1);
      }
      if (p16 <= 12 && v[0] != 0.0f) {
        const size_t gid = s_id[cur][sl];
        float *dst;
        if (p16 < 9) dst = dL_dtransMat + gid * 12 + (p16 / 3) * 4 + (p16 % 3);
        else if (p16 < 12) dst = dL_dnormal + gid * 3 + (p16 - 9);
        else dst = dL_dopacity + gid;
        atomicAdd(dst, v[0]);
      }
    }
    // ---- park the records of the step below in the other buffer ----
    if (nhit) {
      const int slot = __popc(nm & lt);
      s_rec[cur ^ 1][slot][0] = na; s_rec[cur ^ 1][slot][1] = nb;
      s_rec[cur ^ 1][slot][2] = nc; s_rec[cur ^ 1][slot][3] = nd;
      s_cut[cur ^ 1][slot] = rho_cutoff(nd.x);
      s_pos[cur ^ 1][slot] = ((uint32_t)(chunk - 1) << 5) + lane;
      s_id[cur ^ 1][slot] = id1;
    }
    m = nm;
    __syncwarp();
  }
#ifdef RENDER_TRACE
  if (lane == 0 && blockIdx.x < 8192) g_btrace[blockIdx.x][1] = btime();
#endif
}

// Per-surfel VJP: dL/dT, dL/dnormal -> dL/dmean3D, dL/dscale, dL/dquat.
// Writes every output element (zeros for culled surfels) so no memset is needed.
// PUSH: the [10] row also goes to the staging slot of the rank that owns it (NVLink store)
// and the last CTA publishes this rank's arrival to all peers - the reduce-scatter half of the
// window all-reduce, overlapped row by row with the arithmetic above it.
template <bool PUSH>
__global__ void __launch_bounds__(256)
backward_preprocess_kernel(int P, const int *__restrict__ radii,
                           const float *__restrict__ scales,
                           const float *__restrict__ rotations,
                           const float *__restrict__ viewmatrix,
                           const float *__restrict__ dL_dtransMat,
                           const float *__restrict__ dL_dnormal,
                           const float *__restrict__ dL_dopacity, float *dL_dmeans3D,
                           float *dL_dscales, float *dL_drot, float *packed,
                           int packed_accumulate, XchgView xv) {
  static_assert(XCHG_SLICE_ALIGN == 256, "one CTA = one 256-row block with a single owner");
  __shared__ __align__(16) float s_rows[PUSH ? XCHG_SLICE_ALIGN * XCHG_ROW : 4];
  __shared__ float m[16];
  if (threadIdx.x < 16) m[threadIdx.x] = viewmatrix[threadIdx.x];
  __syncthreads();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < P) {
  float3 dmean = make_float3(0.f, 0.f, 0.f);
  float2 dscale = make_float2(0.f, 0.f);
  float4 dq = make_float4(0.f, 0.f, 0.f, 0.f);
  if (radii[idx] > 0) {
    const float4 *T4 = reinterpret_cast<const float4 *>(dL_dtransMat) + 3 * (size_t)idx;
    const float4 du = T4[0], dv = T4[1], dw = T4[2];
    // dL_dmean3D = vec3(dL_dTw) * mat3(world2sensor): dot with each column
    dmean.x = m[0] * dw.x + m[1] * dw.y + m[2] * dw.z;
    dmean.y = m[4] * dw.x + m[5] * dw.y + m[6] * dw.z;
    dmean.z = m[8] * dw.x + m[9] * dw.y + m[10] * dw.z;
    if (scales != nullptr && rotations != nullptr) {
      const float2 sc = reinterpret_cast<const float2 *>(scales)[idx];
      const float4 qq = reinterpret_cast<const float4 *>(rotations)[idx];
      const float nx = dL_dnormal[3 * (size_t)idx], ny = dL_dnormal[3 * (size_t)idx + 1],
                  nz = dL_dnormal[3 * (size_t)idx + 2];
      const float sinv = rsqrtf(qq.x * qq.x + qq.y * qq.y + qq.z * qq.z + qq.w * qq.w);
      const float w = qq.x * sinv, x = qq.y * sinv, y = qq.z * sinv, z = qq.w * sinv;
      // R columns (auxiliary.h:293-297)
      const float r00 = 1.f - 2.f * (y * y + z * z), r01 = 2.f * (x * y + w * z),
                  r02 = 2.f * (x * z - w * y);
      const float r10 = 2.f * (x * y - w * z), r11 = 1.f - 2.f * (x * x + z * z),
                  r12 = 2.f * (y * z + w * x);
      // dL_dR columns before the change of frame; Q5: scale_modifier is NOT applied
      float c0[3] = {sc.x * du.x, sc.x * du.y, sc.x * du.z};
      float c1[3] = {sc.y * dv.x, sc.y * dv.y, sc.y * dv.z};
      float c2[3] = {nx, ny, nz};
      // vR[c][r] = (V3^T * col_c)[r] = dot(V3 column r, col_c)
      float vR[3][3];
#pragma unroll
      for (int r = 0; r < 3; r++) {
        vR[0][r] = m[4 * r] * c0[0] + m[4 * r + 1] * c0[1] + m[4 * r + 2] * c0[2];
        vR[1][r] = m[4 * r] * c1[0] + m[4 * r + 1] * c1[1] + m[4 * r + 2] * c1[2];
        vR[2][r] = m[4 * r] * c2[0] + m[4 * r + 1] * c2[1] + m[4 * r + 2] * c2[2];
      }
      // dL_dscale = dot(dL_dTu, V3 * R[0]), dot(dL_dTv, V3 * R[1])
      const float a0 = m[0] * r00 + m[4] * r01 + m[8] * r02,
                  a1 = m[1] * r00 + m[5] * r01 + m[9] * r02,
                  a2 = m[2] * r00 + m[6] * r01 + m[10] * r02;
      const float b0 = m[0] * r10 + m[4] * r11 + m[8] * r12,
                  b1 = m[1] * r10 + m[5] * r11 + m[9] * r12,
                  b2 = m[2] * r10 + m[6] * r11 + m[10] * r12;
      dscale.x = du.x * a0 + du.y * a1 + du.z * a2;
      dscale.y = dv.x * b0 + dv.y * b1 + dv.z * b2;
      // quat_to_rotmat_vjp (auxiliary.h:300-339); Q6: w.r.t. the normalised quaternion
      dq.x = 2.f * (x * (vR[1][2] - vR[2][1]) + y * (vR[2][0] - vR[0][2]) +
                    z * (vR[0][1] - vR[1][0]));
      dq.y = 2.f * (-2.f * x * (vR[1][1] + vR[2][2]) + y * (vR[0][1] + vR[1][0]) +
                    z * (vR[0][2] + vR[2][0]) + w * (vR[1][2] - vR[2][1]));
      dq.z = 2.f * (x * (vR[0][1] + vR[1][0]) - 2.f * y * (vR[0][0] + vR[2][2]) +
                    z * (vR[1][2] + vR[2][1]) + w * (vR[2][0] - vR[0][2]));
      dq.w = 2.f * (x * (vR[0][2] + vR[2][0]) + y * (vR[1][2] + vR[2][1]) -
                    2.f * z * (vR[0][0] + vR[1][1]) + w * (vR[0][1] - vR[1][0]));
    }
  }
  dL_dmeans3D[3 * (size_t)idx] = dmean.x;
  dL_dmeans3D[3 * (size_t)idx + 1] = dmean.y;
  dL_dmeans3D[3 * (size_t)idx + 2] = dmean.z;
  reinterpret_cast<float2 *>(dL_dscales)[idx] = dscale;
  reinterpret_cast<float4 *>(dL_drot)[idx] = dq;
  if (packed || PUSH) {   // [means3D 3 | opacity 1 | scales 2 | rotations 4]
    float2 r0 = make_float2(dmean.x, dmean.y), r1 = make_float2(dmean.z, dL_dopacity[idx]),
           r2 = dscale, r3 = make_float2(dq.x, dq.y), r4 = make_float2(dq.z, dq.w);
    if (packed) {
      float2 *row = reinterpret_cast<float2 *>(packed + 10 * (size_t)idx);
      if (packed_accumulate) {
        const float2 o0 = row[0], o1 = row[1], o2 = row[2], o3 = row[3], o4 = row[4];
        r0.x += o0.x; r0.y += o0.y; r1.x += o1.x; r1.y += o1.y; r2.x += o2.x; r2.y += o2.y;
        r3.x += o3.x; r3.y += o3.y; r4.x += o4.x; r4.y += o4.y;
      }
      row[0] = r0; row[1] = r1; row[2] = r2; row[3] = r3; row[4] = r4;
    }
    if (PUSH) {
      float2 *row = reinterpret_cast<float2 *>(s_rows + XCHG_ROW * threadIdx.x);
      row[0] = r0; row[1] = r1; row[2] = r2; row[3] = r3; row[4] = r4;
    }
  }
  }
  if (PUSH) {
    const int row0 = blockIdx.x * blockDim.x;
    if (idx >= P) {
      float2 *row = reinterpret_cast<float2 *>(s_rows + XCHG_ROW * threadIdx.x);
      row[0] = row[1] = row[2] = row[3] = row[4] = make_float2(0.f, 0.f);
    }
    __syncthreads();
    xchg_store_block(xv, row0, min((int)blockDim.x, P - row0), s_rows);
    xchg_publish(xv, 0);
  }
}

}  // namespace

void launch_render_backward(const BwdParams &p, const GeomView &g, const ImageView &im,
                            const BinView &b, cudaStream_t s) {
  render_backward_kernel<<<im.tiles * SUBS_PER_TILE, 32, 0, LS(s)>>>(
      im.ranges, im.bwd_order, b.point_list, reinterpret_cast<const uint16_t *>(b.hdr), b.hdr, g.rec,
      p.projmatrix, p.W, p.H, im.tiles_x, (uint32_t)p.P, im.final_T, im.n_contrib, p.dL_dout, p.dL_dtransMat, p.dL_dnormal, p.dL_dopacity);
}

void launch_backward_preprocess(const BwdParams &p, cudaStream_t s) {
  if (p.push)
    backward_preprocess_kernel<true><<<(p.P + 255) / 256, 256, 0, LS(s)>>>(
        p.P, p.radii, p.scales, p.rotations, p.viewmatrix, p.dL_dtransMat, p.dL_dnormal,
        p.dL_dopacity, p.dL_dmeans3D, p.dL_dscales, p.dL_drotations, p.packed,
        p.packed_accumulate, *p.push);
  else
    backward_preprocess_kernel<false><<<(p.P + 255) / 256, 256, 0, LS(s)>>>(
        p.P, p.radii, p.scales, p.rotations, p.viewmatrix, p.dL_dtransMat, p.dL_dnormal,
        p.dL_dopacity, p.dL_dmeans3D, p.dL_dscales, p.dL_drotations, p.packed,
        p.packed_accumulate, XchgView());
}
