// dssr_render_bwd.cu — reverse-order tile compositing backward and the per-surfel VJP.
//
// Replaces (submodules/diff-surfel-spherical-rasterization/cuda_rasterizer/):
//   renderCUDA<3> (backward)    backward.cu:19-301   (launch :413-425)
//   preprocessCUDA<3> (backward) backward.cu:377-396 with compute_transmat_aabb :303-375
//   quat_to_rotmat_vjp          auxiliary.h:300-339
//
// The reference issues 13 global float atomicAdds per (pixel, surfel) pair, all 256
// threads of a tile hitting the same 13 addresses.  Here a warp owns an 8x4 pixel
// sub-tile, walks its tile's list back to front from the furthest entry any of its
// pixels blended, skips entries whose cull box misses the sub-tile (exact, see
// dssr_render_fwd.cu), and for every surviving entry reduces the 13 per-pixel
// contributions across the warp with a 16-shuffle transposed butterfly, after which 13
// lanes issue ONE atomic instruction (one RED per component per warp instead of up to
// 32).  The reference's backward is not the adjoint of its forward (SURVEY.md §8 a10,
// quirks Q1-Q4, Q7); it is reproduced as written.
#include "raster_math.cuh"

namespace {

constexpr int BW_WARPS = 4;
constexpr int SUB_W = 8, SUB_H = 4;

__global__ void __launch_bounds__(BW_WARPS * 32)
render_backward_kernel(const uint2 *__restrict__ ranges,
                       const uint32_t *__restrict__ point_list,
                       const float4 *__restrict__ rec, const float4 *__restrict__ cull,
                       const float *__restrict__ projmatrix, int W, int H, int tiles_x,
                       const float *__restrict__ final_T,
                       const uint32_t *__restrict__ n_contrib,
                       const float *__restrict__ dL_dout, float *__restrict__ dL_dtransMat,
                       float *__restrict__ dL_dnormal, float *__restrict__ dL_dopacity) {
  __shared__ float4 s_rec[BW_WARPS][32][4];
  __shared__ uint32_t s_id[BW_WARPS][32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int x0 = blockIdx.x * SPLAT_TILE + (wid & 1) * SUB_W;
  const int y0 = blockIdx.y * (2 * SUB_H) + (wid >> 1) * SUB_H;
  const int px = x0 + (lane & 7), py = y0 + (lane >> 3);
  const bool inside = px < W && py < H;
  const size_t HW = (size_t)W * H;
  const size_t pix = (size_t)py * W + px;

  const uint32_t last_contributor = inside ? n_contrib[pix] : 0u;
  uint32_t max_last = last_contributor;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    max_last = max(max_last, __shfl_xor_sync(0xffffffffu, max_last, o));
  if (max_last == 0) return;

  const float fx = projmatrix[0], fy = projmatrix[5], cx = projmatrix[8], cy = projmatrix[9];
  const PixelRay ray = make_ray((float)px, (float)py, fx, fy, cx, cy);
  const float inv_period = 1.0f / (6.2831853f * fabsf(fx));
  const float el_a = fabsf((0.0f - cy) / fy), el_b = fabsf(((float)(H - 1) - cy) / fy);
  const bool cull_on = fmaxf(el_a, el_b) < 1.55f;
  const float bx0 = (float)x0, bx1 = (float)(x0 + SUB_W - 1);
  const float by0 = (float)y0, by1 = (float)(y0 + SUB_H - 1);
  const uint2 range = ranges[(y0 / SPLAT_TILE) * tiles_x + blockIdx.x];
  const uint32_t lt = (1u << lane) - 1u;

  const float T_final = inside ? final_T[pix] : 0.0f;
  float T = T_final;
  const float final_D = inside ? final_T[pix + HW] : 0.0f;
  const float final_D2 = inside ? final_T[pix + 2 * HW] : 0.0f;
  const float final_A = 1.0f - T_final;
  const uint32_t median_contributor = inside ? n_contrib[pix + HW] : 0u;
  float gD = 0.f, gA = 0.f, gN0 = 0.f, gN1 = 0.f, gN2 = 0.f, gMed = 0.f, gReg = 0.f;
  if (inside) {
    gD = dL_dout[pix];
    gA = dL_dout[pix + HW];
    gN0 = dL_dout[pix + 2 * HW];
    gN1 = dL_dout[pix + 3 * HW];
    gN2 = dL_dout[pix + 4 * HW];
    gMed = dL_dout[pix + 5 * HW];
    gReg = dL_dout[pix + 6 * HW];
  }
  float accum_depth_rec = 0.f, accum_alpha_rec = 0.f;
  float accum_n0 = 0.f, accum_n1 = 0.f, accum_n2 = 0.f;
  float last_alpha = 0.f, last_depth = 0.f, last_n0 = 0.f, last_n1 = 0.f, last_n2 = 0.f;
  float last_dL_dT = 0.f;

  for (int chunk = (int)((max_last - 1) >> 5); chunk >= 0; chunk--) {
    const uint32_t q0 = (uint32_t)chunk << 5;   // 0-based list position of lane 0
    const uint32_t q = q0 + lane;
    const bool have = q < max_last;
    uint32_t id = 0;
    bool hit = false;
    if (have) {
      id = point_list[range.x + q];
      hit = !cull_on || cull_overlap(cull[id], bx0, bx1, by0, by1, inv_period);
    }
    uint32_t m = __ballot_sync(0xffffffffu, hit);
    if (m == 0) continue;
    if (hit) {
      const int slot = __popc(m & lt);
      const float4 *r = rec + 4 * (size_t)id;
      float4 a = __ldg(r), b = __ldg(r + 1), c = __ldg(r + 2), d = __ldg(r + 3);
      s_rec[wid][slot][0] = a;
      s_rec[wid][slot][1] = b;
      s_rec[wid][slot][2] = c;
      s_rec[wid][slot][3] = d;
      s_id[wid][slot] = id;
    }
    __syncwarp();
    int slot = __popc(m);
    while (m) {
      const int j = 31 - __clz(m);   // back to front
      m &= ~(1u << j);
      slot--;
      const SurfelRec s = unpack_rec(s_rec[wid][slot][0], s_rec[wid][slot][1],
                                     s_rec[wid][slot][2], s_rec[wid][slot][3]);
      const uint32_t pos = q0 + j;   // == the reference's `contributor` after decrement
      float v[16];
#pragma unroll
      for (int i = 0; i < 16; i++) v[i] = 0.0f;
      bool contributes = false;
      Hit h;
      if (pos < last_contributor && intersect(ray, s, h)) {
        contributes = true;
        const float alpha = h.alpha, depth = h.depth;
        T = __fdiv_rn(T, 1.0f - alpha);
        const float w = alpha * T;
        const float dL_dweight =
            (final_D2 + depth * depth * final_A - 2.0f * depth * final_D) * gReg;
        float dL_dalpha = dL_dweight - last_dL_dT;
        last_dL_dT = dL_dweight * alpha + (1.0f - alpha) * last_dL_dT;
        float dLd_depth = 2.0f * (T * alpha) * (depth * final_A - final_D) * gReg;
        if (pos == median_contributor - 1u) dLd_depth += gMed;
        dLd_depth += w * gD;
        accum_depth_rec = last_alpha * last_depth + (1.0f - last_alpha) * accum_depth_rec;
        last_depth = depth;
        dL_dalpha += (depth - accum_depth_rec) * gD;
        accum_n0 = last_alpha * last_n0 + (1.0f - last_alpha) * accum_n0;
        accum_n1 = last_alpha * last_n1 + (1.0f - last_alpha) * accum_n1;
        accum_n2 = last_alpha * last_n2 + (1.0f - last_alpha) * accum_n2;
        last_n0 = s.nx; last_n1 = s.ny; last_n2 = s.nz;
        dL_dalpha += (s.nx - accum_n0) * gN0;
        dL_dalpha += (s.ny - accum_n1) * gN1;
        dL_dalpha += (s.nz - accum_n2) * gN2;
        v[9] = w * gN0;
        v[10] = w * gN1;
        v[11] = w * gN2;
        accum_alpha_rec = last_alpha + (1.0f - last_alpha) * accum_alpha_rec;
        dL_dalpha += (1.0f - accum_alpha_rec) * gA;
        dL_dalpha *= T;
        last_alpha = alpha;
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
          const float inv_ss = 1.0f / ss, inv_snd = 1.0f / (sn * s.d2d),
                      inv_d2d = 1.0f / s.d2d;
          const float gx = fx * ddx, gy = fy * ddy;
          const float cw = dLd_depth * inv_d2d;
          v[6] = gx * (-s.twy * inv_ss) + gy * (-s.twx * s.twz * inv_snd) + cw * s.twx;
          v[7] = gx * (s.twx * inv_ss) + gy * (-s.twy * s.twz * inv_snd) + cw * s.twy;
          v[8] = gy * sn * inv_d2d + cw * s.twz;
        }
      }
      if (__ballot_sync(0xffffffffu, contributes) == 0u) continue;
      // transposed butterfly: 16 values x 32 lanes -> lane l holds the total of v[l>>1]
      {
        const bool hi = lane & 16;
#pragma unroll
        for (int i = 0; i < 8; i++) {
          float send = hi ? v[i] : v[i + 8], keep = hi ? v[i + 8] : v[i];
          v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
      }
      {
        const bool hi = lane & 8;
#pragma unroll
        for (int i = 0; i < 4; i++) {
          float send = hi ? v[i] : v[i + 4], keep = hi ? v[i + 4] : v[i];
          v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
        }
      }
      {
        const bool hi = lane & 4;
#pragma unroll
        for (int i = 0; i < 2; i++) {
          float send = hi ? v[i] : v[i + 2], keep = hi ? v[i + 2] : v[i];
          v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
        }
      }
      {
        const bool hi = lane & 2;
        float send = hi ? v[0] : v[1], keep = hi ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
      }
      v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
      const int comp = lane >> 1;
      if ((lane & 1) == 0 && comp <= 12 && v[0] != 0.0f) {
        const size_t gid = s_id[wid][slot];
        float *dst;
        if (comp < 9) dst = dL_dtransMat + gid * 12 + (comp / 3) * 4 + (comp % 3);
        else if (comp < 12) dst = dL_dnormal + gid * 3 + (comp - 9);
        else dst = dL_dopacity + gid;
        atomicAdd(dst, v[0]);
      }
    }
    __syncwarp();
  }
}

// Per-surfel VJP: dL/dT, dL/dnormal -> dL/dmean3D, dL/dscale, dL/dquat.
// Writes every output element (zeros for culled surfels) so no memset is needed.
__global__ void __launch_bounds__(256)
backward_preprocess_kernel(int P, const int *__restrict__ radii,
                           const float *__restrict__ scales,
                           const float *__restrict__ rotations,
                           const float *__restrict__ viewmatrix,
                           const float *__restrict__ dL_dtransMat,
                           const float *__restrict__ dL_dnormal, float *dL_dmeans3D,
                           float *dL_dscales, float *dL_drot) {
  __shared__ float m[16];
  if (threadIdx.x < 16) m[threadIdx.x] = viewmatrix[threadIdx.x];
  __syncthreads();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
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
}

}  // namespace

void launch_render_backward(const BwdParams &p, const GeomView &g, const ImageView &im,
                            const BinView &b, cudaStream_t s) {
  dim3 grid(im.tiles_x, (p.H + 2 * SUB_H - 1) / (2 * SUB_H));
  render_backward_kernel<<<grid, BW_WARPS * 32, 0, s>>>(
      im.ranges, b.point_list, g.rec, g.cull, p.projmatrix, p.W, p.H, im.tiles_x,
      im.final_T, im.n_contrib, p.dL_dout, p.dL_dtransMat, p.dL_dnormal, p.dL_dopacity);
}

void launch_backward_preprocess(const BwdParams &p, cudaStream_t s) {
  backward_preprocess_kernel<<<(p.P + 255) / 256, 256, 0, s>>>(
      p.P, p.radii, p.scales, p.rotations, p.viewmatrix, p.dL_dtransMat, p.dL_dnormal,
      p.dL_dmeans3D, p.dL_dscales, p.dL_drotations);
}
