// dssr_render_bwd.cu — reverse-order tile compositing backward and the per-surfel VJP.
//
// Replaces (submodules/diff-surfel-spherical-rasterization/cuda_rasterizer/):
//   renderCUDA<3> (backward)    backward.cu:19-301   (launch :413-425)
//   preprocessCUDA<3> (backward) backward.cu:377-396 with compute_transmat_aabb :303-375
//   quat_to_rotmat_vjp          auxiliary.h:300-339
//
// The reference re-walks every tile list with 256 threads, repeats the forward's tests at every
// (pixel, entry) pair and issues 13 global float atomicAdds per surviving pair, all threads of
// a tile hitting the same 13 addresses.  Here the forward has already written down WHICH pairs
// it blended (dssr_render_fwd.cu: one word per 32-entry step and a pixel mask per blended
// entry), and those are exactly the pairs the reference's backward ends up processing
// (backward.cu:132 skips entries past a pixel's last contributor, :152-182 repeat the
// forward's tests on the same floats).  So the backward tests nothing:
//   * a warp owns an 8x4 pixel block (two 4x4 sub-tiles) and walks the BLOCK's list (the
//     tile's entries that can touch it, radix_sort.cu) back to front, as far as the forward
//     got; per 32-entry step it reads the step word, and only the blended entries' list
//     entries (surfel index, list position), records (64 B, L2-resident) and pixel masks;
//   * the blended (pixel, entry) pairs are laid end to end and processed 32 at a time, ONE PAIR
//     PER LANE whatever its pixel and entry (P1: intersection, alpha, depth);
//   * the per-pixel recurrences (backward.cu:183-237: T, the accum_* terms, last_dL_dT), the
//     only sequential part, run on the pixel's own lane over that pixel's pairs of the batch,
//     in list order, from shared memory (P2);
//   * back on the pair lanes the gradient of the pair w.r.t. the surfel's 13 values is
//     finished (P3), the batch's contributions are summed per entry by a segmented shuffle
//     reduction (an entry's pairs sit in consecutive lanes) and one atomic per (entry,
//     component) goes out.
// The reference's backward is not the adjoint of its forward (SURVEY.md §8 a10, quirks Q1-Q4,
// Q7); it is reproduced as written.
#include "peer_exchange.cuh"
#include "raster_math.cuh"

#ifdef RENDER_TRACE
__device__ unsigned long long g_bstage[8192][5];   // cycles in gather / P1 / P2 / P3 / reduce
extern "C" __attribute__((visibility("default"))) int splat_debug_bwd_stages(unsigned long long *out) {
  return (int)cudaMemcpyFromSymbol(out, g_bstage, sizeof(g_bstage));
}
#define STAGE_CLOCK(acc) { const long long c_ = clock64(); acc += c_ - t_stage; t_stage = c_; }
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

#ifndef RENDER_TRACE
#define STAGE_CLOCK(acc)
#endif

namespace {

constexpr int SUB = 4;                                            // 4x4 pixels per sub-tile
constexpr int UNITS_PER_TILE = SPLAT_BLOCKS_PER_TILE;             // a warp = SPLAT_UNIT_SUBS sub-tiles
#ifndef COMPOSITE_ENT_CAP
#define COMPOSITE_ENT_CAP 64
#endif
constexpr int ENT_CAP = COMPOSITE_ENT_CAP;        // blended entries staged per warp
// (measured at C2, forward / backward us: 32 pairs 231 / 265, 64: 219 / 252, 96: 213 / 244, 128: 210 / 242;
//  more staged entries than 64 cost residency: ENT_CAP 128 puts the forward in two waves, 295 us)
#ifndef COMPOSITE_GATHER_PAIRS
#define COMPOSITE_GATHER_PAIRS 128
#endif
constexpr int GATHER_PAIRS = COMPOSITE_GATHER_PAIRS;   // entries are gathered until they hold this many pairs
constexpr int QUEUE_CAP = GATHER_PAIRS + 32 * 32;   // one more step can add 32 entries x 32 pixels

// Per-pixel state of the reverse traversal (backward.cu:62-98) and its constants.
struct BackState {
  float T, last_dL_dT, accum_depth_rec, accum_alpha_rec, accum_n0, accum_n1, accum_n2;
  float last_alpha, last_depth, last_n0, last_n1, last_n2;
};
struct PixelConst {
  float gD, gA, gN0, gN1, gN2, gMed, gReg, final_D, final_D2, final_A;
  uint32_t median_contributor;
};
struct StepOut {
  float T, w, dL_dalpha, dLd_depth;
};

// backward.cu:183-237 for one blended (pixel, surfel) pair, as written (quirks kept)
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

// everything a warp stages in shared memory (one warp per CTA)
struct __align__(16) WarpStage {
  float4 rec[ENT_CAP][4];           // the gathered entries' 64-byte records
  float4 pout[32];                  // P2 -> P3: T, dL_dalpha, dLd_depth, w of the batch's pairs
  uint32_t ent_id[ENT_CAP];         // surfel index
  uint32_t ent_pos[ENT_CAP];        // position in the tile's list (the reference's `contributor`)
  float ray[10][32];                // per pixel: X Y Z ax ay ex ey ez pxf pyf
  float gN[3][32];                  // per pixel: dL_dout of the normal channels
  float pa[32], pd[32];             // P1 -> P2: alpha, depth of the batch's pairs
  uint32_t pe[32];                  // P1 -> P2: entry slot of the pair
  uint32_t pmask[32];               // per pixel: which lanes of the batch hold its pairs
  uint16_t queue[QUEUE_CAP];        // the gathered pairs, (slot << 5) | pixel, back to front
};

// One WARP per CTA = one 8x4 pixel block; CTAs in order of decreasing number of pairs.
__global__ void __launch_bounds__(32, SPLAT_UNIT_SUBS == 2 ? 14 : 24)
render_backward_kernel(const uint2 *__restrict__ ranges, const uint32_t *__restrict__ bwd_order,
                       const uint32_t *__restrict__ blk_top, const char *__restrict__ bin_base,
                       const BinHeader *__restrict__ bhdr, const float4 *__restrict__ rec,
                       const float *__restrict__ projmatrix, int W, int H, int tiles_x, uint32_t P,
                       const float *__restrict__ final_T, const uint32_t *__restrict__ n_contrib,
                       const float *__restrict__ dL_dout, float *__restrict__ dL_dtransMat,
                       float *__restrict__ dL_dnormal, float *__restrict__ dL_dopacity) {
  __shared__ WarpStage st;
  const int lane = threadIdx.x, half = lane >> 4, p16 = lane & 15;
  const uint32_t full = 0xffffffffu;
#ifdef RENDER_TRACE
  if (lane == 0 && blockIdx.x < 8192) {
    g_btrace[blockIdx.x][0] = btime();
    g_btrace[blockIdx.x][1] = 0;
  }
#endif
  const uint32_t unit = bwd_order[blockIdx.x];
  const uint32_t tile = unit / UNITS_PER_TILE;
  const int sub_a = SPLAT_UNIT_SUBS * (int)(unit % UNITS_PER_TILE);           // left sub-tile; right = sub_a + 1
  const int x0 = (tile % tiles_x) * SPLAT_TILE + (sub_a & 3) * SUB;
  const int y0 = (tile / tiles_x) * SPLAT_TILE + (sub_a >> 2) * SUB;
  if (x0 >= W || y0 >= H) return;
  const int px = x0 + SUB * half + (p16 & 3), py = y0 + (p16 >> 2);
  const bool inside = half < SPLAT_UNIT_SUBS && px < W && py < H;
  const size_t HW = (size_t)W * H;
  const size_t pix = (size_t)py * W + px;

  // how far the forward walked the block's list (in 32-entry steps)
  const uint32_t top = blk_top[unit];
#ifdef RENDER_TRACE
  if (lane == 0 && blockIdx.x < 8192) g_btrace[blockIdx.x][2] = top;
#endif
  if (top == 0) return;

  // the block's list and what the forward wrote down about it (dssr_render_fwd.cu)
  uint2 range = ranges[tile];
  range.x = min(range.x, bhdr->capacity);
  range.y = min(range.y, bhdr->capacity);
  const size_t ubase = (size_t)UNITS_PER_TILE * range.x +
                       (size_t)(unit % UNITS_PER_TILE) * (range.y - range.x);
  const uint2 *bl = reinterpret_cast<const uint2 *>(bin_base + bhdr->off_blist) + ubase;
  const uint32_t *bw = reinterpret_cast<const uint32_t *>(bin_base + bhdr->off_bwords) +
                       (ubase >> 5) + unit;
  const uint32_t *bp = reinterpret_cast<const uint32_t *>(bin_base + bhdr->off_bpix) + ubase;

  const float fx = projmatrix[0], fy = projmatrix[5], cx = projmatrix[8], cy = projmatrix[9];
  {
    const PixelRay r = make_ray((float)px, (float)py, fx, fy, cx, cy);
    st.ray[0][lane] = r.X;  st.ray[1][lane] = r.Y;  st.ray[2][lane] = r.Z;
    st.ray[3][lane] = r.ax; st.ray[4][lane] = r.ay;
    st.ray[5][lane] = r.ex; st.ray[6][lane] = r.ey; st.ray[7][lane] = r.ez;
    st.ray[8][lane] = r.pxf; st.ray[9][lane] = r.pyf;
  }

  BackState bs;
  PixelConst pc;
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
  st.gN[0][lane] = pc.gN0; st.gN[1][lane] = pc.gN1; st.gN[2][lane] = pc.gN2;

  int chunk = (int)top - 1;
  // Software pipeline of the walk (back to front): the step words are fetched four steps
  // ahead of their use, the blended entries' list entries and pixel masks three steps ahead,
  // their records one step ahead (into registers) - every load's address comes from a load
  // issued at least two steps earlier, so its L2 latency sits under the work in between.
  // A warp issues in order, so ANY instruction that reads a register with a load in flight
  // stalls it - a move included.  The four stages therefore keep the RAW loaded words (they
  // are combined only when their step comes) in four fixed register sets whose roles rotate:
  // the step body is instantiated once per rotation and nothing is shifted.
  struct WalkStage { uint32_t w, idr, posr, mr; };
  WalkStage s0, s1, s2, s3;
#define LOAD_WORDS(S, c)                                                 \
  {                                                                      \
    const int c_ = (c);                                                  \
    S.w = c_ >= 0 ? bw[c_] : 0u;                                         \
  }
#define LOAD_ENTRY(S, c)                                                 \
  {                                                                      \
    S.idr = 0u; S.posr = 0u; S.mr = 0u;                                  \
    if ((S.w >> lane) & 1u) {                                            \
      const uint32_t q_ = ((uint32_t)(c) << 5) + lane;                   \
      const uint2 e_ = bl[q_];                                           \
      S.idr = e_.x; S.posr = e_.y;                                       \
      S.mr = bp[q_];                                                     \
    }                                                                    \
  }
  LOAD_WORDS(s0, chunk);
  LOAD_WORDS(s1, chunk - 1);
  LOAD_WORDS(s2, chunk - 2);
  LOAD_WORDS(s3, chunk - 3);
  LOAD_ENTRY(s0, chunk);
  LOAD_ENTRY(s1, chunk - 1);
  LOAD_ENTRY(s2, chunk - 2);
  float4 ra, rb, rc, rd;                            // records of chunk
  ra = rb = rc = rd = make_float4(0.f, 0.f, 0.f, 0.f);
  if ((s0.w >> lane) & 1u) {
    const float4 *r = rec + 4 * (size_t)min(s0.idr, P - 1u);
    ra = __ldg(r); rb = __ldg(r + 1); rc = __ldg(r + 2); rd = __ldg(r + 3);
  }
  int nent = 0, rot = 0;
  uint32_t pending = 0;
  __syncwarp();
#ifdef RENDER_TRACE
  long long t_stage = clock64(), c_gather = 0, c_p1 = 0, c_p2 = 0, c_p3 = 0, c_red = 0;
#endif

  // One step of the gather.  A: the stage of `chunk` (consumed, then re-armed for chunk - 4),
  // B: chunk - 1 (its records are requested), D: chunk - 3 (its entries are requested).
#define GATHER_STEP(A, B, C, D)                                                                  \
  {                                                                                              \
    const uint32_t vote = A.w;                                                                   \
    const uint32_t id = min(A.idr, P - 1u), mask = A.mr, pos = A.posr;                           \
    const float4 a = ra, b = rb, c = rc, d = rd;                                                 \
    LOAD_WORDS(A, chunk - 4);                                                                    \
    LOAD_ENTRY(D, chunk - 3);                                                                    \
    if ((B.w >> lane) & 1u) {                                                                    \
      const float4 *r = rec + 4 * (size_t)min(B.idr, P - 1u);                                    \
      ra = __ldg(r); rb = __ldg(r + 1); rc = __ldg(r + 2); rd = __ldg(r + 3);                    \
    }                                                                                            \
    if (vote) {                                                                                  \
      const bool any = (vote >> lane) & 1u;                                                      \
      const int ns = __popc(vote);                                                               \
      const int slot = nent + __popc(vote >> lane) - 1; /* later entries first */                \
      if (any) {                                                                                 \
        st.ent_id[slot] = id;                                                                    \
        st.ent_pos[slot] = pos;                                                                  \
        st.rec[slot][0] = a; st.rec[slot][1] = b; st.rec[slot][2] = c; st.rec[slot][3] = d;      \
      }                                                                                          \
      /* lay the entries' pairs end to end, in slot order: every entry's lane writes its own */  \
      uint32_t m = any ? mask : 0u;                                                              \
      uint32_t suf = __popc(m);                      /* pairs of this and the later lanes */     \
      _Pragma("unroll") for (int o = 1; o < 32; o <<= 1) {                                       \
        const uint32_t t = __shfl_down_sync(full, suf, o);                                       \
        if (lane + o < 32) suf += t;                                                             \
      }                                                                                          \
      uint32_t w = pending + suf - __popc(m);                                                    \
      const uint32_t tag = (uint32_t)slot << 5;                                                  \
      for (int it = (int)__reduce_max_sync(full, (unsigned)__popc(m)); it > 0; it--) {           \
        if (m) {                                                                                 \
          st.queue[w++] = (uint16_t)(tag | (uint32_t)(__ffs(m) - 1));                            \
          m &= m - 1u;                                                                           \
        }                                                                                        \
      }                                                                                          \
      pending += __shfl_sync(full, suf, 0);                                                      \
      nent += ns;                                                                                \
    }                                                                                            \
    chunk--;                                                                                     \
  }

  while (true) {
    // ---- gather blended entries, back to front, until they hold enough pairs ----
    while (chunk >= 0 && pending < GATHER_PAIRS && nent + 32 <= ENT_CAP) {
      switch (rot) {
        case 0: GATHER_STEP(s0, s1, s2, s3); break;
        case 1: GATHER_STEP(s1, s2, s3, s0); break;
        case 2: GATHER_STEP(s2, s3, s0, s1); break;
        default: GATHER_STEP(s3, s0, s1, s2); break;
      }
      rot = (rot + 1) & 3;
    }
    if (pending == 0) {
      if (chunk < 0) break;
      continue;
    }
    __syncwarp();
    STAGE_CLOCK(c_gather)

    // ---- composite the gathered pairs, 32 at a time ----
    for (uint32_t base = 0; base < pending; base += 32) {
      const uint32_t g = base + lane;
      const bool valid = g < pending;
      // P1: pair g of the queue: intersection, alpha, depth
      const uint32_t item = valid ? st.queue[g] : (uint32_t)lane;
      const int j = item >> 5, pixel = item & 31;
      const SurfelRec s = unpack_rec(st.rec[j][0], st.rec[j][1], st.rec[j][2], st.rec[j][3]);
      PixelRay ray;
      ray.X = st.ray[0][pixel];  ray.Y = st.ray[1][pixel];  ray.Z = st.ray[2][pixel];
      ray.ax = st.ray[3][pixel]; ray.ay = st.ray[4][pixel];
      ray.ex = st.ray[5][pixel]; ray.ey = st.ray[6][pixel]; ray.ez = st.ray[7][pixel];
      ray.pxf = st.ray[8][pixel]; ray.pyf = st.ray[9][pixel];
      Hit h;
      // the forward blended this pair, so the same arithmetic passes the same tests
      const bool ok = intersect_all(ray, s, h) && valid;
      st.pa[lane] = h.alpha;
      st.pd[lane] = h.depth;
      st.pe[lane] = (uint32_t)j;
      st.pmask[lane] = 0u;
      __syncwarp();
      if (ok) atomicOr(&st.pmask[pixel], 1u << lane);
      __syncwarp();
      STAGE_CLOCK(c_p1)
      // P2: every pixel's lane runs its recurrences over its pairs of the batch, in list order
      {
        uint32_t m = st.pmask[lane];
        while (__any_sync(full, m != 0u)) {
          if (m) {
            const int q = __ffs(m) - 1;
            m &= m - 1u;
            const uint32_t e = st.pe[q];
            const float4 c = st.rec[e][2];
            const StepOut o = recur(bs, pc, st.pa[q], st.pd[q], c.y, c.z, c.w, st.ent_pos[e]);
            st.pout[q] = make_float4(o.T, o.dL_dalpha, o.dLd_depth, o.w);
          }
        }
      }
      __syncwarp();
      STAGE_CLOCK(c_p2)
      // P3: the pair's contribution to its surfel's 13 values
      {
        float v[13];
#pragma unroll
        for (int i = 0; i < 13; i++) v[i] = 0.0f;
        if (ok) {
          const float4 o = st.pout[lane];
          const float T = o.x, dL_dalpha = o.y, dLd_depth = o.z, w = o.w;
          const float alpha = h.alpha, depth = h.depth;
          v[9] = w * st.gN[0][pixel];
          v[10] = w * st.gN[1][pixel];
          v[11] = w * st.gN[2][pixel];
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
        STAGE_CLOCK(c_p3)
        // The batch's pairs of one entry sit in consecutive lanes: a segmented reduction by
        // shuffles (13 independent chains, no shared-memory round trip) leaves every entry's
        // sums in the first lane of its segment, which issues the 13 atomics.
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const int jd = __shfl_down_sync(full, j, d);
          const bool take = lane + d < 32 && jd == j;
#pragma unroll
          for (int i = 0; i < 13; i++) {
            const float o = __shfl_down_sync(full, v[i], d);
            if (take) v[i] += o;
          }
        }
        const int jp = __shfl_up_sync(full, j, 1);
        if (valid && (lane == 0 || jp != j)) {
          const size_t gid = st.ent_id[j];
          float *tm = dL_dtransMat + gid * 12, *nr = dL_dnormal + gid * 3;
          // (unconditional: the sums are zero only in the first six of a 2D-filter entry, and
          // thirteen compare-and-branch pairs cost more than those few atomics)
#pragma unroll
          for (int i = 0; i < 9; i++) atomicAdd(tm + (i / 3) * 4 + (i % 3), v[i]);
#pragma unroll
          for (int i = 0; i < 3; i++) atomicAdd(nr + i, v[9 + i]);
          atomicAdd(dL_dopacity + gid, v[12]);
        }
      }
      __syncwarp();
      STAGE_CLOCK(c_red)
    }
    nent = 0;
    pending = 0;
    if (chunk < 0) break;
  }
#ifdef RENDER_TRACE
  if (lane == 0 && blockIdx.x < 8192) {
    g_btrace[blockIdx.x][1] = btime();
    g_bstage[blockIdx.x][0] = c_gather; g_bstage[blockIdx.x][1] = c_p1; g_bstage[blockIdx.x][2] = c_p2;
    g_bstage[blockIdx.x][3] = c_p3;     g_bstage[blockIdx.x][4] = c_red;
  }
#endif
}

#undef GATHER_STEP
#undef LOAD_ENTRY
#undef LOAD_WORDS

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
  grid_dependency_wait();            // (see launch_pdl, common.cuh)
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
  static unsigned long long seen = 0;
  if (first_use_on_device(seen))   // one-warp CTAs with ~10 KB each: residency needs the carve-out
    cudaFuncSetAttribute(render_backward_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                         cudaSharedmemCarveoutMaxShared);
  // (launched the plain way: scheduled while the fill blocks of the work-order kernel still run,
  // its CTAs - numbered by decreasing cost for a round-robin placement on an empty machine -
  // landed unevenly and the kernel took 189 instead of 170 us)
  render_backward_kernel<<<im.tiles * UNITS_PER_TILE, 32, 0, LS(s)>>>(im.ranges, im.bwd_order, im.blk_top, reinterpret_cast<const char *>(b.hdr), b.hdr, g.rec,
      p.projmatrix, p.W, p.H, im.tiles_x, (uint32_t)p.P, im.final_T, im.n_contrib, p.dL_dout, p.dL_dtransMat, p.dL_dnormal, p.dL_dopacity);
}

void launch_backward_preprocess(const BwdParams &p, cudaStream_t s) {
  if (p.push)
    launch_pdl(backward_preprocess_kernel<true>, dim3((p.P + 255) / 256), dim3(256), 0, s,
        p.P, p.radii, p.scales, p.rotations, p.viewmatrix, p.dL_dtransMat, p.dL_dnormal,
        p.dL_dopacity, p.dL_dmeans3D, p.dL_dscales, p.dL_drotations, p.packed,
        p.packed_accumulate, *p.push);
  else
    launch_pdl(backward_preprocess_kernel<false>, dim3((p.P + 255) / 256), dim3(256), 0, s,
        p.P, p.radii, p.scales, p.rotations, p.viewmatrix, p.dL_dtransMat, p.dL_dnormal,
        p.dL_dopacity, p.dL_dmeans3D, p.dL_dscales, p.dL_drotations, p.packed,
        p.packed_accumulate, XchgView());
}
