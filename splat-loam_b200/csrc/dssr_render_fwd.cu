// dssr_render_fwd.cu — forward tile compositing.
//
// Replaces renderCUDA<3>, submodules/diff-surfel-spherical-rasterization/
// cuda_rasterizer/forward.cu:172-351 (launch :353-362).
//
// The reference runs one 256-thread CTA per 16x16 tile; every pixel walks the tile's whole
// depth-sorted list, evaluates every entry and the CTA leaves only when all 256 pixels are
// done.  Here the unit of work is a WARP that owns an 8x4 pixel block (two 4x4 sub-tiles) and
// works through its own list of the tile's entries, in three stages that each keep the 32 lanes busy:
//   * gather: 32 entries of the BLOCK's list per step (block_list_kernel, radix_sort.cu: the
//     tile's entries whose 16-bit sub-tile mask touches the block, a few per cent of the tile's
//     list); they are staged in shared memory - 64-byte record and 16-byte cull box, 128-bit
//     loads, L2-resident - and every pixel tests itself against each staged entry's cull box;
//     the (pixel, entry) pairs that pass are appended to a queue, in list order;
//   * evaluate: the queue is consumed 32 pairs at a time, ONE PAIR PER LANE whatever its pixel
//     and entry: ray-splat intersection, alpha, depth (forward.cu:249-301);
//   * blend: every pixel's own lane blends its pairs of the batch in list order
//     (forward.cu:302-330), with early termination; the warp stops when its 32 pixels are done.
// Skipping a pair whose pixel lies outside the entry's cull box is exact: the box bounds where
// alpha >= 1/255 is possible, so such a pair is one the reference `continue`s over without
// touching any state.  `contributor` keeps counting list positions, so n_contrib /
// median_contributor are the reference's.
// The kernel also writes down what it blended - per 32-entry step of the block's list one word
// (which entries) and per blended entry the 32-bit mask of the block's pixels - which is all
// the backward needs to know (dssr_render_bwd.cu).
#include "raster_math.cuh"

#ifdef RENDER_TRACE
__device__ unsigned long long g_rstage[8192][4];   // cycles in gather / evaluate / blend / write-down
extern "C" __attribute__((visibility("default"))) int splat_debug_render_stages(unsigned long long *out) {
  return (int)cudaMemcpyFromSymbol(out, g_rstage, sizeof(g_rstage));
}
#define STAGE_CLOCK(acc) { const long long c_ = clock64(); acc += c_ - t_stage; t_stage = c_; }
__device__ unsigned long long g_rtrace[8192][4];
extern "C" __attribute__((visibility("default"))) int splat_debug_render_trace(unsigned long long *out) {
  return (int)cudaMemcpyFromSymbol(out, g_rtrace, sizeof(g_rtrace));
}
__device__ __forceinline__ unsigned long long rtime() {
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
constexpr int ENT_CAP = COMPOSITE_ENT_CAP;        // entries staged per gather phase
constexpr int STEP_CAP = 16;       // 32-entry steps per gather phase
// (measured at C2, forward / backward us: 32 pairs 231 / 265, 64: 219 / 252, 96: 213 / 244, 128: 210 / 242;
//  more staged entries than 64 cost residency: ENT_CAP 128 puts the forward in two waves, 295 us)
#ifndef COMPOSITE_GATHER_PAIRS
#define COMPOSITE_GATHER_PAIRS 128
#endif
constexpr int GATHER_PAIRS = COMPOSITE_GATHER_PAIRS;   // a phase is evaluated once it holds this many pairs
constexpr int QUEUE_CAP = GATHER_PAIRS + 32 * 32;   // one more step can add 32 entries x 32 pixels

struct PixelState {
  float T, D, N0, N1, N2, M1, M2, dist, median_depth;
  uint32_t last_contributor, median_contributor;
  bool done;
};

// forward.cu:302-330 for one surviving (pixel, surfel) pair; false: the pixel terminated instead
__device__ __forceinline__ bool blend(PixelState &p, float alpha, float depth, float nx, float ny,
                                      float nz, uint32_t contributor) {
  const float test_T = FM(p.T, FS(1.0f, alpha));
  if (test_T < 0.0001f) {
    p.done = true;
    return false;
  }
  const float w = FM(p.T, alpha);
  const float A = FS(1.0f, p.T);
  const float mm = FM(depth, depth), m2 = FA(depth, depth);
  float tmp = FF(A, mm, p.M2);
  tmp = FF(-p.M1, m2, tmp);
  p.dist = FF(w, tmp, p.dist);
  p.M2 = FF(w, mm, p.M2);
  p.M1 = FF(depth, w, p.M1);
  p.D = FF(depth, w, p.D);
  if (p.T > 0.5f) {
    p.median_depth = depth;
    p.median_contributor = contributor;
  }
  p.N0 = FF(nx, w, p.N0);
  p.N1 = FF(ny, w, p.N1);
  p.N2 = FF(nz, w, p.N2);
  p.T = test_T;
  p.last_contributor = contributor;
  return true;
}

// everything a warp stages in shared memory (one warp per CTA)
struct __align__(16) FwdStage {
  float4 rec[ENT_CAP][4];        // the staged entries' 64-byte records
  uint32_t ent_pos[ENT_CAP];     // the reference's `contributor` (1-based list position)
  uint32_t ent_mask[ENT_CAP];    // pixels that blended the entry (bit = lane of the pixel)
  uint32_t word[STEP_CAP];       // the phase's step words
  float ray[10][32];             // per pixel: X Y Z ax ay ex ey ez pxf pyf
  float pa[64], pd[64];          // evaluate -> blend: alpha, depth of the two batches' pairs
  uint32_t pe[64];               // evaluate -> blend: entry slot of the pair
  uint32_t pmask[64];            // per batch and pixel: which lanes of the batch hold its pairs
  uint16_t queue[QUEUE_CAP];     // (slot << 5) | pixel, in list order
};

// One WARP per CTA = one 8x4 pixel block.  CTAs are numbered in order of decreasing tile list
// length (tile_order), so the hardware's round-robin placement hands every SM a balanced mix
// of long and short lists.
__global__ void __launch_bounds__(32, SPLAT_UNIT_SUBS == 2 ? 16 : 24)
render_forward_kernel(const uint2 *__restrict__ ranges, const uint32_t *__restrict__ tile_order,
                      const uint2 *__restrict__ blist, const uint32_t *__restrict__ blk_len,
                      const float4 *__restrict__ rec,
                      const float4 *__restrict__ cull, const float *__restrict__ projmatrix, int W,
                      int H, int tiles_x, uint32_t capacity, uint32_t P,
                      float *__restrict__ final_T, uint32_t *__restrict__ n_contrib,
                      float *__restrict__ out, uint32_t *__restrict__ unit_cost,
                      uint32_t *__restrict__ blk_top, uint32_t *__restrict__ bwords,
                      uint32_t *__restrict__ bpix) {
  grid_dependency_wait();            // (see launch_pdl, common.cuh)
  __shared__ FwdStage st;
  const int lane = threadIdx.x, half = lane >> 4, p16 = lane & 15;
  const uint32_t full = 0xffffffffu;
  const uint32_t work = blockIdx.x;
#ifdef RENDER_TRACE
  const unsigned long long t_start = rtime();
  uint32_t n_steps = 0, n_pairs = 0;
  long long t_stage = clock64(), c_gather = 0, c_eval = 0, c_blend = 0, c_write = 0;
#endif
  const uint32_t tile = tile_order ? tile_order[work / UNITS_PER_TILE] : work / UNITS_PER_TILE;
  const int unit = work % UNITS_PER_TILE;
  const int sub_a = SPLAT_UNIT_SUBS * unit;                                   // left sub-tile; right = sub_a + 1
  const int x0 = (tile % tiles_x) * SPLAT_TILE + (sub_a & 3) * SUB;
  const int y0 = (tile / tiles_x) * SPLAT_TILE + (sub_a >> 2) * SUB;
  if (x0 >= W || y0 >= H) {
    if (lane == 0) unit_cost[tile * UNITS_PER_TILE + unit] = blk_top[tile * UNITS_PER_TILE + unit] = 0u;
    return;
  }
  const int px = x0 + SUB * half + (p16 & 3), py = y0 + (p16 >> 2);
  const bool inside = half < SPLAT_UNIT_SUBS && px < W && py < H;
  const float fx = projmatrix[0], fy = projmatrix[5], cx = projmatrix[8], cy = projmatrix[9];
  const float pxf = (float)px, pyf = (float)py;
  {
    const PixelRay r = make_ray(pxf, pyf, fx, fy, cx, cy);
    st.ray[0][lane] = r.X;  st.ray[1][lane] = r.Y;  st.ray[2][lane] = r.Z;
    st.ray[3][lane] = r.ax; st.ray[4][lane] = r.ay;
    st.ray[5][lane] = r.ex; st.ray[6][lane] = r.ey; st.ray[7][lane] = r.ez;
    st.ray[8][lane] = r.pxf; st.ray[9][lane] = r.pyf;
  }
  // the cull boxes are only meaningful while |elevation| < pi/2 on every image row (the last
  // sort pass sets every mask bit otherwise, radix_sort.cu)
  const float period = 6.2831853f * fabsf(fx);
  const float inv_period = 1.0f / period;
  const bool cull_on =
      fmaxf(fabsf((0.0f - cy) / fy), fabsf(((float)(H - 1) - cy) / fy)) < 1.55f;

  uint2 range = ranges[tile];
  range.x = min(range.x, capacity);   // only bites when the capacity hint was too small
  range.y = min(range.y, capacity);
  // the block's list (see block_list_kernel) and, indexed like it, what the backward reads
  const size_t ubase = (size_t)UNITS_PER_TILE * range.x + (size_t)unit * (range.y - range.x);
  const uint32_t len = min(blk_len[tile * UNITS_PER_TILE + unit], range.y - range.x);
  const uint2 *bl = blist + ubase;
  uint32_t *bw = bwords + (ubase >> 5) + tile * UNITS_PER_TILE + unit;
  uint32_t *bp = bpix + ubase;

  PixelState ps;
  ps.T = 1.0f;
  ps.D = ps.N0 = ps.N1 = ps.N2 = ps.M1 = ps.M2 = ps.dist = ps.median_depth = 0.0f;
  ps.last_contributor = ps.median_contributor = 0;
  ps.done = !inside;
  uint32_t npairs = 0;

  const uint32_t nsteps = (len + 31) >> 5;
  uint32_t step = 0;
  // Software pipeline of the walk: the list entries (surfel index, list position) are fetched
  // three steps ahead of their use, the records and cull boxes of a step's entries one step
  // ahead (into registers): a record's address comes from an index loaded two steps earlier,
  // so every L2 latency sits under the tests of the steps in between and under the evaluation.
  uint2 cent = make_uint2(0u, 0u), nent_r = cent, n2ent = cent;
  if (lane < len) cent = bl[lane];
  if (32 + lane < len) nent_r = bl[32 + lane];
  if (64 + lane < len) n2ent = bl[64 + lane];
  float4 ra, rb, rc, rd, rcb;
  ra = rb = rc = rd = rcb = make_float4(0.f, 0.f, 0.f, 0.f);
  if (lane < len) {
    // (id < P only fails on the garbage a too-small capacity hint leaves behind; that frame is
    // re-run with the exact size)
    const uint32_t id = min(cent.x, P - 1u);
    const float4 *r = rec + 4 * (size_t)id;
    ra = __ldg(r); rb = __ldg(r + 1); rc = __ldg(r + 2); rd = __ldg(r + 3);
    rcb = __ldg(cull + id);
  }
  __syncwarp();

  while (step < nsteps && !__all_sync(full, ps.done)) {
    // ---- gather: stage the step's entries, queue the candidate pairs ----
    int nent = 0, gsteps = 0;
    uint32_t pending = 0;
    const uint32_t step0 = step;
    while (step < nsteps && pending < GATHER_PAIRS && nent + 32 <= ENT_CAP && gsteps < STEP_CAP) {
      const bool hit = (step << 5) + lane < len;
      const int ns = (int)min(32u, len - (step << 5));       // the hits are the low lanes
      const float4 a = ra, b = rb, c = rc, d = rd;
      float4 cb = rcb;
      const uint32_t pos = cent.y;
      // advance the pipeline: this step's prefetched registers are consumed below
      cent = nent_r;
      nent_r = n2ent;
      {
        const uint32_t q = ((step + 3) << 5) + lane;
        if (q < len) n2ent = bl[q];
      }
      if (((step + 1) << 5) + lane < len) {
        const uint32_t id = min(cent.x, P - 1u);
        const float4 *r = rec + 4 * (size_t)id;
        ra = __ldg(r); rb = __ldg(r + 1); rc = __ldg(r + 2); rd = __ldg(r + 3);
        rcb = __ldg(cull + id);
      }
      if (lane == 0) st.word[gsteps] = 0u;
      const uint32_t not_done = ~__ballot_sync(full, ps.done);
      uint32_t cm = 0u;
      if (hit) {
        const int slot = nent + lane;
        st.rec[slot][0] = a; st.rec[slot][1] = b; st.rec[slot][2] = c; st.rec[slot][3] = d;
        // x is periodic: bring the box into the first period in which it can reach this block
        // (the smallest shift that puts its right edge at or beyond the block's left edge),
        // once per entry; a pixel of the block is then inside the box in that period or - a
        // box almost a period wide - beyond the left edge of the next one.  A box without
        // culling information (+-inf, or not meaningful at all) accepts every pixel.
        if (!cull_on) cb = make_float4(-3.0e38f, 3.0e38f, -3.0e38f, 3.0e38f);
        if (cb.x > -3.0e38f && cb.y < 3.0e38f) {
          const float k = ceilf(((float)x0 - cb.y) * inv_period) * period;
          cb.x += k;
          cb.y += k;
        }
        // the quarter-pixel margin of the box (dssr_preprocess.cu:cull_box) dwarfs the
        // rounding of the shift (< 1e-3 pixel on images up to 8192 wide)
        // The entry's own lane works out which of the block's 32 pixels lie in the box - the
        // 4 rows and 8 columns against the box, the comparisons written so that a NaN in the
        // box answers "candidate" - instead of every pixel testing itself against every entry
        // (32 x 32 tests with a ballot each: a fifth of the kernel's instructions).
        uint32_t rows = 0u, cols = 0u;
#pragma unroll
        for (int r = 0; r < SUB; r++) {
          const float yf = (float)(y0 + r);
          if (!(yf < cb.z) && !(yf > cb.w)) rows |= 1u << r;
        }
#pragma unroll
        for (int q = 0; q < SPLAT_UNIT_SUBS * SUB; q++) {
          const float xf = (float)(x0 + q);
          if ((!(xf < cb.x) && !(xf > cb.y)) || !(xf < cb.x + period)) cols |= 1u << q;
        }
        // lane of pixel (row r, column q) = 16 * (q / 4) + 4 * r + q % 4
        const uint32_t row0 = (cols & 0xfu) | ((cols >> 4) << 16);
        cm = ((rows & 1u) ? row0 : 0u) | ((rows & 2u) ? row0 << 4 : 0u) |
             ((rows & 4u) ? row0 << 8 : 0u) | ((rows & 8u) ? row0 << 12 : 0u);
        cm &= not_done;
        st.ent_pos[slot] = pos + 1u;                       // the reference's `contributor`
        st.ent_mask[slot] = 0u;
      }
      // where each entry's pairs start in the queue: entries in list order, pixels in lane order
      uint32_t inc = __popc(cm);
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(full, inc, o);
        if (lane >= o) inc += t;
      }
      // the entry's lane writes the entry's pairs, pixel by pixel
      {
        uint32_t m = cm, w = pending + inc - __popc(cm);
        const uint32_t tag = (uint32_t)(nent + lane) << 5;
        for (int it = (int)__reduce_max_sync(full, (unsigned)__popc(cm)); it > 0; it--) {
          if (m) {
            st.queue[w++] = (uint16_t)(tag | (uint32_t)(__ffs(m) - 1));
            m &= m - 1u;
          }
        }
      }
      pending += __shfl_sync(full, inc, 31);
      nent += ns;
      step++;
      gsteps++;
    }
#ifdef RENDER_TRACE
    n_steps += gsteps;
    n_pairs += pending;
#endif
    __syncwarp();
    STAGE_CLOCK(c_gather)

    // ---- evaluate and blend the queued pairs, 64 at a time ----
    // Two pairs per lane (g and g + 32), evaluated by the same straight-line code side by side:
    // the two intersection chains are independent, which doubles the instruction-level
    // parallelism of the stage that is bound by the latency of its dependent FP32 chain.
    for (uint32_t base = 0; base < pending; base += 64) {
      bool ok[2];
      float alpha[2], depth[2];
      int slot[2], pixel[2];
#pragma unroll
      for (int u = 0; u < 2; u++) {
        const uint32_t g = base + 32 * u + lane;
        const bool valid = g < pending;
        const uint32_t item = valid ? st.queue[g] : (uint32_t)lane;
        slot[u] = item >> 5;
        pixel[u] = item & 31;
        const SurfelRec s = unpack_rec(st.rec[slot[u]][0], st.rec[slot[u]][1], st.rec[slot[u]][2],
                                       st.rec[slot[u]][3]);
        PixelRay ray;
        ray.X = st.ray[0][pixel[u]];  ray.Y = st.ray[1][pixel[u]];  ray.Z = st.ray[2][pixel[u]];
        ray.ax = st.ray[3][pixel[u]]; ray.ay = st.ray[4][pixel[u]];
        ray.ex = st.ray[5][pixel[u]]; ray.ey = st.ray[6][pixel[u]]; ray.ez = st.ray[7][pixel[u]];
        ray.pxf = st.ray[8][pixel[u]]; ray.pyf = st.ray[9][pixel[u]];
        Hit h;
        // (branch-free: in a warp of 32 unrelated pairs an early exit is never taken by all)
        ok[u] = intersect_all(ray, s, h) && valid;
        alpha[u] = h.alpha;
        depth[u] = h.depth;
      }
#pragma unroll
      for (int u = 0; u < 2; u++) {
        st.pa[32 * u + lane] = alpha[u];
        st.pd[32 * u + lane] = depth[u];
        st.pe[32 * u + lane] = (uint32_t)slot[u];
        st.pmask[32 * u + lane] = 0u;
      }
      __syncwarp();
#pragma unroll
      for (int u = 0; u < 2; u++)
        if (ok[u]) atomicOr(&st.pmask[32 * u + pixel[u]], 1u << lane);
      __syncwarp();
      STAGE_CLOCK(c_eval)
      // every pixel's own lane blends its pairs of the two batches, in list order
#pragma unroll
      for (int u = 0; u < 2; u++) {
        uint32_t m = st.pmask[32 * u + lane];
        while (__any_sync(full, m != 0u && !ps.done)) {
          if (m != 0u && !ps.done) {
            const int q = 32 * u + __ffs(m) - 1;
            m &= m - 1u;
            const uint32_t e = st.pe[q];
            const float4 c = st.rec[e][2];
            // book-keeping for the backward: this pixel blended entry e
            if (blend(ps, st.pa[q], st.pd[q], c.y, c.z, c.w, st.ent_pos[e])) {
              atomicOr(&st.ent_mask[e], 1u << lane);
              npairs++;
            }
          }
        }
      }
      __syncwarp();
      STAGE_CLOCK(c_blend)
      if (__all_sync(full, ps.done)) break;
    }

    // ---- write down what was blended in this phase's steps ----
    // (entry e of the phase is entry (step0 << 5) + e of the block's list: only a list's last
    // step can be short)
    for (int e = lane; e < nent; e += 32) {
      const uint32_t m = st.ent_mask[e];
      if (m) {
        bp[(step0 << 5) + e] = m;
        atomicOr(&st.word[e >> 5], 1u << (e & 31));
      }
    }
    __syncwarp();
    if (lane < gsteps) bw[step0 + lane] = st.word[lane];
    __syncwarp();
    STAGE_CLOCK(c_write)
  }

#ifdef RENDER_TRACE
  if (lane == 0 && work < 8192) {
    g_rtrace[work][0] = t_start;
    g_rtrace[work][1] = rtime();
    g_rtrace[work][2] = ((unsigned long long)n_steps << 32) | n_pairs;
    g_rtrace[work][3] = range.y - range.x;
    g_rstage[work][0] = c_gather; g_rstage[work][1] = c_eval;
    g_rstage[work][2] = c_blend;  g_rstage[work][3] = c_write;
  }
#endif
  // what the backward will have to do for this block: the number of blended (pixel, entry)
  // pairs (its warps are numbered by decreasing cost, see bwd_order_kernel)
  npairs = __reduce_add_sync(full, npairs);
  if (lane == 0) {
    unit_cost[tile * UNITS_PER_TILE + unit] = npairs;
    blk_top[tile * UNITS_PER_TILE + unit] = step;     // steps of the block's list that were walked
  }
  if (inside) {
    const size_t HW = (size_t)W * H;
    const size_t pix = (size_t)py * W + px;
    final_T[pix] = ps.T;
    final_T[pix + HW] = ps.M1;
    final_T[pix + 2 * HW] = ps.M2;
    n_contrib[pix] = ps.last_contributor;
    n_contrib[pix + HW] = ps.median_contributor;
    out[pix] = ps.D;                      // DEPTH_OFFSET
    out[pix + HW] = FS(1.0f, ps.T);       // ALPHA_OFFSET
    out[pix + 2 * HW] = ps.N0;            // NORMAL_OFFSET..+2
    out[pix + 3 * HW] = ps.N1;
    out[pix + 4 * HW] = ps.N2;
    out[pix + 5 * HW] = ps.median_depth;  // MIDDEPTH_OFFSET
    out[pix + 6 * HW] = ps.dist;          // DISTORTION_OFFSET
  }
}

// bwd_order[rank] = backward work unit (the forward's 8x4 pixel blocks), by decreasing cost =
// blended pairs (counting sort on cost / 32, one CTA).  On large images the backward's warps run in more than one wave; longest-first
// numbering packs the waves.
// Blocks 1.. of the same launch zero the backward's accumulation buffers (`fill`: four float
// arrays) meanwhile: one launch instead of fills and a kernel one after the other.
constexpr int COST_BINS = 2048;
__global__ void __launch_bounds__(1024)
bwd_order_kernel(const uint32_t *__restrict__ sub_cost, int units, uint32_t *__restrict__ order,
                 FillList fill) {
  if (blockIdx.x > 0) {
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    const size_t t0 = (size_t)(blockIdx.x - 1) * 1024 + threadIdx.x;
    const size_t stride = (size_t)(gridDim.x - 1) * 1024;
#pragma unroll
    for (int a = 0; a < 4; a++) {
      float *p = fill.ptr[a];
      const size_t n = fill.n[a];
      // scalar head up to the first 16-byte boundary, float4 body, scalar tail
      size_t head = ((16u - ((uintptr_t)p & 15u)) & 15u) / 4u;
      if (head > n) head = n;
      const size_t body = (n - head) / 4;
      float4 *p4 = reinterpret_cast<float4 *>(p + head);
      for (size_t k = t0; k < body; k += stride) p4[k] = z;
      if (t0 < head) p[t0] = 0.f;
      const size_t tail0 = head + 4 * body;
      if (t0 < n - tail0) p[tail0 + t0] = 0.f;
    }
    return;
  }
  __shared__ uint32_t s_bin[COST_BINS];
  __shared__ uint32_t s_warp[32];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  for (int b = tid; b < COST_BINS; b += 1024) s_bin[b] = 0u;
  __syncthreads();
  for (int u = tid; u < units; u += 1024)
    atomicAdd(&s_bin[COST_BINS - 1 - min(sub_cost[u] >> 5, (uint32_t)COST_BINS - 1u)], 1u);
  __syncthreads();
  // exclusive scan, two bins per thread
  const uint32_t v0 = s_bin[2 * tid], v1 = s_bin[2 * tid + 1];
  uint32_t inc = v0 + v1;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  if (lane == 31) s_warp[wid] = inc;
  __syncthreads();
  uint32_t before = 0;
  for (int w = 0; w < wid; w++) before += s_warp[w];
  __syncthreads();
  s_bin[2 * tid] = before + inc - v0 - v1;
  s_bin[2 * tid + 1] = before + inc - v1;
  __syncthreads();
  for (int u = tid; u < units; u += 1024)
    order[atomicAdd(&s_bin[COST_BINS - 1 - min(sub_cost[u] >> 5, (uint32_t)COST_BINS - 1u)], 1u)] =
        (uint32_t)u;
}

}  // namespace

void launch_bwd_order(const ImageView &im, const FillList &fill, cudaStream_t s) {
  bwd_order_kernel<<<1 + 2 * splat_sm_count(), 1024, 0, LS(s)>>>(
      im.sub_cost, im.tiles * UNITS_PER_TILE, im.bwd_order, fill);
}

void launch_render_forward(const FwdParams &p, const GeomView &g, const ImageView &im,
                           const BinView &b, uint32_t capacity, cudaStream_t s) {
  const uint32_t *order = im.tiles <= SPLAT_MAX_ORDERED_TILES ? im.tile_order : nullptr;
  static unsigned long long seen = 0;
  if (first_use_on_device(seen))   // one-warp CTAs: residency is limited by the carve-out
    cudaFuncSetAttribute(render_forward_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                         cudaSharedmemCarveoutMaxShared);
  launch_pdl(render_forward_kernel, dim3(im.tiles * UNITS_PER_TILE), dim3(32), 0, s,
      im.ranges, order, b.blist, im.blk_len, g.rec, g.cull, p.projmatrix, p.W, p.H, im.tiles_x,
      capacity, (uint32_t)p.P, im.final_T, im.n_contrib, p.out_maps, im.sub_cost, im.blk_top,
      b.bwords, b.bpix);
}
