// dssr_render_fwd_steps.cu — forward tile compositing, step-synchronous variant (kept for A/B).
//
// Replaces renderCUDA<3>, submodules/diff-surfel-spherical-rasterization/
// cuda_rasterizer/forward.cu:172-351 (launch :353-362).
//
// The reference runs one 256-thread CTA per 16x16 tile; every pixel walks the tile's
// whole depth-sorted list and the CTA leaves only when all 256 pixels are done.  On a
// 64x1024 range image that is 256 CTAs on 148 SMs, and a warp's dependent-instruction
// chain is as long as the tile's list.  Here the unit of work is a WARP that owns a 4x4
// pixel sub-tile (4096 warps at 64x1024) and walks its tile's list on its own:
//   * 32 list entries per step: each lane reads its entry's 16-bit sub-tile mask (written
//     by the last sort pass from the surfel's conservative cull box) and a ballot compacts
//     the entries that can touch this sub-tile; their 64-byte records are gathered
//     (128-bit loads, L2-resident) into a warp-private shared-memory ring;
//   * survivors are evaluated TWO AT A TIME: lanes 0-15 intersect the 16 pixels with one
//     surfel, lanes 16-31 with the next; the results are swapped by shuffle and both
//     halves blend both surfels in list order, so every lane carries its pixel's state;
//   * the warp stops as soon as ITS 16 pixels are done.
// Skipping an entry whose cull box misses the sub-tile is exact: such an entry fails the
// reference's alpha >= 1/255 test at every pixel of the sub-tile, i.e. it is one the
// reference `continue`s over without touching any state.  `contributor` keeps counting
// list positions, so n_contrib / median_contributor are the reference's.
#include "raster_math.cuh"

#ifdef RENDER_TRACE
__device__ unsigned long long g_rtrace_steps[8192][4];
extern "C" __attribute__((visibility("default"))) int splat_debug_render_trace_steps(unsigned long long *out) {
  return (int)cudaMemcpyFromSymbol(out, g_rtrace_steps, sizeof(g_rtrace_steps));
}
__device__ __forceinline__ unsigned long long rtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#endif

#ifndef FWD_WARPS_PER_SM
#define FWD_WARPS_PER_SM 28
#endif

namespace {

constexpr int SUB = 4;                                            // 4x4 pixels per warp
constexpr int SUBS_PER_TILE = (SPLAT_TILE / SUB) * (SPLAT_TILE / SUB);   // 16

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

// One WARP per CTA = one 4x4 sub-tile.  CTAs are numbered in order of decreasing tile list
// length (tile_order), so the hardware's round-robin placement hands every SM a balanced
// mix of long and short lists.
__global__ void __launch_bounds__(32, FWD_WARPS_PER_SM)
render_forward_steps_kernel(const uint2 *__restrict__ ranges, const uint32_t *__restrict__ tile_order,
                      const uint32_t *__restrict__ point_list,
                      const uint16_t *__restrict__ sub_mask, const float4 *__restrict__ rec,
                      const float *__restrict__ projmatrix, int W, int H, int tiles_x,
                      uint32_t capacity, uint32_t P, float *__restrict__ final_T,
                      uint32_t *__restrict__ n_contrib, float *__restrict__ out,
                      uint32_t *__restrict__ sub_cost, uint32_t *__restrict__ blend_bits,
                      uint32_t bits_stride, uint16_t *__restrict__ pix_mask, uint32_t pix_stride) {
  // double-buffered staging ring: while the survivors of one 32-entry step are being
  // composited, the records of the next step's survivors are already in flight
  __shared__ float4 s_rec[2][32][4];
  __shared__ float s_cut[2][32];
  __shared__ uint32_t s_pos[2][32];
  const int lane = threadIdx.x, half = lane >> 4, p16 = lane & 15;
  const uint32_t work = blockIdx.x;
#ifdef RENDER_TRACE
  const unsigned long long t_start = rtime();
  uint32_t n_steps = 0, n_pairs = 0;
#endif
  const uint32_t tile = tile_order ? tile_order[work / SUBS_PER_TILE] : work / SUBS_PER_TILE;
  const int sub = work % SUBS_PER_TILE;
  const int x0 = (tile % tiles_x) * SPLAT_TILE + (sub & 3) * SUB;
  const int y0 = (tile / tiles_x) * SPLAT_TILE + (sub >> 2) * SUB;
  if (x0 >= W || y0 >= H) return;
  const int px = x0 + (p16 & 3), py = y0 + (p16 >> 2);
  const bool inside = px < W && py < H;
  const float fx = projmatrix[0], fy = projmatrix[5], cx = projmatrix[8], cy = projmatrix[9];
  const PixelRay ray = make_ray((float)px, (float)py, fx, fy, cx, cy);

  uint2 range = ranges[tile];
  range.x = min(range.x, capacity);   // only bites when the capacity hint was too small
  range.y = min(range.y, capacity);
  const uint32_t lt = (1u << lane) - 1u;
  // The record of what was blended, for the backward: one word per 32-entry step (which of its
  // entries some pixel of this sub-tile blended) and, per such entry, the mask of those pixels.
  // The backward then composites exactly these (pixel, entry) pairs and tests nothing.
  uint32_t *step_word = blend_bits + (size_t)sub * bits_stride + (range.x >> 5) + tile;
  uint16_t *pix = pix_mask + (size_t)sub * pix_stride + range.x;
  uint32_t step_bits = 0u, npairs = 0u;

  PixelState ps;
  ps.T = 1.0f;
  ps.D = ps.N0 = ps.N1 = ps.N2 = ps.M1 = ps.M2 = ps.dist = ps.median_depth = 0.0f;
  ps.last_contributor = ps.median_contributor = 0;
  ps.done = !inside;

  // ---- prologue: stage step 0, fetch the masks of step 1 ----
  uint32_t m;   // survivors of the step about to be composited (staged in buffer `cur`)
  {
    const uint32_t i0 = range.x + lane;
    const bool hit = (i0 < range.y) && ((sub_mask[i0] >> sub) & 1u);
    m = __ballot_sync(0xffffffffu, hit);
    if (hit) {
      const int slot = __popc(m & lt);
      // (id < P only fails on the garbage a too-small capacity hint leaves behind; that
      // frame is re-run with the exact size)
      const uint32_t id = min(point_list[i0], P - 1u);
      const float4 *r = rec + 4 * (size_t)id;
      const float4 a = __ldg(r), b = __ldg(r + 1), c = __ldg(r + 2), d = __ldg(r + 3);
      s_rec[0][slot][0] = a; s_rec[0][slot][1] = b; s_rec[0][slot][2] = c; s_rec[0][slot][3] = d;
      s_cut[0][slot] = rho_cutoff(d.x);
      s_pos[0][slot] = lane + 1;   // the reference's `contributor`
    }
  }
  // masks and surfel indices run two steps ahead, records one step ahead
  uint32_t nmask = 0u, nid = 0u;
  if (range.x + 32 + lane < range.y) {
    nmask = sub_mask[range.x + 32 + lane];
    nid = point_list[range.x + 32 + lane];
  }
  __syncwarp();
  int cur = 0;
  for (uint32_t base = range.x; base < range.y; base += 32, cur ^= 1) {
    if (__all_sync(0xffffffffu, ps.done)) break;
    // ---- issue the loads of the NEXT step (consumed after this step's compositing) ----
    const bool nhit = (nmask >> sub) & 1u;
    const uint32_t nm = __ballot_sync(0xffffffffu, nhit);
    const uint32_t id1 = min(nid, P - 1u);
    {
      const uint32_t ni = base + 64 + lane;
      nmask = 0u;
      if (ni < range.y) {
        nmask = sub_mask[ni];
        nid = point_list[ni];
      }
    }
    float4 na, nb, nc, nd;
    if (nhit) {
      const float4 *r = rec + 4 * (size_t)id1;
      na = __ldg(r); nb = __ldg(r + 1); nc = __ldg(r + 2); nd = __ldg(r + 3);
    }
    // ---- composite this step's survivors, two at a time ----
    const int nsurv = __popc(m);
#ifdef RENDER_TRACE
    n_steps++;
    n_pairs += (nsurv + 1) / 2;
#endif
    for (int pair = 0; pair < nsurv; pair += 2) {
      // lanes 0-15 take survivor `pair`, lanes 16-31 survivor `pair + 1`
      const int slot = pair + half;
      const bool mine = slot < nsurv;
      const int sl = mine ? slot : pair;
      const SurfelRec s = unpack_rec(s_rec[cur][sl][0], s_rec[cur][sl][1], s_rec[cur][sl][2],
                                     s_rec[cur][sl][3]);
      Hit h;
      const bool maybe = mine && !ps.done && intersect_begin(ray, s, s_cut[cur][sl], h);
      // the cull box is only a box: skip the expensive half when no pixel can pass
      if (!__any_sync(0xffffffffu, maybe)) continue;
      const bool ok = maybe && intersect_finish(ray, s, h);
      // swap results between the halves: every lane gets (first, second) for its pixel
      const bool ok_o = __shfl_xor_sync(0xffffffffu, (int)ok, 16) != 0;
      const float alpha_o = __shfl_xor_sync(0xffffffffu, h.alpha, 16);
      const float depth_o = __shfl_xor_sync(0xffffffffu, h.depth, 16);
      const bool ok1 = half ? ok_o : ok, ok2 = half ? ok : ok_o;
      bool b1 = false, b2 = false;
      if (ok1) {
        const float4 c = s_rec[cur][pair][2];
        b1 = blend(ps, half ? alpha_o : h.alpha, half ? depth_o : h.depth, c.y, c.z, c.w,
                   s_pos[cur][pair]);
      }
      if (ok2 && !ps.done) {
        const float4 c = s_rec[cur][pair + 1][2];
        b2 = blend(ps, half ? h.alpha : alpha_o, half ? h.depth : depth_o, c.y, c.z, c.w,
                   s_pos[cur][pair + 1]);
      }
      // (both halves carry the same 16 pixel states: the low half's votes are the pixel mask)
      const uint32_t m1 = __ballot_sync(0xffffffffu, b1) & 0xffffu;
      const uint32_t m2 = __ballot_sync(0xffffffffu, b2) & 0xffffu;
      if (m1) {
        const uint32_t q = s_pos[cur][pair] - 1u;
        if (lane == 0) pix[q] = (uint16_t)m1;
        step_bits |= 1u << (q & 31u);
        npairs += __popc(m1);
      }
      if (m2) {
        const uint32_t q = s_pos[cur][pair + 1] - 1u;
        if (lane == 0) pix[q] = (uint16_t)m2;
        step_bits |= 1u << (q & 31u);
        npairs += __popc(m2);
      }
      if (__all_sync(0xffffffffu, ps.done)) break;
    }
    if (lane == 0) step_word[(base - range.x) >> 5] = step_bits;
    step_bits = 0u;
    // ---- park the next step's records in the other buffer ----
    if (nhit) {
      const int slot = __popc(nm & lt);
      s_rec[cur ^ 1][slot][0] = na; s_rec[cur ^ 1][slot][1] = nb;
      s_rec[cur ^ 1][slot][2] = nc; s_rec[cur ^ 1][slot][3] = nd;
      s_cut[cur ^ 1][slot] = rho_cutoff(nd.x);
      s_pos[cur ^ 1][slot] = base - range.x + 32 + lane + 1;
    }
    m = nm;
    __syncwarp();
  }

#ifdef RENDER_TRACE
  if (lane == 0 && work < 8192) {
    g_rtrace_steps[work][0] = t_start;
    g_rtrace_steps[work][1] = rtime();
    g_rtrace_steps[work][2] = ((unsigned long long)n_steps << 32) | n_pairs;
    g_rtrace_steps[work][3] = range.y - range.x;
  }
#endif
  // what the backward will have to do for this sub-tile: the number of blended (pixel, entry)
  // pairs (its warps are numbered by decreasing cost, see bwd_order_kernel)
  if (lane == 0 && npairs) atomicAdd(&sub_cost[tile * (SUBS_PER_TILE / 2) + (sub >> 1)], npairs);
  if (inside && half == 0) {
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

}  // namespace

void launch_render_forward_steps(const FwdParams &p, const GeomView &g, const ImageView &im,
                           const BinView &b, uint32_t capacity, cudaStream_t s) {
  const uint32_t *order = im.tiles <= SPLAT_MAX_ORDERED_TILES ? im.tile_order : nullptr;
  static unsigned long long seen = 0;
  if (first_use_on_device(seen))   // one-warp CTAs: residency is limited by the carve-out
    cudaFuncSetAttribute(render_forward_steps_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                         cudaSharedmemCarveoutMaxShared);
  // the cost of a backward unit (two sub-tiles) is accumulated by both of its warps
  cudaMemsetAsync(im.sub_cost, 0, (size_t)im.tiles * (SUBS_PER_TILE / 2) * sizeof(uint32_t), s);
  render_forward_steps_kernel<<<im.tiles * SUBS_PER_TILE, 32, 0, LS(s)>>>(
      im.ranges, order, b.point_list, b.sub_mask, g.rec, p.projmatrix, p.W, p.H, im.tiles_x,
      capacity, (uint32_t)p.P, im.final_T, im.n_contrib, p.out_maps, im.sub_cost, b.blend_bits,
      b.bits_stride, b.pix_mask, b.pix_stride);
}
