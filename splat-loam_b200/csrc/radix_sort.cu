// radix_sort.cu — the sort of the instance words and what is derived from it: the per-tile
// ranges, point_list and the per-block lists the compositing kernels walk.
//
// Replaces (submodules/diff-surfel-spherical-rasterization/cuda_rasterizer/):
//   cub::DeviceRadixSort::SortPairs(keys u64, values u32, N, 0, 32+bit)
//                                           rasterizer_impl.cu:232-238
//   cudaMemset(ranges) + identifyTileRanges rasterizer_impl.cu:240-248, :94-114
//
// What is sorted.  The reference sorts (key = tile<<32 | float_bits(depth), value =
// surfel idx) pairs stably on the low 32+bit key bits.  Emission order is ascending in
// surfel idx, so "stable by emission order" == "ties broken by idx" (a surfel that
// emits the same tile twice emits identical pairs).  We therefore sort ONE 64-bit word
//   [ tile | depth_bits - bits(0.1f) : 27 | idx ]
// stably on its [tile|depth27] bits only (visible depths lie in (0.1, 120], so 27 bits cover
// them).  Two paths, same result - the reference's point_list / point_list_keys, bit for bit:
//   * tile-first (<= 1024 tiles, <= 2 M surfels; second half of this file): one stable
//     partition of the words by tile (the tile histogram comes from the emit kernel), then
//     one CTA per tile sorts the tile's range on the 27 depth bits in shared memory;
//   * global (anything else): stable LSD radix sort, one kernel per digit ("onesweep":
//     chained scan with decoupled look-back) - 3 passes x 9 bits for the depth and one or two
//     passes for the tile, 8 B/instance each instead of CUB's 6 passes of 12 B/instance -
//     after a digit-base scan that also yields the per-tile ranges.
// The pass kernels also serve simple-knn's Morton sort (sknn.cu).
#include <cstdlib>
#include <cstring>

#include "raster_math.cuh"

namespace {

constexpr uint32_t FLAG_LOCAL = 1u << 30;
constexpr uint32_t FLAG_INCL = 2u << 30;
constexpr uint32_t VAL_MASK = (1u << 30) - 1u;
constexpr int WARPS = SORT_THREADS / 32;

__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_volatile_u32(uint32_t *p, uint32_t v) {
  asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Exclusive scan of each pass's global digit histogram -> digit_base.  When the tile
// id is a single digit its histogram IS the per-tile instance count, so the same scan
// writes ranges[tile] = [start, end) (and (0,0) for empty tiles, which is what the
// reference's memset + identifyTileRanges leaves).
__global__ void __launch_bounds__(SORT_MAX_BINS)
digit_scan_kernel(const uint32_t *__restrict__ hist, uint32_t *__restrict__ digit_base,
                  uint2 *ranges, uint32_t *tile_order, int tiles, int tile_pass_single,
                  int nrepl) {
  __shared__ uint32_t s_warp[32];
  __shared__ uint32_t s_len[SORT_MAX_BINS];
  const int pass = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  uint32_t v = 0;
#pragma unroll 8
  for (int r = 0; r < nrepl; r++)
    v += hist[(size_t)r * (SORT_MAX_PASSES * SORT_MAX_BINS) + pass * SORT_MAX_BINS + tid];
  uint32_t inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  if (lane == 31) s_warp[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    uint32_t w = s_warp[lane], winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t n = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += n;
    }
    s_warp[lane] = winc - w;
  }
  __syncthreads();
  uint32_t excl = s_warp[wid] + inc - v;
  digit_base[pass * SORT_MAX_BINS + tid] = excl;
  if (pass != tile_pass_single) return;
  if (tid < tiles) ranges[tid] = v ? make_uint2(excl, excl + v) : make_uint2(0u, 0u);
  // tile_order[rank] = tile, by decreasing list length (see tile_order_kernel)
  s_len[tid] = v;
  __syncthreads();
  if (tid < tiles && tile_order) {
    uint32_t rank = 0;
    for (int o = 0; o < tiles; o++) {
      const uint32_t l = s_len[o];
      rank += (l > v) || (l == v && o < tid);
    }
    tile_order[rank] = (uint32_t)tid;
  }
}

// identifyTileRanges (rasterizer_impl.cu:94-114) for the > 1024 tiles case.
// When the capacity hint was too small (*num_ptr > capacity) the passes above dropped words
// and `keys` has holes of stale bytes: nothing is written then (ranges stay (0,0), the
// compositing of this attempt is a no-op) and the frame is re-run with the exact size.  A tile
// id beyond the grid can then never index `ranges`; the bound test stays as a second guard.
__global__ void identify_ranges_kernel(const uint32_t *num_ptr, uint32_t capacity,
                                       const unsigned long long *__restrict__ keys,
                                       uint2 *ranges, uint32_t tiles) {
  if (*num_ptr > capacity) return;
  uint32_t L = *num_ptr;
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= L) return;
  uint32_t cur = (uint32_t)(keys[i] >> 32);
  if (cur >= tiles) return;
  if (i == 0) ranges[cur].x = 0;
  else {
    uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
    if (cur != prev) {
      if (prev < tiles) ranges[prev].y = i;
      ranges[cur].x = i;
    }
  }
  if (i == L - 1) ranges[cur].y = L;
}

// tile_order[rank] = tile, rank by decreasing list length (ties by index): the compositing
// kernels number their CTAs through it so that long and short lists are spread evenly.
__global__ void tile_order_kernel(const uint2 *__restrict__ ranges, int tiles,
                                  uint32_t *tile_order) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= tiles) return;
  const uint2 me = ranges[t];
  const uint32_t len = me.y - me.x;
  uint32_t rank = 0;
  for (int o = 0; o < tiles; o++) {
    const uint2 r = ranges[o];
    const uint32_t l = r.y - r.x;
    rank += (l > len) || (l == len && o < t);
  }
  tile_order[rank] = (uint32_t)t;
}

// One radix pass.  Each CTA takes a ticket (so CTA b only ever waits on CTAs that
// have already started), ranks its 512 x ITEMS words stably per digit with warp match-any,
// publishes its digit counts, resolves its global offsets by looking back over the
// predecessors' published counts, and scatters.
template <int ITEMS>
__global__ void __launch_bounds__(SORT_THREADS)
sort_pass_kernel(const unsigned long long *__restrict__ in, unsigned long long *out,
                 const uint32_t *__restrict__ num_ptr, uint32_t capacity, uint32_t *status,
                 uint32_t *ticket, const uint32_t *__restrict__ digit_base, int shift,
                 int bits) {
  extern __shared__ uint32_t smem[];
  const int bins = 1 << bits;
  uint16_t *s_cnt = reinterpret_cast<uint16_t *>(smem);            // [WARPS][bins]
  uint32_t *s_base = smem + (WARPS * bins) / 2;                    // [bins]
  uint32_t *s_tbl = s_base + bins;                                 // [WARPS][bins]
  __shared__ uint32_t s_bid;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const uint32_t N = min(*num_ptr, capacity);
  if (tid == 0) s_bid = atomicAdd(ticket, 1u);
  for (int i = tid; i < (WARPS * bins) / 2; i += SORT_THREADS) smem[i] = 0u;
  for (int i = tid; i < WARPS * bins; i += SORT_THREADS) s_tbl[i] = 0u;
  __syncthreads();
  const uint32_t bid = s_bid;
  const uint32_t start = bid * (SORT_THREADS * ITEMS);
  if (start >= N) return;

  const uint32_t mask = (uint32_t)bins - 1u;
  const uint32_t lt = (1u << lane) - 1u;
  unsigned long long key[ITEMS];
  uint16_t rank[ITEMS];
  uint16_t *my_cnt = s_cnt + wid * bins;
  uint32_t *my_tbl = s_tbl + wid * bins;
  const uint32_t wstart = start + wid * (ITEMS * 32);
#pragma unroll
  for (int r = 0; r < ITEMS; r++) {
    uint32_t i = wstart + r * 32 + lane;
    key[r] = (i < N) ? in[i] : ~0ull;
  }
  // The lanes of a warp that hold the same digit find each other through a per-warp table in
  // shared memory (atomicOr of the lane bit, read back).  MATCH.ANY does it in one instruction
  // but was measured at ~60 cycles per warp and SM on B200: 3x the time of this sequence.
#pragma unroll
  for (int r = 0; r < ITEMS; r++) {
    const uint32_t i = wstart + r * 32 + lane;
    const bool valid = i < N;
    const uint32_t digit = (uint32_t)(key[r] >> shift) & mask;
    if (valid) atomicOr(&my_tbl[digit], 1u << lane);
    __syncwarp();
    uint32_t peers = 0u, old = 0u;
    if (valid) {
      peers = my_tbl[digit];
      old = my_cnt[digit];
    }
    __syncwarp();
    if (valid && (peers & lt) == 0u) {         // the group's first lane
      my_cnt[digit] = (uint16_t)(old + __popc(peers));
      my_tbl[digit] = 0u;
    }
    rank[r] = (uint16_t)(old + __popc(peers & lt));
    __syncwarp();
  }
  __syncthreads();

  // per digit: exclusive prefix over the warps, CTA total, publish, look back
  uint32_t *my_status = status + (size_t)bid * bins;
  for (int d = tid; d < bins; d += SORT_THREADS) {
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < WARPS; w++) {
      uint32_t c = s_cnt[w * bins + d];
      s_cnt[w * bins + d] = (uint16_t)run;
      run += c;
    }
    s_base[d] = run;  // CTA total for now
    st_volatile_u32(&my_status[d], (bid == 0 ? FLAG_INCL : FLAG_LOCAL) | run);
  }
  for (int d = tid; d < bins; d += SORT_THREADS) {
    uint32_t total = s_base[d];
    uint32_t excl = 0;
    if (bid > 0) {
      int p = (int)bid - 1;
      while (true) {
        uint32_t sv;
        do {
          sv = ld_volatile_u32(&status[(size_t)p * bins + d]);
        } while ((sv >> 30) == 0u);
        excl += sv & VAL_MASK;
        if ((sv >> 30) == 2u) break;
        p--;
      }
      st_volatile_u32(&my_status[d], FLAG_INCL | (excl + total));
    }
    s_base[d] = digit_base[d] + excl;
  }
  __syncthreads();

#pragma unroll
  for (int r = 0; r < ITEMS; r++) {
    uint32_t i = wstart + r * 32 + lane;
    if (i < N) {
      uint32_t digit = (uint32_t)(key[r] >> shift) & mask;
      uint32_t pos = s_base[digit] + my_cnt[digit] + rank[r];
      if (pos >= capacity) continue;  // only when the capacity hint was too small
      out[pos] = key[r];
    }
  }
}

// Developer facility (build with -DSORT_TRACE, read with tests/dev_sort_trace.py): wall-clock
// timestamps of the phases of every CTA of the last pass.
#ifdef SORT_TRACE
__device__ unsigned long long g_trace[1024][16];
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define TRACE(slot) if (threadIdx.x == 0 && bid < 1024) g_trace[bid][slot] = gtime();
#else
#define TRACE(slot)
#endif

// The final pass of the rasterizer's sort (the tile digit): same chained scan, but
//   * ranking by ballots - on tile digits MATCH.ANY was measured ~3x slower than on the
//     depth digits (its cost depends on the group structure of the operands);
//   * the CTA's words are first put in CTA-sorted order in shared memory and written out
//     from there, so consecutive threads store consecutive addresses: the direct scatter
//     of this pass ran at a fifth of the store rate of the depth passes;
//   * the unpacking to the reference's point_list_keys / point_list and the 16-bit
//     sub-tile mask of every instance (previously a kernel of its own that re-read the keys)
//     are done on the way out.
// UNPACK = false: the same pass as a plain stable partition - the words go out as words
// (keys_out is then the destination word array); used to split the instances by tile BEFORE
// their depth is sorted (tile_sort_kernel below).
template <int ITEMS, bool UNPACK>
__global__ void __launch_bounds__(SORT_THREADS)
sort_last_pass_kernel(const unsigned long long *__restrict__ in,
                      unsigned long long *__restrict__ keys_out, uint32_t *__restrict__ plist_out,
                      MaskParams mp, uint32_t P, const uint32_t *__restrict__ num_ptr,
                      uint32_t capacity, uint32_t *status, uint32_t *ticket,
                      const uint32_t *__restrict__ digit_base, int shift, int bits, int idx_bits,
                      const uint16_t *__restrict__ mask_in, const uint32_t *__restrict__ hist,
                      uint2 *ranges, uint32_t *tile_order, int tiles) {
  grid_dependency_wait();            // (see launch_pdl, common.cuh)
  grid_launch_dependents();
  extern __shared__ uint32_t smem[];
  const int bins = 1 << bits;
  uint16_t *s_cnt = reinterpret_cast<uint16_t *>(smem);            // [WARPS][bins]
  uint32_t *s_base = smem + (WARPS * bins) / 2;                    // [bins] CTA total, then delta
  uint32_t *s_loc = s_base + bins;                                 // [bins] CTA-local exclusive
  unsigned long long *s_stage =
      reinterpret_cast<unsigned long long *>(s_loc + bins);        // [(SORT_THREADS * ITEMS)]
  // the ranking's per-warp peer table [WARPS][bins] lies over the staging area (used before
  // it; the launch sizes the area for the larger of the two)
  uint32_t *s_tbl = reinterpret_cast<uint32_t *>(s_stage);
  // partition mode: the instances' sub-tile masks travel with the words
  uint16_t *s_mask = reinterpret_cast<uint16_t *>(
      reinterpret_cast<char *>(s_stage) +
      (SORT_THREADS * ITEMS * 8 > WARPS * bins * 4 ? SORT_THREADS * ITEMS * 8 : WARPS * bins * 4));
  // partition mode: every CTA scans the tile histogram itself (digit_base of its digits, the
  // tile ranges), which saves the digit-scan kernel in front of this one
  uint32_t *s_dbase = reinterpret_cast<uint32_t *>(s_mask + SORT_THREADS * ITEMS);   // [bins]
  uint32_t *s_len = s_dbase + bins;                                                  // [bins]
  __shared__ uint32_t s_bid;
  __shared__ uint32_t s_warp[WARPS];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const uint32_t N = min(*num_ptr, capacity);
  if (tid == 0) s_bid = atomicAdd(ticket, 1u);
  for (int i = tid; i < (WARPS * bins) / 2; i += SORT_THREADS) smem[i] = 0u;
  for (int i = tid; i < WARPS * bins; i += SORT_THREADS) s_tbl[i] = 0u;
  if (!UNPACK) {
    const int ipt = (bins + SORT_THREADS - 1) / SORT_THREADS;     // 1 or 2
    const int d0 = tid * ipt;
    uint32_t v0 = 0u, v1 = 0u;
    if (d0 < bins) {
#pragma unroll 8
      for (int r = 0; r < HIST_REPL; r++) v0 += hist[(size_t)r * (SORT_MAX_PASSES * SORT_MAX_BINS) + d0];
    }
    if (ipt > 1 && d0 + 1 < bins) {
#pragma unroll 8
      for (int r = 0; r < HIST_REPL; r++) v1 += hist[(size_t)r * (SORT_MAX_PASSES * SORT_MAX_BINS) + d0 + 1];
    }
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
    const uint32_t excl = before + inc - (v0 + v1);
    if (d0 < bins) { s_dbase[d0] = excl; s_len[d0] = v0; }
    if (ipt > 1 && d0 + 1 < bins) { s_dbase[d0 + 1] = excl + v0; s_len[d0 + 1] = v1; }
  }
  __syncthreads();
  const uint32_t bid = s_bid;
  // ranges[tile] = [start, end) ((0,0) for an empty tile: what the reference's memset +
  // identifyTileRanges leaves) and the work order, tile_order[rank] = tile by decreasing list
  // length - written by the first CTA when it is done with its words
  auto write_ranges = [&]() {
    for (int d = tid; d < tiles; d += SORT_THREADS) {
      const uint32_t v = s_len[d], b0 = s_dbase[d];
      ranges[d] = v ? make_uint2(b0, b0 + v) : make_uint2(0u, 0u);
      if (tile_order) {
        uint32_t rank = 0;
        for (int o = 0; o < tiles; o++) {
          const uint32_t l = s_len[o];
          rank += (l > v) || (l == v && o < d);
        }
        tile_order[rank] = (uint32_t)d;
      }
    }
  };
  const uint32_t start = bid * (SORT_THREADS * ITEMS);
  if (start >= N) {
    if (!UNPACK && bid == 0) write_ranges();
    return;
  }
  const uint32_t count = min((uint32_t)(SORT_THREADS * ITEMS), N - start);
  TRACE(0)

  const uint32_t mask = (uint32_t)bins - 1u;
  const uint32_t lt = (1u << lane) - 1u;
  unsigned long long key[ITEMS];
  uint16_t rank[ITEMS];
  uint16_t *my_cnt = s_cnt + wid * bins;
  uint32_t *my_tbl = s_tbl + wid * bins;
  const uint32_t wstart = start + wid * (ITEMS * 32);
#pragma unroll
  for (int r = 0; r < ITEMS; r++) {
    uint32_t i = wstart + r * 32 + lane;
    key[r] = (i < N) ? in[i] : ~0ull;
  }
  uint16_t msk[UNPACK ? 1 : ITEMS];
  if (!UNPACK) {
#pragma unroll
    for (int r = 0; r < ITEMS; r++) {
      const uint32_t i = wstart + r * 32 + lane;
      msk[r] = (i < N) ? mask_in[i] : (uint16_t)0;
    }
  }
#ifdef SORT_TRACE
  {
    unsigned long long acc = 0;
#pragma unroll
    for (int r = 0; r < ITEMS; r++) acc ^= key[r];
    if (acc == 0xdeadbeefdeadbeefull) g_trace[0][15] = acc;
    __syncthreads();
    TRACE(8)
  }
#endif
#pragma unroll
  for (int r = 0; r < ITEMS; r++) {
    const uint32_t i = wstart + r * 32 + lane;
    const bool valid = i < N;
    const uint32_t digit = (uint32_t)(key[r] >> shift) & mask;
    if (valid) atomicOr(&my_tbl[digit], 1u << lane);
    __syncwarp();
    uint32_t peers = 0u, old = 0u;
    if (valid) {
      peers = my_tbl[digit];
      old = my_cnt[digit];
    }
    __syncwarp();
    if (valid && (peers & lt) == 0u) {         // the group's first lane
      my_cnt[digit] = (uint16_t)(old + __popc(peers));
      my_tbl[digit] = 0u;
    }
    rank[r] = (uint16_t)(old + __popc(peers & lt));
    __syncwarp();
  }
  __syncthreads();
  TRACE(1)

  // per digit: exclusive prefix over the warps, CTA total, publish
  uint32_t *my_status = status + (size_t)bid * bins;
  for (int d = tid; d < bins; d += SORT_THREADS) {
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < WARPS; w++) {
      uint32_t c = s_cnt[w * bins + d];
      s_cnt[w * bins + d] = (uint16_t)run;
      run += c;
    }
    s_base[d] = run;
    st_volatile_u32(&my_status[d], (bid == 0 ? FLAG_INCL : FLAG_LOCAL) | run);
  }
  __syncthreads();
  TRACE(2)
  // CTA-local exclusive scan of the totals over the digits (thread t owns bins [t*ipt, ..))
  {
    const int ipt = (bins + SORT_THREADS - 1) / SORT_THREADS;     // 1 or 2
    const int d0 = tid * ipt;
    const uint32_t v0 = d0 < bins ? s_base[d0] : 0u;
    const uint32_t v1 = (ipt > 1 && d0 + 1 < bins) ? s_base[d0 + 1] : 0u;
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
    const uint32_t excl = before + inc - (v0 + v1);
    if (d0 < bins) s_loc[d0] = excl;
    if (ipt > 1 && d0 + 1 < bins) s_loc[d0 + 1] = excl + v0;
  }
  __syncthreads();
  TRACE(3)
  // stage in CTA-sorted order (needs no global information: overlaps the look-back)
#pragma unroll
  for (int r = 0; r < ITEMS; r++) {
    const uint32_t i = wstart + r * 32 + lane;
    if (i < N) {
      const uint32_t digit = (uint32_t)(key[r] >> shift) & mask;
      const uint32_t slot = s_loc[digit] + my_cnt[digit] + rank[r];
      s_stage[slot] = key[r];
      if (!UNPACK) s_mask[slot] = msk[r];
    }
  }
  TRACE(4)
  // look back; s_base[d] becomes (global position of the CTA's first word of digit d) - s_loc[d]
  for (int d = tid; d < bins; d += SORT_THREADS) {
    const uint32_t total = s_base[d];
    uint32_t excl = 0;
    if (bid > 0) {
      int p = (int)bid - 1;
      while (true) {
        uint32_t sv;
        do {
          sv = ld_volatile_u32(&status[(size_t)p * bins + d]);
        } while ((sv >> 30) == 0u);
        excl += sv & VAL_MASK;
        if ((sv >> 30) == 2u) break;
        p--;
      }
      st_volatile_u32(&my_status[d], FLAG_INCL | (excl + total));
    }
    s_base[d] = (UNPACK ? digit_base[d] : s_dbase[d]) + excl - s_loc[d];
  }
  TRACE(5)
  __syncthreads();
  TRACE(6)

  if (!UNPACK) {
    // the tile is the word's range from here on: its bits make room for the sub-tile mask
    const unsigned long long low = (1ull << 48) - 1ull;
    for (uint32_t j = tid; j < count; j += SORT_THREADS) {
      const unsigned long long k = s_stage[j];
      const uint32_t pos = j + s_base[(uint32_t)(k >> shift) & mask];
      if (pos < capacity) keys_out[pos] = ((unsigned long long)s_mask[j] << 48) | (k & low);
    }
    if (bid == 0) write_ranges();
    return;
  }
  const unsigned long long idx_mask = (1ull << idx_bits) - 1ull;
  bool cull_on = false;
  float inv_period = 0.0f;
  if (mp.sub_mask) {
    const float fx = mp.projmatrix[0], fy = mp.projmatrix[5], cy = mp.projmatrix[9];
    inv_period = 1.0f / (6.2831853f * fabsf(fx));
    // culling assumes |elevation| < pi/2 on every image row
    cull_on = fmaxf(fabsf((0.0f - cy) / fy), fabsf(((float)(mp.H - 1) - cy) / fy)) < 1.55f;
  }
  // four words per thread and step, their cull boxes fetched together: the gathers are
  // independent L2 round trips, not a chain
  constexpr int WU = 4;
  for (uint32_t j0 = tid; j0 < count; j0 += WU * SORT_THREADS) {
    unsigned long long k[WU];
    float4 cb[WU];
#pragma unroll
    for (int u = 0; u < WU; u++) {
      const uint32_t j = j0 + u * SORT_THREADS;
      k[u] = j < count ? s_stage[j] : 0ull;
    }
    if (cull_on) {
#pragma unroll
      for (int u = 0; u < WU; u++)
        cb[u] = __ldg(&mp.cull[min((uint32_t)(k[u] & idx_mask), P - 1u)]);
    }
#pragma unroll
    for (int u = 0; u < WU; u++) {
      const uint32_t j = j0 + u * SORT_THREADS;
      if (j >= count) continue;
      const uint32_t digit = (uint32_t)(k[u] >> shift) & mask;
      const uint32_t pos = j + s_base[digit];
      if (pos >= capacity) continue;   // only when the capacity hint was too small
      const uint32_t idx = (uint32_t)(k[u] & idx_mask);
      const uint32_t d27 = (uint32_t)(k[u] >> idx_bits) & ((1u << SPLAT_DEPTH_BITS) - 1u);
      const uint32_t tile = (uint32_t)(k[u] >> (idx_bits + SPLAT_DEPTH_BITS));
      keys_out[pos] =
          ((unsigned long long)tile << 32) | (unsigned long long)(d27 + SPLAT_DEPTH_BASE);
      plist_out[pos] = idx;
      if (mp.sub_mask) {
        uint32_t sm = 0xffffu;
        if (cull_on)
          sm = subtile_mask(cb[u], (float)((tile % mp.tiles_x) * SPLAT_TILE),
                            (float)((tile / mp.tiles_x) * SPLAT_TILE), inv_period);
        mp.sub_mask[pos] = (uint16_t)sm;
      }
    }
  }
  TRACE(7)
}


// ---------------------------------------------------------------------------------------------
// Tile-first sort (images of up to 1024 tiles).  The tile digit's histogram gives every tile's
// range before anything is sorted, so the instance words are first split by tile with ONE
// stable partition pass (sort_last_pass_kernel<.., false>) and each tile's range - a few
// thousand words - is then sorted on its 27 depth bits by one CTA in shared memory: three
// 9-bit LSD passes that never leave the SM, instead of three passes over all instances through
// L2 with a chained scan each.  The write-out produces the reference's point_list and the
// tile's 8 block lists; the sub-tile masks it needs ride in the words' top 16 bits (computed by
// the emit kernel, packed by the partition pass: the tile bits are redundant once a word lies
// in its tile's range).  point_list_keys is not written (rebuild_keys_kernel).
// Ranges longer than TS_CAP words are sorted by the same CTA with the same passes in global
// memory (the tile's range of the two word arrays as ping and pong).
constexpr int TS_THREADS = 1024;
constexpr int TS_WARPS = TS_THREADS / 32;
constexpr int TS_BINS = 512;
constexpr int TS_ITEMS = 12;
constexpr int TS_CAP = TS_THREADS * TS_ITEMS;
// words | per-warp peer table (u32; the oversize path's u32 counts) | per-warp counts (u16) | scan
constexpr size_t TS_SMEM = (size_t)TS_CAP * 8 + (size_t)TS_WARPS * TS_BINS * 4 +
                           (size_t)TS_WARPS * TS_BINS * 2 + 3 * TS_BINS * 4;

// counts -> offsets: cnt[w][d] becomes the position of warp w's first word with digit d
// (words with a smaller digit + words with digit d in the warps before w).  1024 threads:
// (digit, half of the warps).
template <typename CT>
__device__ __forceinline__ void ts_scan(CT *cnt, uint32_t *s_half, uint32_t *s_base,
                                        uint32_t *s_warp) {
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int d = tid & (TS_BINS - 1), h = tid >> 9;
  uint32_t c[16], run = 0;
#pragma unroll
  for (int j = 0; j < 16; j++) c[j] = cnt[(16 * h + j) * TS_BINS + d];
#pragma unroll
  for (int j = 0; j < 16; j++) {
    const uint32_t t = c[j];
    c[j] = run;
    run += t;
  }
  s_half[h * TS_BINS + d] = run;
  __syncthreads();
  const uint32_t add = h ? s_half[d] : 0u;
  uint32_t tot = 0, inc = 0;
  if (tid < TS_BINS) {
    tot = s_half[tid] + s_half[TS_BINS + tid];
    inc = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += n;
    }
    if (lane == 31) s_warp[wid] = inc;
  }
  __syncthreads();
  if (tid < TS_BINS) {
    uint32_t before = 0;
    for (int w = 0; w < wid; w++) before += s_warp[w];
    s_base[tid] = before + inc - tot;
  }
  __syncthreads();
  // cnt[w][d] = words with a smaller digit + words with digit d in the warps before w
  const uint32_t base = s_base[d] + add;
#pragma unroll
  for (int j = 0; j < 16; j++) cnt[(16 * h + j) * TS_BINS + d] = (CT)(c[j] + base);
  __syncthreads();
}

__global__ void __launch_bounds__(TS_THREADS, 1)
tile_sort_kernel(unsigned long long *words, unsigned long long *alt,
                 const uint2 *__restrict__ ranges, const uint32_t *__restrict__ tile_order,
                 uint32_t capacity, uint32_t *__restrict__ plist_out, int idx_bits,
                 uint2 *__restrict__ blist, uint32_t *__restrict__ blk_len) {
  grid_dependency_wait();            // (see launch_pdl, common.cuh)
  grid_launch_dependents();
  extern __shared__ unsigned long long ts_smem[];
  unsigned long long *s_buf = ts_smem;                                   // [TS_CAP]
  uint32_t *tbl = reinterpret_cast<uint32_t *>(s_buf + TS_CAP);          // [TS_WARPS][TS_BINS]
  uint16_t *cnt16 = reinterpret_cast<uint16_t *>(tbl + TS_WARPS * TS_BINS);   // [TS_WARPS][TS_BINS]
  uint32_t *s_half = reinterpret_cast<uint32_t *>(cnt16 + TS_WARPS * TS_BINS);   // [2][TS_BINS]
  uint32_t *s_base = s_half + 2 * TS_BINS;                               // [TS_BINS]
  __shared__ uint32_t s_warp[32];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const uint32_t lt = (1u << lane) - 1u;
  const uint32_t tile = tile_order ? tile_order[blockIdx.x] : blockIdx.x;
  uint2 rg = ranges[tile];
  rg.x = min(rg.x, capacity);          // only bites when the capacity hint was too small
  rg.y = min(rg.y, capacity);
  const uint32_t n = rg.y - rg.x;
  if (n == 0) {
    if (tid < SPLAT_BLOCKS_PER_TILE) blk_len[tile * SPLAT_BLOCKS_PER_TILE + tid] = 0u;
    return;
  }
#ifdef SORT_TRACE
  const uint32_t bid = blockIdx.x;
  if (tid == 0 && bid < 1024) g_trace[bid][12] = n;
#endif
  TRACE(0)
  const unsigned long long *fin;       // the sorted words of the tile

  if (n <= (uint32_t)TS_CAP) {
    // ---- the whole range in registers + shared memory ----
    const int items = (int)((n + TS_THREADS - 1) / TS_THREADS);
    const uint32_t wbase = (uint32_t)wid * items * 32;       // the warp's chunk, in list order
    unsigned long long key[TS_ITEMS];
    uint32_t rank[TS_ITEMS];
    const unsigned long long *src = words + rg.x;
#pragma unroll
    for (int r = 0; r < TS_ITEMS; r++) {
      const uint32_t i = wbase + r * 32 + lane;
      key[r] = (r < items && i < n) ? src[i] : ~0ull;
    }
#ifdef SORT_TRACE
    {
      unsigned long long acc = 0;
#pragma unroll
      for (int r = 0; r < TS_ITEMS; r++) acc ^= key[r];
      if (acc == 0xdeadbeefdeadbeefull) g_trace[0][15] = acc;
    }
#endif
    TRACE(1)
    // Ranking: the lanes of a warp that hold the same digit find each other through a
    // per-warp table in shared memory (atomicOr of the lane bit, read back) - MATCH.ANY, which
    // does the same in one instruction, was measured at ~60 cycles per warp and SM here: 7.4 us
    // of every 10 us pass.
    uint32_t *my_tbl = tbl + wid * TS_BINS;
    uint16_t *my_cnt = cnt16 + wid * TS_BINS;
    for (int i = tid; i < TS_WARPS * TS_BINS; i += TS_THREADS) tbl[i] = 0u;
    for (int pass = 0; pass < 3; pass++) {
      const int shift = idx_bits + 9 * pass;
      for (int i = tid; i < TS_WARPS * TS_BINS / 2; i += TS_THREADS)
        reinterpret_cast<uint32_t *>(cnt16)[i] = 0u;
      __syncthreads();
#pragma unroll
      for (int r = 0; r < TS_ITEMS; r++) {
        if (r < items) {
          const uint32_t i = wbase + r * 32 + lane;
          const bool valid = i < n;
          const uint32_t digit = (uint32_t)(key[r] >> shift) & (TS_BINS - 1);
          if (valid) atomicOr(&my_tbl[digit], 1u << lane);
          __syncwarp();
          uint32_t peers = 0u, old = 0u;
          if (valid) {
            peers = my_tbl[digit];
            old = my_cnt[digit];
          }
          __syncwarp();
          if (valid && (peers & lt) == 0u) {       // the group's first lane
            my_cnt[digit] = (uint16_t)(old + __popc(peers));
            my_tbl[digit] = 0u;
          }
          rank[r] = old + __popc(peers & lt);
          __syncwarp();
        }
      }
      __syncthreads();
#ifdef SORT_TRACE
      if (tid == 0 && bid < 1024) g_trace[bid][2 + 3 * pass] = gtime();
#endif
      ts_scan(cnt16, s_half, s_base, s_warp);
#ifdef SORT_TRACE
      if (tid == 0 && bid < 1024) g_trace[bid][3 + 3 * pass] = gtime();
#endif
      // (positions first, stores after: the compiler must assume the tables and the word
      // buffer alias and would otherwise order every table load behind the previous store)
#pragma unroll
      for (int r = 0; r < TS_ITEMS; r++) {
        if (r < items) {
          const uint32_t i = wbase + r * 32 + lane;
          if (i < n) rank[r] += my_cnt[(uint32_t)(key[r] >> shift) & (TS_BINS - 1)];
        }
      }
#pragma unroll
      for (int r = 0; r < TS_ITEMS; r++) {
        if (r < items) {
          const uint32_t i = wbase + r * 32 + lane;
          if (i < n) s_buf[rank[r]] = key[r];
        }
      }
      __syncthreads();
#ifdef SORT_TRACE
      if (tid == 0 && bid < 1024) g_trace[bid][4 + 3 * pass] = gtime();
#endif
      if (pass < 2) {
#pragma unroll
        for (int r = 0; r < TS_ITEMS; r++) {
          const uint32_t i = wbase + r * 32 + lane;
          if (r < items && i < n) key[r] = s_buf[i];
        }
      }
    }
    fin = s_buf;
  } else {
    // ---- oversize range: the same three passes through global memory ----
    uint32_t *cnt = tbl;
    uint32_t *my_cnt = cnt + wid * TS_BINS;
    unsigned long long *a = words + rg.x, *b = alt + rg.x;
    const uint32_t chunk = ((n + TS_WARPS - 1) / TS_WARPS + 31u) & ~31u;
    const uint32_t w0 = (uint32_t)wid * chunk, w1 = min(n, w0 + chunk);
    for (int pass = 0; pass < 3; pass++) {
      const int shift = idx_bits + 9 * pass;
      for (int i = tid; i < TS_WARPS * TS_BINS; i += TS_THREADS) cnt[i] = 0u;
      __syncthreads();
      for (uint32_t i0 = w0; i0 < w1; i0 += 32) {
        const uint32_t i = i0 + lane;
        const bool valid = i < w1;
        const uint32_t digit = valid ? ((uint32_t)(a[i] >> shift) & (TS_BINS - 1)) : 0xffffffffu;
        const uint32_t peers = __match_any_sync(0xffffffffu, digit);
        if (lane == __ffs(peers) - 1 && valid) my_cnt[digit] += __popc(peers);
        __syncwarp();
      }
      __syncthreads();
      ts_scan(cnt, s_half, s_base, s_warp);
      for (uint32_t i0 = w0; i0 < w1; i0 += 32) {
        const uint32_t i = i0 + lane;
        const bool valid = i < w1;
        const unsigned long long k = valid ? a[i] : 0ull;
        const uint32_t digit = valid ? ((uint32_t)(k >> shift) & (TS_BINS - 1)) : 0xffffffffu;
        const uint32_t peers = __match_any_sync(0xffffffffu, digit);
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (lane == leader && valid) {
          old = my_cnt[digit];
          my_cnt[digit] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        if (valid) b[old + __popc(peers & lt)] = k;
        __syncwarp();
      }
      __threadfence_block();
      __syncthreads();
      unsigned long long *t = a;
      a = b;
      b = t;
    }
    fin = a;
  }

  // ---- write-out: the reference's point_list, and the 8 block lists of the tile (see
  // block_list_kernel below; the sub-tile mask rides in the word's top bits) ----
  // (point_list_keys is not materialised on this path: every key is (tile << 32) | bits of
  // depths[point_list[i]], which rebuild_keys_kernel writes out when dssr_debug_state asks)
  const unsigned long long idx_mask = (1ull << idx_bits) - 1ull;
  for (uint32_t j = tid; j < n; j += TS_THREADS) plist_out[rg.x + j] = (uint32_t)(fin[j] & idx_mask);
  TRACE(13)
  // Every thread takes a run of consecutive entries and counts, per block, the entries of its
  // run that touch it; a scan over the threads turns the counts into each run's first position
  // in the 8 lists.  (A version with one warp ballot per block and 32 entries spent 8 us of
  // the CTA's 25 in this phase: VOTE and POPC issue far below the rate of plain arithmetic.)
  uint32_t(*s_bcnt)[SPLAT_BLOCKS_PER_TILE] =
      reinterpret_cast<uint32_t(*)[SPLAT_BLOCKS_PER_TILE]>(s_half);     // [TS_WARPS][8], scan scratch
  const uint32_t ipt = (n + TS_THREADS - 1) / TS_THREADS;
  const uint32_t e0 = min(n, (uint32_t)tid * ipt), e1 = min(n, e0 + ipt);
  uint32_t bcnt[SPLAT_BLOCKS_PER_TILE], binc[SPLAT_BLOCKS_PER_TILE];
#pragma unroll
  for (int u = 0; u < SPLAT_BLOCKS_PER_TILE; u++) bcnt[u] = 0u;
  for (uint32_t i = e0; i < e1; i++) {
    const uint32_t m = (uint32_t)(fin[i] >> 48);
#pragma unroll
    for (int u = 0; u < SPLAT_BLOCKS_PER_TILE; u++) bcnt[u] += ((m >> (SPLAT_UNIT_SUBS * u)) & ((1u << SPLAT_UNIT_SUBS) - 1u)) != 0u;
  }
#pragma unroll
  for (int u = 0; u < SPLAT_BLOCKS_PER_TILE; u++) {
    uint32_t inc = bcnt[u];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    binc[u] = inc;
    if (lane == 31) s_bcnt[wid][u] = inc;
  }
  TRACE(14)
  __syncthreads();
  if (tid < SPLAT_BLOCKS_PER_TILE) {
    uint32_t run = 0;
    for (int w = 0; w < TS_WARPS; w++) {
      const uint32_t c = s_bcnt[w][tid];
      s_bcnt[w][tid] = run;
      run += c;
    }
    blk_len[tile * SPLAT_BLOCKS_PER_TILE + tid] = run;
  }
  __syncthreads();
#pragma unroll
  for (int u = 0; u < SPLAT_BLOCKS_PER_TILE; u++) bcnt[u] = s_bcnt[wid][u] + binc[u] - bcnt[u];
  uint2 *bbase = blist + (size_t)SPLAT_BLOCKS_PER_TILE * rg.x;
  TRACE(15)
  for (uint32_t i = e0; i < e1; i++) {
    const unsigned long long k = fin[i];
    const uint32_t m = (uint32_t)(k >> 48);
    const uint2 e = make_uint2((uint32_t)(k & idx_mask), i);
#pragma unroll
    for (int u = 0; u < SPLAT_BLOCKS_PER_TILE; u++)
      if (((m >> (SPLAT_UNIT_SUBS * u)) & ((1u << SPLAT_UNIT_SUBS) - 1u)) != 0u) bbase[(size_t)u * n + bcnt[u]++] = e;
  }
  TRACE(11)
}


// ---------------------------------------------------------------------------------------------
// Block lists.  The compositing kernels work on 8x4-pixel blocks, 8 per tile; a block is
// touched by a few per cent of its tile's entries (the sub-tile masks say which).  One CTA per
// tile splits the tile's sorted list into the 8 lists of its blocks, in list order, so that a
// block's warp reads only what concerns it instead of scanning the whole tile list:
//   blist[8 * ranges[t].x + u * len(t) + c] = (surfel index, position in the tile's list)
// for the c-th entry whose mask has one of the block's two bits (2u, 2u + 1).
constexpr int BL_THREADS = 1024;
constexpr int BL_WARPS = BL_THREADS / 32;

__global__ void __launch_bounds__(BL_THREADS)
block_list_kernel(const uint2 *__restrict__ ranges, uint32_t capacity,
                  const uint32_t *__restrict__ point_list, const uint16_t *__restrict__ sub_mask,
                  uint2 *__restrict__ blist, uint32_t *__restrict__ blk_len) {
  __shared__ uint32_t s_cnt[BL_WARPS][SPLAT_BLOCKS_PER_TILE];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const uint32_t tile = blockIdx.x;
  uint2 rg = ranges[tile];
  rg.x = min(rg.x, capacity);          // only bites when the capacity hint was too small
  rg.y = min(rg.y, capacity);
  const uint32_t n = rg.y - rg.x;
  if (n == 0) {
    if (tid < SPLAT_BLOCKS_PER_TILE) blk_len[tile * SPLAT_BLOCKS_PER_TILE + tid] = 0u;
    return;
  }
  const uint16_t *sm = sub_mask + rg.x;
  const uint32_t *pl = point_list + rg.x;
  const uint32_t chunk = ((n + BL_WARPS - 1) / BL_WARPS + 31u) & ~31u;   // entries per warp
  const uint32_t w0 = min(n, (uint32_t)wid * chunk), w1 = min(n, w0 + chunk);
  const uint32_t lt = (1u << lane) - 1u;
  // pass 1: how many entries of this warp's part go to each block
  uint32_t cnt[SPLAT_BLOCKS_PER_TILE];
#pragma unroll
  for (int u = 0; u < SPLAT_BLOCKS_PER_TILE; u++) cnt[u] = 0u;
  for (uint32_t i0 = w0; i0 < w1; i0 += 32) {
    const uint32_t i = i0 + lane;
    const uint32_t m = i < w1 ? (uint32_t)sm[i] : 0u;
#pragma unroll
    for (int u = 0; u < SPLAT_BLOCKS_PER_TILE; u++)
      cnt[u] += __popc(__ballot_sync(0xffffffffu, ((m >> (SPLAT_UNIT_SUBS * u)) & ((1u << SPLAT_UNIT_SUBS) - 1u)) != 0u));
  }
  if (lane == 0) {
#pragma unroll
    for (int u = 0; u < SPLAT_BLOCKS_PER_TILE; u++) s_cnt[wid][u] = cnt[u];
  }
  __syncthreads();
  // exclusive prefix over the warps, per block (one thread per block)
  if (tid < SPLAT_BLOCKS_PER_TILE) {
    uint32_t run = 0;
    for (int w = 0; w < BL_WARPS; w++) {
      const uint32_t c = s_cnt[w][tid];
      s_cnt[w][tid] = run;
      run += c;
    }
    blk_len[tile * SPLAT_BLOCKS_PER_TILE + tid] = run;
  }
  __syncthreads();
  // pass 2: write
  uint32_t off[SPLAT_BLOCKS_PER_TILE];
#pragma unroll
  for (int u = 0; u < SPLAT_BLOCKS_PER_TILE; u++) off[u] = s_cnt[wid][u];
  uint2 *base = blist + (size_t)SPLAT_BLOCKS_PER_TILE * rg.x;
  for (uint32_t i0 = w0; i0 < w1; i0 += 32) {
    const uint32_t i = i0 + lane;
    const bool in = i < w1;
    const uint32_t m = in ? (uint32_t)sm[i] : 0u;
    const uint32_t id = in ? pl[i] : 0u;
#pragma unroll
    for (int u = 0; u < SPLAT_BLOCKS_PER_TILE; u++) {
      const bool hit = ((m >> (SPLAT_UNIT_SUBS * u)) & ((1u << SPLAT_UNIT_SUBS) - 1u)) != 0u;
      const uint32_t vote = __ballot_sync(0xffffffffu, hit);
      if (hit) base[(size_t)u * n + off[u] + __popc(vote & lt)] = make_uint2(id, i);
      off[u] += __popc(vote);
    }
  }
}

}  // namespace

#ifdef SORT_TRACE
extern "C" __attribute__((visibility("default"))) int splat_debug_sort_trace(unsigned long long *out) {
  return (int)cudaMemcpyFromSymbol(out, g_trace, sizeof(g_trace));
}
#endif

void launch_digit_scan(const uint32_t *hist, uint32_t *digit_base, int npass, uint2 *ranges,
                       uint32_t *tile_order, int tiles, int tile_pass_single, int nrepl,
                       cudaStream_t s) {
  digit_scan_kernel<<<npass, SORT_MAX_BINS, 0, LS(s)>>>(hist, digit_base, ranges, tile_order, tiles,
                                                    tile_pass_single, nrepl);
}

// Keys per thread of a pass: all CTAs of a pass are resident one per SM, so a pass runs in
// ceil(#CTAs / #SMs) waves of (fixed cost + per-key cost * keys per thread).  215 CTAs of 16
// keys per thread are two waves, the second a third full; 287 CTAs of 12 are two full, shorter
// ones.  `n_estimate`: the expected number of words (the capacity may be a generous hint).
static int pick_items(uint32_t n_estimate) {
  // SPLAT_B200_SORT_ITEMS=8|12|16 overrides the choice (developer A/B switch)
  static const int forced = [] {
    const char *e = getenv("SPLAT_B200_SORT_ITEMS");
    const int v = e ? atoi(e) : 0;
    return (v == 8 || v == 12 || v == 16) ? v : 0;
  }();
  if (forced) return forced;
  const int sms = splat_sm_count();
  const int choices[3] = {16, 12, 8};
  int best = 16;
  long long best_cost = -1;
  for (int c = 0; c < 3; c++) {
    const long long ctas = ((long long)n_estimate + SORT_THREADS * choices[c] - 1) / (SORT_THREADS * choices[c]);
    const long long waves = (ctas + sms - 1) / sms;
    const long long cost = (waves < 1 ? 1 : waves) * (10 + choices[c]);   // fixed part ~ 10 keys
    if (best_cost < 0 || cost < best_cost) {
      best_cost = cost;
      best = choices[c];
    }
  }
  return best;
}

template <int ITEMS>
static void launch_one_pass(int last /* 1: unpacking last pass, 2: staged partition */,
                            const unsigned long long *src, unsigned long long *dst,
                            unsigned long long *keys_out, uint32_t *plist_out,
                            const MaskParams &mp, uint32_t P, const uint32_t *num_ptr,
                            uint32_t capacity, uint32_t *status, uint32_t *ticket,
                            const uint32_t *digit_base, int shift, int bits, int idx_bits,
                            cudaStream_t s, const uint16_t *mask_in = nullptr,
                            const uint32_t *hist = nullptr, uint2 *ranges = nullptr,
                            uint32_t *tile_order = nullptr, int tiles = 0) {
  static unsigned long long seen = 0;
  if (first_use_on_device(seen)) {   // the widest pass (10 bits) needs more than the default 48 KB
    const int big = WARPS * SORT_MAX_BINS * 2 + SORT_MAX_BINS * 4;
    cudaFuncSetAttribute(sort_pass_kernel<ITEMS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         big + WARPS * SORT_MAX_BINS * 4);
    const int stage_max = SORT_THREADS * ITEMS * 8 > WARPS * SORT_MAX_BINS * 4
                              ? SORT_THREADS * ITEMS * 8 : WARPS * SORT_MAX_BINS * 4;
    cudaFuncSetAttribute(sort_last_pass_kernel<ITEMS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         big + SORT_MAX_BINS * 4 + stage_max);
    cudaFuncSetAttribute(sort_last_pass_kernel<ITEMS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         big + SORT_MAX_BINS * 4 + stage_max + SORT_THREADS * ITEMS * 2 +
                             2 * SORT_MAX_BINS * 4);
  }
  const int nbins = 1 << bits;
  int nblk = (int)(((size_t)capacity + SORT_THREADS * ITEMS - 1) / (SORT_THREADS * ITEMS));
  if (last == 2 && nblk < 1) nblk = 1;   // the partition pass also writes the tile ranges
  const size_t smem = (size_t)WARPS * nbins * 2 + (size_t)nbins * 4;
  const size_t stage = (size_t)SORT_THREADS * ITEMS * 8 > (size_t)WARPS * nbins * 4
                           ? (size_t)SORT_THREADS * ITEMS * 8 : (size_t)WARPS * nbins * 4;
  if (nblk < 1) return;
  if (last == 1)
    sort_last_pass_kernel<ITEMS, true><<<nblk, SORT_THREADS, smem + (size_t)nbins * 4 + stage, LS(s)>>>(
        src, keys_out, plist_out, mp, P, num_ptr, capacity, status, ticket, digit_base, shift, bits,
        idx_bits, nullptr, nullptr, nullptr, nullptr, 0);
  else if (last == 2)
    launch_pdl(sort_last_pass_kernel<ITEMS, false>, dim3(nblk), dim3(SORT_THREADS),
               smem + (size_t)nbins * 4 + stage + (size_t)SORT_THREADS * ITEMS * 2 + (size_t)nbins * 8,
               s, src, dst, nullptr, mp, P, num_ptr, capacity, status, ticket, digit_base, shift,
               bits, idx_bits, mask_in, hist, ranges, tile_order, tiles);
  else
    sort_pass_kernel<ITEMS><<<nblk, SORT_THREADS, smem + (size_t)WARPS * nbins * 4, LS(s)>>>(src, dst, num_ptr, capacity, status,
                                                             ticket, digit_base, shift, bits);
}

// Runs `npass` passes ping-ponging between a and b (pass 0 reads a).  When
// `unpack_last` the final pass writes keys_out / plist_out (and mp->sub_mask) instead of a
// word array.  Returns the array that holds the sorted words otherwise.  The status arrays
// must hold SORT_MAX_BLOCKS(capacity) entries per digit (the layouts size them so).
unsigned long long *launch_sort_passes(unsigned long long *a, unsigned long long *b,
                                       unsigned long long *keys_out, uint32_t *plist_out,
                                       const uint32_t *num_ptr, uint32_t capacity,
                                       uint32_t *status, const size_t *status_off,
                                       uint32_t *tickets, const uint32_t *digit_base,
                                       int npass, const int *shift, const int *bits,
                                       int unpack_last, int idx_bits, uint32_t n_estimate,
                                       cudaStream_t s, const MaskParams *mp, uint32_t P) {
  unsigned long long *src = a, *dst = b;
  const int items = pick_items(n_estimate ? n_estimate : capacity);
  const MaskParams none{};
  for (int p = 0; p < npass; p++) {
    const int last = (unpack_last && p == npass - 1) ? 1 : 0;
#define ONE_PASS(I)                                                                             \
  launch_one_pass<I>(last, src, dst, keys_out, plist_out, mp ? *mp : none, P, num_ptr, capacity, \
                     status + status_off[p], tickets + p, digit_base + p * SORT_MAX_BINS,        \
                     shift[p], bits[p], idx_bits, s)
    if (items == 16) ONE_PASS(16);
    else if (items == 12) ONE_PASS(12);
    else ONE_PASS(8);
#undef ONE_PASS
    unsigned long long *t = dst;
    dst = src;
    src = t;
  }
  return src;
}

// point_list_keys for dssr_debug_state when the sort did not write them: key i of tile t is
// (t << 32) | bits(depth of surfel point_list[i]) - the word the reference sorts.
__global__ void rebuild_keys_kernel(const uint2 *__restrict__ ranges, uint32_t capacity,
                                    const uint32_t *__restrict__ point_list,
                                    const float *__restrict__ depths, uint32_t P,
                                    unsigned long long *__restrict__ keys) {
  const uint32_t tile = blockIdx.x;
  uint2 rg = ranges[tile];
  rg.x = min(rg.x, capacity);
  rg.y = min(rg.y, capacity);
  for (uint32_t i = rg.x + threadIdx.x; i < rg.y; i += blockDim.x)
    keys[i] = ((unsigned long long)tile << 32) |
              (unsigned long long)__float_as_uint(depths[min(point_list[i], P - 1u)]);
}

void launch_rebuild_keys(const ImageView &im, const GeomView &g, const BinView &b, uint32_t P,
                         uint32_t capacity, cudaStream_t s) {
  rebuild_keys_kernel<<<im.tiles, 256, 0, LS(s)>>>(im.ranges, capacity, b.point_list,
                                                  g.depths, P, b.keys);
}

// Tile-first needs the tile id in one digit (<= 1024 tiles) and 16 free bits above
// [depth27 | idx] for the sub-tile mask (<= 2 M surfels); SPLAT_B200_SORT=global keeps the four
// global passes (developer A/B switch).
bool sort_tile_first(const SortPlan &sp, int tiles) {
  static const bool enabled = [] {
    const char *e = getenv("SPLAT_B200_SORT");
    return !(e && strcmp(e, "global") == 0);
  }();
  return enabled && sp.npass == 4 && tiles <= SPLAT_MAX_ORDERED_TILES &&
         sp.idx_bits + SPLAT_DEPTH_BITS + 16 <= 64;
}

void launch_sort(const FwdParams &p, const GeomView &g, const ImageView &im, const BinView &b,
                 const SortPlan &sp, uint32_t capacity, uint32_t n_estimate, cudaStream_t s) {
  MaskParams mp;
  mp.sub_mask = b.sub_mask;
  mp.cull = g.cull;
  mp.projmatrix = p.projmatrix;
  mp.tiles_x = im.tiles_x;
  mp.H = p.H;
  const int single = (sp.npass == 4) ? 3 : -1;
  if (!sort_tile_first(sp, im.tiles))
    launch_digit_scan(b.hist, b.digit_base, sp.npass, im.ranges, im.tile_order, im.tiles, single,
                      HIST_REPL, s);
  if (sort_tile_first(sp, im.tiles)) {
    static unsigned long long seen = 0;
    if (first_use_on_device(seen))
      cudaFuncSetAttribute(tile_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TS_SMEM);
    const int items = pick_items(n_estimate ? n_estimate : capacity);
    const MaskParams none{};
#define PART_PASS(I)                                                                              \
  launch_one_pass<I>(2, b.words_a, b.words_b, nullptr, nullptr, none, (uint32_t)p.P,             \
                     &g.hdr->num_rendered, capacity, b.status + b.status_off[single],            \
                     b.tickets + single, b.digit_base + single * SORT_MAX_BINS, sp.shift[single], \
                     sp.bits[single], sp.idx_bits, s, reinterpret_cast<const uint16_t *>(b.keys), \
                     b.hist + single * SORT_MAX_BINS, im.ranges, im.tile_order, im.tiles)
    if (items == 16) PART_PASS(16);
    else if (items == 12) PART_PASS(12);
    else PART_PASS(8);
#undef PART_PASS
    launch_pdl(tile_sort_kernel, dim3(im.tiles), dim3(TS_THREADS), TS_SMEM, s, b.words_b,
               b.words_a, im.ranges, im.tile_order, capacity, b.point_list, sp.idx_bits, b.blist,
               im.blk_len);
    return;
  }
  launch_sort_passes(b.words_a, b.words_b, b.keys, b.point_list, &g.hdr->num_rendered,
                     capacity, b.status, b.status_off, b.tickets, b.digit_base, sp.npass,
                     sp.shift, sp.bits, 1, sp.idx_bits, n_estimate, s, &mp, (uint32_t)p.P);
  if (single < 0) {
    cudaMemsetAsync(im.ranges, 0, (size_t)im.tiles * sizeof(uint2), s);
    uint32_t blocks = (capacity + 255) / 256;
    if (blocks)
      identify_ranges_kernel<<<blocks, 256, 0, LS(s)>>>(&g.hdr->num_rendered, capacity, b.keys,
                                                    im.ranges, (uint32_t)im.tiles);
  }
  if (single < 0 && im.tiles <= SPLAT_MAX_ORDERED_TILES)   // beyond that: identity order
    tile_order_kernel<<<(im.tiles + 127) / 128, 128, 0, LS(s)>>>(im.ranges, im.tiles, im.tile_order);
  block_list_kernel<<<im.tiles, BL_THREADS, 0, LS(s)>>>(im.ranges, capacity, b.point_list,
                                                     b.sub_mask, b.blist, im.blk_len);
}
