// radix_sort.cu — hand-written stable LSD radix sort of the instance words, one
// kernel per digit ("onesweep": chained scan with decoupled look-back), plus the
// digit-base scan that also yields the per-tile ranges.
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
// stably on its [tile|depth27] bits only: 3 passes x 9 bits for the depth (visible
// depths lie in (0.1, 120] so 27 bits cover them) and one pass of <= 10 bits for the
// tile (two when there are more than 1024 tiles) — 4 passes of 8 B/instance instead
// of CUB's 6 passes of 12 B/instance.  The last pass unpacks to the reference's
// point_list_keys (u64) and point_list (u32), which are therefore bit-identical.
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
__global__ void identify_ranges_kernel(const uint32_t *num_ptr, uint32_t capacity,
                                       const unsigned long long *__restrict__ keys,
                                       uint2 *ranges) {
  uint32_t L = min(*num_ptr, capacity);
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= L) return;
  uint32_t cur = (uint32_t)(keys[i] >> 32);
  if (i == 0) ranges[cur].x = 0;
  else {
    uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
    if (cur != prev) {
      ranges[prev].y = i;
      ranges[cur].x = i;
    }
  }
  if (i == L - 1) ranges[cur].y = L;
}

// sub_mask[i]: which 4x4 sub-tiles of its tile the i-th sorted instance's cull box can touch
// (consumed by the compositing kernels instead of re-testing boxes in their inner loops).
__global__ void __launch_bounds__(256)
sub_mask_kernel(const uint32_t *__restrict__ num_ptr, uint32_t capacity,
                const unsigned long long *__restrict__ keys,
                const uint32_t *__restrict__ point_list, MaskParams mp, uint32_t P) {
  const uint32_t N = min(*num_ptr, capacity);
  const uint32_t i = blockIdx.x * 256 + threadIdx.x;
  if (i >= N) return;
  const float fx = mp.projmatrix[0], fy = mp.projmatrix[5], cy = mp.projmatrix[9];
  const float inv_period = 1.0f / (6.2831853f * fabsf(fx));
  // culling assumes |elevation| < pi/2 on every image row
  const bool cull_on = fmaxf(fabsf((0.0f - cy) / fy), fabsf(((float)(mp.H - 1) - cy) / fy)) < 1.55f;
  uint32_t sm = 0xffffu;
  if (cull_on) {
    const uint32_t t = (uint32_t)(keys[i] >> 32);
    const uint32_t id = min(point_list[i], P - 1u);
    sm = subtile_mask(mp.cull[id], (float)((t % mp.tiles_x) * SPLAT_TILE),
                      (float)((t / mp.tiles_x) * SPLAT_TILE), inv_period);
  }
  mp.sub_mask[i] = (uint16_t)sm;
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
// have already started), ranks its 4096 words stably per digit with warp match-any,
// publishes its digit counts, resolves its global offsets by looking back over the
// predecessors' published counts, and scatters.
template <bool LAST>
__global__ void __launch_bounds__(SORT_THREADS)
sort_pass_kernel(const unsigned long long *__restrict__ in, unsigned long long *out,
                 unsigned long long *keys_out, uint32_t *plist_out,
                 const uint32_t *__restrict__ num_ptr, uint32_t capacity, uint32_t *status,
                 uint32_t *ticket, const uint32_t *__restrict__ digit_base, int shift,
                 int bits, int idx_bits) {
  extern __shared__ uint32_t smem[];
  const int bins = 1 << bits;
  uint16_t *s_cnt = reinterpret_cast<uint16_t *>(smem);            // [WARPS][bins]
  uint32_t *s_base = smem + (WARPS * bins) / 2;                    // [bins]
  __shared__ uint32_t s_bid;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const uint32_t N = min(*num_ptr, capacity);
  if (tid == 0) s_bid = atomicAdd(ticket, 1u);
  for (int i = tid; i < (WARPS * bins) / 2; i += SORT_THREADS) smem[i] = 0u;
  __syncthreads();
  const uint32_t bid = s_bid;
  const uint32_t start = bid * SORT_BLOCK_ELEMS;
  if (start >= N) return;

  const uint32_t mask = (uint32_t)bins - 1u;
  const uint32_t lt = (1u << lane) - 1u;
  unsigned long long key[SORT_ITEMS];
  uint16_t rank[SORT_ITEMS];
  uint16_t *my_cnt = s_cnt + wid * bins;
  const uint32_t wstart = start + wid * (SORT_ITEMS * 32);
#pragma unroll
  for (int r = 0; r < SORT_ITEMS; r++) {
    uint32_t i = wstart + r * 32 + lane;
    key[r] = (i < N) ? in[i] : ~0ull;
  }
#pragma unroll
  for (int r = 0; r < SORT_ITEMS; r++) {
    uint32_t i = wstart + r * 32 + lane;
    bool valid = i < N;
    uint32_t digit = valid ? ((uint32_t)(key[r] >> shift) & mask) : 0xffffffffu;
    uint32_t peers = __match_any_sync(0xffffffffu, digit);
    int leader = __ffs(peers) - 1;
    uint32_t old = 0;
    if (lane == leader && valid) {
      old = my_cnt[digit];
      my_cnt[digit] = (uint16_t)(old + __popc(peers));
    }
    old = __shfl_sync(0xffffffffu, old, leader);
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

  const unsigned long long idx_mask = (1ull << idx_bits) - 1ull;
#pragma unroll
  for (int r = 0; r < SORT_ITEMS; r++) {
    uint32_t i = wstart + r * 32 + lane;
    if (i < N) {
      uint32_t digit = (uint32_t)(key[r] >> shift) & mask;
      uint32_t pos = s_base[digit] + my_cnt[digit] + rank[r];
      if (pos >= capacity) continue;  // only when the capacity hint was too small
      if (LAST) {
        unsigned long long k = key[r];
        uint32_t idx = (uint32_t)(k & idx_mask);
        uint32_t d27 = (uint32_t)(k >> idx_bits) & ((1u << SPLAT_DEPTH_BITS) - 1u);
        unsigned long long tile = k >> (idx_bits + SPLAT_DEPTH_BITS);
        keys_out[pos] = (tile << 32) | (unsigned long long)(d27 + SPLAT_DEPTH_BASE);
        plist_out[pos] = idx;
      } else {
        out[pos] = key[r];
      }
    }
  }
}

}  // namespace

void launch_digit_scan(const uint32_t *hist, uint32_t *digit_base, int npass, uint2 *ranges,
                       uint32_t *tile_order, int tiles, int tile_pass_single, int nrepl,
                       cudaStream_t s) {
  digit_scan_kernel<<<npass, SORT_MAX_BINS, 0, s>>>(hist, digit_base, ranges, tile_order, tiles,
                                                    tile_pass_single, nrepl);
}

// Runs `npass` passes ping-ponging between a and b (pass 0 reads a).  When
// `unpack_last` the final pass writes keys_out / plist_out instead of a word array.
// Returns the array that holds the sorted words otherwise.
unsigned long long *launch_sort_passes(unsigned long long *a, unsigned long long *b,
                                       unsigned long long *keys_out, uint32_t *plist_out,
                                       const uint32_t *num_ptr, uint32_t capacity,
                                       uint32_t *status, const size_t *status_off,
                                       uint32_t *tickets, const uint32_t *digit_base,
                                       int npass, const int *shift, const int *bits,
                                       int unpack_last, int idx_bits, int nblk,
                                       cudaStream_t s) {
  unsigned long long *src = a, *dst = b;
  static bool attr_set = false;
  if (!attr_set) {   // the widest pass (10 bits) needs more than the default 48 KB
    const int big = WARPS * SORT_MAX_BINS * 2 + SORT_MAX_BINS * 4;
    cudaFuncSetAttribute(sort_pass_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
    cudaFuncSetAttribute(sort_pass_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
    attr_set = true;
  }
  for (int p = 0; p < npass; p++) {
    const int nbins = 1 << bits[p];
    const size_t smem = (size_t)WARPS * nbins * 2 + (size_t)nbins * 4;
    if (unpack_last && p == npass - 1)
      sort_pass_kernel<true><<<nblk, SORT_THREADS, smem, s>>>(
          src, dst, keys_out, plist_out, num_ptr, capacity, status + status_off[p], tickets + p,
          digit_base + p * SORT_MAX_BINS, shift[p], bits[p], idx_bits);
    else
      sort_pass_kernel<false><<<nblk, SORT_THREADS, smem, s>>>(
          src, dst, keys_out, plist_out, num_ptr, capacity, status + status_off[p], tickets + p,
          digit_base + p * SORT_MAX_BINS, shift[p], bits[p], idx_bits);
    unsigned long long *t = dst;
    dst = src;
    src = t;
  }
  return src;
}

void launch_sort(const FwdParams &p, const GeomView &g, const ImageView &im, const BinView &b,
                 const SortPlan &sp, uint32_t capacity, cudaStream_t s) {
  MaskParams mp;
  mp.sub_mask = b.sub_mask;
  mp.cull = g.cull;
  mp.projmatrix = p.projmatrix;
  mp.tiles_x = im.tiles_x;
  mp.H = p.H;
  const int single = (sp.npass == 4) ? 3 : -1;
  launch_digit_scan(b.hist, b.digit_base, sp.npass, im.ranges, im.tile_order, im.tiles, single,
                    HIST_REPL, s);
  launch_sort_passes(b.words_a, b.words_b, b.keys, b.point_list, &g.hdr->num_rendered,
                     capacity, b.status, b.status_off, b.tickets, b.digit_base, sp.npass,
                     sp.shift, sp.bits, 1, sp.idx_bits, b.nblk, s);
  if (single < 0) {
    cudaMemsetAsync(im.ranges, 0, (size_t)im.tiles * sizeof(uint2), s);
    uint32_t blocks = (capacity + 255) / 256;
    if (blocks)
      identify_ranges_kernel<<<blocks, 256, 0, s>>>(&g.hdr->num_rendered, capacity, b.keys,
                                                    im.ranges);
  }
  if (single < 0 && im.tiles <= SPLAT_MAX_ORDERED_TILES)   // beyond that: identity order
    tile_order_kernel<<<(im.tiles + 127) / 128, 128, 0, s>>>(im.ranges, im.tiles, im.tile_order);
  if (capacity)
    sub_mask_kernel<<<(capacity + 255) / 256, 256, 0, s>>>(&g.hdr->num_rendered, capacity, b.keys,
                                                           b.point_list, mp, (uint32_t)p.P);
}
