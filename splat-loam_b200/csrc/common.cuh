// common.cuh — shared declarations for the sm_100a surfel rasterizer kernels.
//
// Buffer layouts (all carved out of the three caller-allocated byte buffers that the
// reference calls geomBuffer / binningBuffer / imgBuffer,
// DSSR/cuda_rasterizer/rasterizer_impl.h:18-56 — our internal layout is our own):
//
//  geometry buffer  [GeomHeader | rec float4[4P] | cull float4[P] | depths f32[P] |
//                    xy float2[P] | tiles_touched u32[P] | point_offsets u32[P] |
//                    chunk_sums u32[nchunk] | chunk_base u32[nchunk]]
//    rec (64 B / surfel, the only thing the compositing kernels gather):
//        f4[0] = Tu.x Tu.y Tu.z Tv.x     f4[1] = Tv.y Tv.z Tw.x Tw.y
//        f4[2] = Tw.z n.x  n.y  n.z      f4[3] = opacity cx cy |Tw|
//    cull (16 B / surfel): conservative pixel-space box {xmin,xmax,ymin,ymax} outside
//        of which the surfel provably fails the alpha >= 1/255 test.
//  image buffer     [final_T f32[3HW] (T, M1, M2) | n_contrib u32[2HW] | ranges uint2[tiles] |
//                    tile_order | sub_cost | bwd_order | blk_len | blk_top]  (ImageLayout)
//  binning buffer   [BinHeader | point_list u32[C] | keys u64[C] | words_a u64[C] |
//                    words_b u64[C] | sub_mask u16[C] | blist uint2[8C] | bpix u32[8C] |
//                    bwords | (keys_unsorted u64[C]) | sort tables]          (BinLayout)
#pragma once
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define SPLAT_TILE 16             // BLOCK_X = BLOCK_Y = 16 (DSSR auxiliary.h:19-20)
#define SPLAT_NEAR 0.1f           // near_n  (auxiliary.h:37)
#define SPLAT_FAR 120.0f          // far_n   (auxiliary.h:38)
#define SPLAT_DEPTH_BASE 0x3DCCCCCDu  // float bits of 0.1f; visible depths are > this
#define SPLAT_DEPTH_BITS 27       // bits(120.0f) - bits(0.1f) < 2^27
#define SPLAT_ALPHA_MIN 0x3B808081u   // smallest float that is >= 1/255

// The compositing kernels' unit of work ("block"): SPLAT_UNIT_SUBS 4x4 sub-tiles side by side -
// 2: 8x4 pixels, 8 blocks per tile, every lane of the warp a pixel; 1: 4x4 pixels, 16 blocks per
// tile, twice the warps with the upper 16 lanes free of pixels (they work in the pair stages)
#ifndef SPLAT_UNIT_SUBS
#define SPLAT_UNIT_SUBS 2
#endif
#define SPLAT_BLOCKS_PER_TILE (16 / SPLAT_UNIT_SUBS)
#define SPLAT_MAX_ORDERED_TILES 8192   // the O(tiles^2) work ordering is skipped beyond this

#define PRE_CHUNK 256             // surfels per preprocess block (one thread each)

#define SORT_THREADS 512
#define SORT_ITEMS 16                                  // keys per thread per pass: 8, 12 or 16
#define SORT_BLOCK_ELEMS (SORT_THREADS * SORT_ITEMS)   // 8192 (largest CTA chunk)
#define SORT_MIN_BLOCK_ELEMS (SORT_THREADS * 8)        // smallest chunk: sizes the status arrays
#define SORT_MAX_BINS 1024
#define SORT_MAX_PASSES 6
#define HIST_REPL 32              // replicas of the global digit histograms

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct GeomHeader {
  uint32_t magic;
  int P, W, H;
  uint32_t num_rendered;   // written by the chunk-scan kernel
  uint32_t status;         // bit0: prefiltered violation
  uint32_t pad[2];
};

struct GeomLayout {
  size_t rec, cull, depths, xy, tiles, offsets, chunk_sums, chunk_base, total;
  int nchunk;
  __host__ explicit GeomLayout(int P) {
    nchunk = (P + PRE_CHUNK - 1) / PRE_CHUNK;
    size_t o = align_up(sizeof(GeomHeader), 256);
    rec = o;        o = align_up(o + (size_t)P * 64, 256);
    cull = o;       o = align_up(o + (size_t)P * 16, 256);
    depths = o;     o = align_up(o + (size_t)P * 4, 256);
    xy = o;         o = align_up(o + (size_t)P * 8, 256);
    tiles = o;      o = align_up(o + (size_t)P * 4, 256);
    offsets = o;    o = align_up(o + (size_t)P * 4, 256);
    chunk_sums = o; o = align_up(o + (size_t)nchunk * 4, 256);
    chunk_base = o; o = align_up(o + (size_t)nchunk * 4, 256);
    total = o + 256;  // slack for aligning the caller's pointer
  }
};

struct ImageLayout {
  size_t final_T, n_contrib, ranges, tile_order, sub_cost, bwd_order, blk_len, blk_top, total;
  int tiles_x, tiles_y, tiles;
  __host__ ImageLayout(int W, int H) {
    tiles_x = (W + SPLAT_TILE - 1) / SPLAT_TILE;
    tiles_y = (H + SPLAT_TILE - 1) / SPLAT_TILE;
    tiles = tiles_x * tiles_y;
    size_t HW = (size_t)W * H;
    size_t o = 0;
    final_T = o;   o = align_up(o + HW * 12, 256);
    n_contrib = o; o = align_up(o + HW * 8, 256);
    ranges = o;    o = align_up(o + (size_t)tiles * 8, 256);
    tile_order = o; o = align_up(o + (size_t)tiles * 4, 256);
    sub_cost = o;  o = align_up(o + (size_t)tiles * 16 * 4, 256);   // per 4x4 sub-tile
    bwd_order = o; o = align_up(o + (size_t)tiles * 16 * 4, 256);
    blk_len = o;   o = align_up(o + (size_t)tiles * SPLAT_BLOCKS_PER_TILE * 4, 256);
    blk_top = o;   o = align_up(o + (size_t)tiles * SPLAT_BLOCKS_PER_TILE * 4, 256);
    total = o + 256;
  }
};

struct SortPlan {
  int idx_bits, tile_bits, npass;
  int shift[SORT_MAX_PASSES], bits[SORT_MAX_PASSES];
  bool ok;
  __host__ SortPlan(int P, int tiles) {
    idx_bits = 1;
    while ((1ll << idx_bits) < (long long)P) idx_bits++;
    tile_bits = 1;
    while ((1 << tile_bits) < tiles) tile_bits++;
    ok = idx_bits + SPLAT_DEPTH_BITS + tile_bits <= 64;
    npass = 0;
    for (int i = 0; i < 3; i++) {  // 27 depth bits = 3 x 9
      shift[npass] = idx_bits + 9 * i;
      bits[npass] = 9;
      npass++;
    }
    int tb = tile_bits, sh = idx_bits + SPLAT_DEPTH_BITS;
    int ntp = (tb + 9) / 10;
    for (int i = 0; i < ntp; i++) {
      int b = (tb + (ntp - i) - 1) / (ntp - i);  // spread evenly, <= 10
      shift[npass] = sh;
      bits[npass] = b;
      sh += b;
      tb -= b;
      npass++;
    }
    if (npass > SORT_MAX_PASSES) ok = false;
  }
};

struct BinHeader {
  uint32_t magic;
  uint32_t capacity;
  uint32_t keep_unsorted;
  // 1: the sort did not materialise point_list_keys (tile-first path); dssr_debug_state
  // rebuilds them from point_list, ranges and the surfels' depths (launch_rebuild_keys)
  uint32_t keys_deferred;
  // byte offset (from the aligned buffer start) of sub_mask: the backward only knows
  // num_rendered, not the capacity the forward laid the buffer out with
  unsigned long long off_sub_mask;
  // Block lists (block_list_kernel, radix_sort.cu): every 8x4 pixel block - the compositing
  // kernels' unit of work, 8 per tile - has its own list of the tile's entries whose sub-tile
  // mask touches it, in list order: blist[] = (surfel index, position in the tile's list).
  // The list of block u of tile t starts at element 8 * ranges[t].x + u * (length of t's list).
  // What the forward compositing leaves for the backward, indexed like blist: bpix[] = the
  // 32-bit mask of the block's pixels that blended the entry; bwords[] = one word per 32
  // entries (bit i: entry i of the step was blended by some pixel), the word of step k of the
  // block at (start >> 5) + 8 * t + u + k.
  unsigned long long off_blist, off_bpix, off_bwords;
};

struct BinLayout {
  size_t words_a, words_b, keys, point_list, sub_mask, blist, bpix, bwords, keys_unsorted, hist,
      digit_base, status, tickets, total;
  size_t status_bytes, hist_bytes;
  size_t status_off[SORT_MAX_PASSES];  // element offset of each pass in `status`
  int nblk;
  __host__ BinLayout(size_t C, int keep_unsorted, const SortPlan &sp, int ntiles) {
    nblk = (int)((C + SORT_MIN_BLOCK_ELEMS - 1) / SORT_MIN_BLOCK_ELEMS);
    if (nblk < 1) nblk = 1;
    size_t o = align_up(sizeof(BinHeader), 256);
    // point_list first: its offset must not depend on the capacity, the backward
    // only knows num_rendered
    point_list = o; o = align_up(o + C * 4, 256);
    keys = o;       o = align_up(o + C * 8, 256);
    words_a = o;    o = align_up(o + C * 8, 256);
    words_b = o;    o = align_up(o + C * 8, 256);
    sub_mask = o;   o = align_up(o + C * 2, 256);
    // (worst case: every entry of a tile touches all 8 blocks)
    blist = o;      o = align_up(o + C * SPLAT_BLOCKS_PER_TILE * 8, 256);
    bpix = o;       o = align_up(o + C * SPLAT_BLOCKS_PER_TILE * 4, 256);
    bwords = o;
    o = align_up(o + ((C * SPLAT_BLOCKS_PER_TILE >> 5) + (size_t)ntiles * SPLAT_BLOCKS_PER_TILE + 2) * 4, 256);
    keys_unsorted = o;
    if (keep_unsorted) o = align_up(o + C * 8, 256);
    // [hist | tickets | status] are zeroed with ONE memset per forward
    hist = o;
    hist_bytes = (size_t)HIST_REPL * SORT_MAX_PASSES * SORT_MAX_BINS * 4;
    o = align_up(o + hist_bytes, 256);
    tickets = o;    o = align_up(o + SORT_MAX_PASSES * 4, 256);
    status = o;
    size_t ent = 0;
    for (int p = 0; p < SORT_MAX_PASSES; p++) {
      status_off[p] = ent;
      if (p < sp.npass) ent += (size_t)nblk << sp.bits[p];
    }
    status_bytes = ent * 4;
    o = align_up(o + status_bytes, 256);
    digit_base = o; o = align_up(o + (size_t)SORT_MAX_PASSES * SORT_MAX_BINS * 4, 256);
    total = o + 256;
  }
};

static inline char *align_ptr(const char *p) {
  return (char *)(((uintptr_t)p + 255) & ~(uintptr_t)255);
}

// Kernel-side views ---------------------------------------------------------------
struct GeomView {
  GeomHeader *hdr;
  float4 *rec;
  float4 *cull;
  float *depths;
  float2 *xy;
  uint32_t *tiles;
  uint32_t *offsets;
  uint32_t *chunk_sums;
  uint32_t *chunk_base;
  int nchunk;
};
struct ImageView {
  float *final_T;
  uint32_t *n_contrib;
  uint2 *ranges;
  uint32_t *tile_order;   // tiles sorted by decreasing list length (work scheduling)
  uint32_t *sub_cost;     // per sub-tile (tile * 16 + sub): list entries the forward consumed
  uint32_t *bwd_order;    // sub-tiles by decreasing sub_cost: the backward's CTA numbering
  uint32_t *blk_len;      // per 8x4 block (tile * 8 + block): length of its list
  uint32_t *blk_top;      // per block: 32-entry steps of its list the forward walked
  int tiles_x, tiles_y, tiles;
};
struct BinView {
  BinHeader *hdr;
  unsigned long long *words_a, *words_b, *keys, *keys_unsorted;
  uint32_t *point_list;
  uint16_t *sub_mask;     // per sorted instance: which 4x4 sub-tiles of its tile it can touch
  size_t off_sub_mask, off_blist, off_bpix, off_bwords;
  uint2 *blist;           // per block: (surfel index, position in the tile's list)
  uint32_t *bpix;         // per block-list entry: the pixels of the block that blended it
  uint32_t *bwords;       // per 32 block-list entries: which were blended
  uint32_t *hist, *digit_base, *status, *tickets;
  size_t status_off[SORT_MAX_PASSES];
  size_t zero_bytes;  // [hist .. end of status) is one contiguous region to clear
  int nblk;
};

static inline GeomView make_geom_view(const char *buf, int P) {
  GeomLayout L(P);
  char *b = align_ptr(buf);
  GeomView v;
  v.hdr = (GeomHeader *)b;
  v.rec = (float4 *)(b + L.rec);
  v.cull = (float4 *)(b + L.cull);
  v.depths = (float *)(b + L.depths);
  v.xy = (float2 *)(b + L.xy);
  v.tiles = (uint32_t *)(b + L.tiles);
  v.offsets = (uint32_t *)(b + L.offsets);
  v.chunk_sums = (uint32_t *)(b + L.chunk_sums);
  v.chunk_base = (uint32_t *)(b + L.chunk_base);
  v.nchunk = L.nchunk;
  return v;
}
static inline ImageView make_image_view(const char *buf, int W, int H) {
  ImageLayout L(W, H);
  char *b = align_ptr(buf);
  ImageView v;
  v.final_T = (float *)(b + L.final_T);
  v.n_contrib = (uint32_t *)(b + L.n_contrib);
  v.ranges = (uint2 *)(b + L.ranges);
  v.tile_order = (uint32_t *)(b + L.tile_order);
  v.sub_cost = (uint32_t *)(b + L.sub_cost);
  v.bwd_order = (uint32_t *)(b + L.bwd_order);
  v.blk_len = (uint32_t *)(b + L.blk_len);
  v.blk_top = (uint32_t *)(b + L.blk_top);
  v.tiles_x = L.tiles_x;
  v.tiles_y = L.tiles_y;
  v.tiles = L.tiles;
  return v;
}
static inline BinView make_bin_view(const char *buf, size_t C, int keep_unsorted,
                                    const SortPlan &sp, int ntiles) {
  BinLayout L(C, keep_unsorted, sp, ntiles);
  char *b = align_ptr(buf);
  BinView v;
  v.hdr = (BinHeader *)b;
  v.words_a = (unsigned long long *)(b + L.words_a);
  v.words_b = (unsigned long long *)(b + L.words_b);
  v.keys = (unsigned long long *)(b + L.keys);
  v.keys_unsorted = keep_unsorted ? (unsigned long long *)(b + L.keys_unsorted) : nullptr;
  v.point_list = (uint32_t *)(b + L.point_list);
  v.sub_mask = (uint16_t *)(b + L.sub_mask);
  v.off_sub_mask = L.sub_mask;
  v.off_blist = L.blist;
  v.off_bpix = L.bpix;
  v.off_bwords = L.bwords;
  v.blist = (uint2 *)(b + L.blist);
  v.bpix = (uint32_t *)(b + L.bpix);
  v.bwords = (uint32_t *)(b + L.bwords);
  v.hist = (uint32_t *)(b + L.hist);
  v.digit_base = (uint32_t *)(b + L.digit_base);
  v.status = (uint32_t *)(b + L.status);
  v.tickets = (uint32_t *)(b + L.tickets);
  for (int p = 0; p < SORT_MAX_PASSES; p++) v.status_off[p] = L.status_off[p];
  v.zero_bytes = L.status + L.status_bytes - L.hist;
  v.nblk = L.nblk;
  return v;
}

// records the message returned by splat_last_error(); returns 1
int splat_fail(const char *fmt, ...);

// Launch-side entry points implemented in the .cu files --------------------------------
struct FwdParams {
  int P, W, H;
  float scale_modifier;
  int prefiltered;
  const float *means3D, *opacities, *scales, *rotations, *transMat_precomp;
  const float *viewmatrix, *projmatrix;
  float *out_maps;
  int *radii;
};

void launch_preprocess(const FwdParams &p, const GeomView &g, const ImageView &im,
                       cudaStream_t s);
void launch_chunk_scan(const GeomView &g, uint32_t *host_out, cudaStream_t s);
void launch_emit(const FwdParams &p, const GeomView &g, const ImageView &im,
                 const BinView &b, const SortPlan &sp, uint32_t capacity,
                 cudaStream_t s);
void launch_sort(const FwdParams &p, const GeomView &g, const ImageView &im, const BinView &b,
                 const SortPlan &sp, uint32_t capacity, uint32_t n_estimate, cudaStream_t s);
// true: launch_sort takes the tile-first path (one partition pass by tile, then one CTA per
// tile sorts the tile's range in shared memory); launch_emit then also writes every instance's
// 16-bit sub-tile mask (into the `keys` array, which nothing else uses before the sort's last
// kernel), and the partition pass packs it into the word's top 16 bits.
bool sort_tile_first(const SortPlan &sp, int tiles);
void launch_rebuild_keys(const ImageView &im, const GeomView &g, const BinView &b, uint32_t P,
                         uint32_t capacity, cudaStream_t s);
void launch_digit_scan(const uint32_t *hist, uint32_t *digit_base, int npass, uint2 *ranges,
                       uint32_t *tile_order, int tiles, int tile_pass_single, int nrepl,
                       cudaStream_t s);
struct MaskParams {          // what the last sort pass needs to emit sub_mask (may be all 0)
  uint16_t *sub_mask;
  const float4 *cull;
  const float *projmatrix;
  int tiles_x, H;
};
unsigned long long *launch_sort_passes(unsigned long long *a, unsigned long long *b,
                                       unsigned long long *keys_out, uint32_t *plist_out,
                                       const uint32_t *num_ptr, uint32_t capacity,
                                       uint32_t *status, const size_t *status_off,
                                       uint32_t *tickets, const uint32_t *digit_base,
                                       int npass, const int *shift, const int *bits,
                                       int unpack_last, int idx_bits, uint32_t n_estimate,
                                       cudaStream_t s, const MaskParams *mp = nullptr,
                                       uint32_t P = 0);
// Number of SMs of the current device (queried once per device, then cached): grids of
// persistent / grid-stride kernels are sized from it.
int splat_sm_count();
// Every kernel launch of the library is counted (splat_launch_count() reads the total), so a
// harness can report how many of OUR kernels ran inside a timed region.
void splat_count_launch(int n = 1);
// LS(stream): the stream argument of a <<<>>> launch, counting the launch on the way
#define LS(s) (splat_count_launch(1), (s))

// Programmatic dependent launch (sm_90+): consecutive kernels of a frame are launched with
// cudaLaunchAttributeProgrammaticStreamSerialization, so that a kernel's CTAs are scheduled while
// the previous kernel's last CTAs still run (its launch latency and prologue hide under that
// tail).  Every such kernel starts with grid_dependency_wait() - nothing the previous kernel
// wrote is touched before it has completed and flushed - and lets its own successor go at once
// with grid_launch_dependents() (the successor is only scheduled once all CTAs of this kernel
// have started, and waits at its own top).  Both are no-ops in a kernel launched the plain way.
// SPLAT_B200_PDL=0 launches everything the plain way (developer A/B switch).
#ifdef __CUDACC__
__device__ __forceinline__ void grid_dependency_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
__device__ __forceinline__ void grid_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
bool splat_pdl_enabled();
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                     cudaStream_t s, Args &&...args) {
  splat_count_launch(1);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = splat_pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

// Function attributes (opt-in shared memory, carve-out) are per device: true the first time a
// call site runs on the current device (one bit per device in the caller's static mask).
static inline bool first_use_on_device(unsigned long long &seen) {
  int dev = 0;
  cudaGetDevice(&dev);
  const unsigned long long bit = 1ull << (dev & 63);
  if (seen & bit) return false;
  seen |= bit;
  return true;
}
struct FillList {           // float arrays to clear (any 4-byte alignment, any placement)
  float *ptr[4];
  size_t n[4];
};
void launch_bwd_order(const ImageView &im, const FillList &fill, cudaStream_t s);
void launch_render_forward(const FwdParams &p, const GeomView &g, const ImageView &im,
                           const BinView &b, uint32_t capacity, cudaStream_t s);

struct BwdParams {
  int P, W, H;
  const float *scales, *rotations, *viewmatrix, *projmatrix;
  const int *radii;
  const float *dL_dout;
  float *dL_dmeans2D, *dL_dnormal, *dL_dopacity, *dL_dmeans3D, *dL_dtransMat,
      *dL_dscales, *dL_drotations;
  float *packed;          // optional [P,10]
  int packed_accumulate;
  const struct XchgView *push;   // optional fused push into the window exchange
};
void launch_render_backward(const BwdParams &p, const GeomView &g, const ImageView &im,
                            const BinView &b, cudaStream_t s);
void launch_backward_preprocess(const BwdParams &p, cudaStream_t s);
void launch_mark_visible(int P, const float *means3D, const float *viewmatrix,
                         unsigned char *present, cudaStream_t s);
void launch_unpack_debug(const GeomView &g, int P, float *transMat, float *normal_opacity,
                         cudaStream_t s);

// Exactly-rounded single operations: the compiler must not re-associate or contract
// these, they pin the reference's compiled arithmetic (see DESIGN.md "bit-exactness").
#define FM(a, b) __fmul_rn((a), (b))
#define FA(a, b) __fadd_rn((a), (b))
#define FS(a, b) __fsub_rn((a), (b))
#define FF(a, b, c) __fmaf_rn((a), (b), (c))
