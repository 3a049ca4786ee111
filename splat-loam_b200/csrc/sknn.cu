// sknn.cu — distCUDA2: mean squared distance to the 3 nearest neighbours.
//
// Replaces SimpleKNN::knn, submodules/simple-knn/simple_knn.cu:186-222 (coord2Morton
// :46-71, boxMinMax :79-118, distBoxPoint :120-130, updateKBest :132-146, boxMeanDist
// :148-184), called from distCUDA2 (submodules/simple-knn/spatial.cu:15-26).
//
// The reference sorts the points by 30-bit Morton code, cuts the sorted order into
// boxes of 1024 points and lets EVERY point test EVERY box (P/1024 box tests per point,
// 1024 distance evaluations per surviving box).  Its pruning is conservative, so what
// it returns is the exact 3-NN; any exact 3-NN search that evaluates a distance as
//   d = p - q;  dist = fma(d.z, d.z, fma(d.x, d.x, d.y*d.y))      (the compiled form)
// and keeps the three smallest returns the same floats.  Here: a Morton order too (our own
// radix sort), but at up to 20 bits per axis instead of 10 (the reference's 10-bit cells
// are ~25 cm on an accumulated LiDAR cloud, hundreds of points each), and an implicit
// 8-ary box hierarchy over KNN_LEAF-point (16) leaves that each point descends with a small stack, so
// the work per point is O(log P) box tests and a handful of leaves, and no cudaMalloc /
// host round trip happens inside the call.
#include <cfloat>

#include "../../include/splat_b200.h"
#include "common.cuh"

namespace {

#ifndef KNN_LEAF
#define KNN_LEAF 16
#endif
constexpr int LEAF = KNN_LEAF;   // points per leaf box (divides 32); a warp of queries = 32 / LEAF leaves
constexpr int FAN = 8;
constexpr int MAXL = 12;
#ifndef KNN_NEAR
#define KNN_NEAR 1
#endif
constexpr uint32_t NEAR = KNN_NEAR;   // Morton-neighbour leaves scanned before the descent
constexpr int BB_BLOCKS = 256;

// Morton code width: as many bits per axis as fit a 64-bit [code | index] word, at most 20.
struct KnnPlan {
  int idx_bits, axis_bits, npass;
  int shift[SORT_MAX_PASSES], bits[SORT_MAX_PASSES];
  __host__ explicit KnnPlan(int P) {
    idx_bits = 1;
    while ((1ll << idx_bits) < (long long)P) idx_bits++;
    axis_bits = (64 - idx_bits) / 3;
    if (axis_bits > 20) axis_bits = 20;
    const int total = 3 * axis_bits;
    npass = (total + 9) / 10;
    int sh = idx_bits, left = total;
    for (int i = 0; i < npass; i++) {
      const int b = (left + (npass - i) - 1) / (npass - i);
      shift[i] = sh;
      bits[i] = b;
      sh += b;
      left -= b;
    }
  }
};

struct KnnLayout {
  size_t hdr, partial, words_a, words_b, sorted, boxes, hist, tickets, status, digit_base,
      total, zero_bytes;
  size_t status_off[SORT_MAX_PASSES];
  int nblk, nlevel;
  uint32_t cnt[MAXL];
  size_t box_off[MAXL];  // in nodes
  __host__ explicit KnnLayout(int P) {
    nblk = (P + SORT_MIN_BLOCK_ELEMS - 1) / SORT_MIN_BLOCK_ELEMS;
    if (nblk < 1) nblk = 1;
    nlevel = 0;
    uint32_t c = (uint32_t)((P + LEAF - 1) / LEAF);
    if (c < 1) c = 1;
    size_t nodes = 0;
    while (true) {
      cnt[nlevel] = c;
      box_off[nlevel] = nodes;
      nodes += c;
      nlevel++;
      if (c <= FAN || nlevel == MAXL) break;
      c = (c + FAN - 1) / FAN;
    }
    size_t o = 0;
    hdr = o;      o = align_up(o + 256, 256);
    partial = o;  o = align_up(o + (size_t)(BB_BLOCKS + 1) * 6 * 4, 256);
    words_a = o;  o = align_up(o + (size_t)P * 8, 256);
    words_b = o;  o = align_up(o + (size_t)P * 8, 256);
    sorted = o;   o = align_up(o + (size_t)P * 16, 256);
    boxes = o;    o = align_up(o + nodes * 32, 256);
    hist = o;     o = align_up(o + (size_t)HIST_REPL * SORT_MAX_PASSES * SORT_MAX_BINS * 4, 256);
    tickets = o;  o = align_up(o + 64, 256);
    status = o;
    for (int p = 0; p < SORT_MAX_PASSES; p++) status_off[p] = (size_t)p * nblk * SORT_MAX_BINS;
    o = align_up(o + (size_t)SORT_MAX_PASSES * nblk * SORT_MAX_BINS * 4, 256);
    zero_bytes = o - hist;
    digit_base = o; o = align_up(o + (size_t)SORT_MAX_PASSES * SORT_MAX_BINS * 4, 256);
    total = o + 256;
  }
};

struct KnnLevels {
  int nlevel;
  uint32_t cnt[MAXL];
  uint32_t off[MAXL];
};

__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// cub::DeviceReduce with CustomMin / CustomMax and init (0,0,0) (simple_knn.cu:192-201)
__global__ void __launch_bounds__(256)
bbox_partial_kernel(int P, const float *__restrict__ pts, float *partial) {
  __shared__ float s[8][6];
  float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  for (int i = blockIdx.x * 256 + threadIdx.x; i < P; i += gridDim.x * 256) {
#pragma unroll
    for (int c = 0; c < 3; c++) {
      float v = pts[3 * (size_t)i + c];
      mn[c] = fminf(mn[c], v);
      mx[c] = fmaxf(mx[c], v);
    }
  }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int c = 0; c < 3; c++) {
    float a = warp_min(mn[c]), b = warp_max(mx[c]);
    if (lane == 0) {
      s[wid][c] = a;
      s[wid][3 + c] = b;
    }
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    float r = s[0][threadIdx.x];
    for (int w = 1; w < 8; w++)
      r = threadIdx.x < 3 ? fminf(r, s[w][threadIdx.x]) : fmaxf(r, s[w][threadIdx.x]);
    partial[blockIdx.x * 6 + threadIdx.x] = r;
  }
}

__global__ void bbox_final_kernel(int nblocks, float *partial, uint32_t *num, int P) {
  const int c = threadIdx.x;
  if (c < 6) {
    // (the reference's reduction starts from the origin; the order we sort in is ours to
    // choose, the result does not depend on it, and a tight box gives finer cells)
    float r = partial[c];
    for (int b = 1; b < nblocks; b++)
      r = c < 3 ? fminf(r, partial[b * 6 + c]) : fmaxf(r, partial[b * 6 + c]);
    partial[BB_BLOCKS * 6 + c] = r;
  }
  if (c == 6) *num = (uint32_t)P;
}

// Morton code in the manner of coord2Morton (simple_knn.cu:46-62), `axis_bits` per axis.
__device__ __forceinline__ unsigned long long split3(uint32_t a) {
  unsigned long long x = a & 0x1fffffu;
  x = (x | x << 32) & 0x1f00000000ffffull;
  x = (x | x << 16) & 0x1f0000ff0000ffull;
  x = (x | x << 8) & 0x100f00f00f00f00full;
  x = (x | x << 4) & 0x10c30c30c30c30c3ull;
  x = (x | x << 2) & 0x1249249249249249ull;
  return x;
}

struct MortonParams {
  int idx_bits, axis_bits, npass;
  int shift[SORT_MAX_PASSES], bits[SORT_MAX_PASSES];
};

__global__ void __launch_bounds__(256)
morton_kernel(int P, const float *__restrict__ pts, const float *__restrict__ bbox,
              unsigned long long *words, uint32_t *hist, MortonParams mp) {
  __shared__ uint32_t s_hist[SORT_MAX_PASSES * SORT_MAX_BINS];   // 24 KB
  for (int i = threadIdx.x; i < mp.npass * SORT_MAX_BINS; i += 256) s_hist[i] = 0u;
  __syncthreads();
  const float mnx = bbox[0], mny = bbox[1], mnz = bbox[2];
  const float sx = 1.0f / (bbox[3] - mnx), sy = 1.0f / (bbox[4] - mny), sz = 1.0f / (bbox[5] - mnz);
  const float qmax = (float)((1u << mp.axis_bits) - 1u);
  for (int i = blockIdx.x * 256 + threadIdx.x; i < P; i += gridDim.x * 256) {
    const float x = pts[3 * (size_t)i], y = pts[3 * (size_t)i + 1], z = pts[3 * (size_t)i + 2];
    // fminf/fmaxf drop NaNs (a flat axis gives 0 * inf): such coordinates quantise to 0
    const uint32_t qx = (uint32_t)fminf(fmaxf((x - mnx) * sx * qmax, 0.0f), qmax);
    const uint32_t qy = (uint32_t)fminf(fmaxf((y - mny) * sy * qmax, 0.0f), qmax);
    const uint32_t qz = (uint32_t)fminf(fmaxf((z - mnz) * sz * qmax, 0.0f), qmax);
    const unsigned long long code = split3(qx) | (split3(qy) << 1) | (split3(qz) << 2);
    const unsigned long long w = (code << mp.idx_bits) | (unsigned long long)(uint32_t)i;
    words[i] = w;
    for (int p = 0; p < mp.npass; p++)
      atomicAdd(&s_hist[p * SORT_MAX_BINS + ((uint32_t)(w >> mp.shift[p]) & ((1u << mp.bits[p]) - 1u))], 1u);
  }
  __syncthreads();
  uint32_t *gh = hist + (size_t)(blockIdx.x % HIST_REPL) * (SORT_MAX_PASSES * SORT_MAX_BINS);
  for (int i = threadIdx.x; i < mp.npass * SORT_MAX_BINS; i += 256) {
    const uint32_t v = s_hist[i];
    if (v) atomicAdd(&gh[i], v);
  }
}

// sorted order -> float4(x, y, z, bits(original index)) and the LEAF-point leaf boxes
__global__ void __launch_bounds__(256)
gather_leaf_kernel(int P, const float *__restrict__ pts,
                   const unsigned long long *__restrict__ words, int idx_bits, float4 *sorted,
                   float4 *boxes) {
  const int i = blockIdx.x * 256 + threadIdx.x;  // grid covers whole leaves
  float x = 0.f, y = 0.f, z = 0.f;
  const bool have = i < P;
  if (have) {
    const uint32_t idx = (uint32_t)(words[i] & ((1ull << idx_bits) - 1ull));
    x = pts[3 * (size_t)idx];
    y = pts[3 * (size_t)idx + 1];
    z = pts[3 * (size_t)idx + 2];
    sorted[i] = make_float4(x, y, z, __uint_as_float(idx));
  }
  float lx = have ? x : FLT_MAX, ly = have ? y : FLT_MAX, lz = have ? z : FLT_MAX;
  float hx = have ? x : -FLT_MAX, hy = have ? y : -FLT_MAX, hz = have ? z : -FLT_MAX;
#pragma unroll
  for (int o = LEAF / 2; o > 0; o >>= 1) {   // min / max over the LEAF lanes of a leaf
    lx = fminf(lx, __shfl_xor_sync(0xffffffffu, lx, o));
    ly = fminf(ly, __shfl_xor_sync(0xffffffffu, ly, o));
    lz = fminf(lz, __shfl_xor_sync(0xffffffffu, lz, o));
    hx = fmaxf(hx, __shfl_xor_sync(0xffffffffu, hx, o));
    hy = fmaxf(hy, __shfl_xor_sync(0xffffffffu, hy, o));
    hz = fmaxf(hz, __shfl_xor_sync(0xffffffffu, hz, o));
  }
  if ((threadIdx.x & (LEAF - 1)) == 0 && have) {
    boxes[2 * (size_t)(i / LEAF)] = make_float4(lx, ly, lz, 0.f);
    boxes[2 * (size_t)(i / LEAF) + 1] = make_float4(hx, hy, hz, 0.f);
  }
}

__global__ void __launch_bounds__(256)
level_kernel(uint32_t n, uint32_t nchild, const float4 *__restrict__ child, float4 *node) {
  const uint32_t i = blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  float4 lo = make_float4(FLT_MAX, FLT_MAX, FLT_MAX, 0.f);
  float4 hi = make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, 0.f);
  for (uint32_t c = i * FAN; c < min(i * FAN + FAN, nchild); c++) {
    const float4 a = child[2 * (size_t)c], b = child[2 * (size_t)c + 1];
    lo.x = fminf(lo.x, a.x); lo.y = fminf(lo.y, a.y); lo.z = fminf(lo.z, a.z);
    hi.x = fmaxf(hi.x, b.x); hi.y = fmaxf(hi.y, b.y); hi.z = fmaxf(hi.z, b.z);
  }
  node[2 * (size_t)i] = lo;
  node[2 * (size_t)i + 1] = hi;
}

// updateKBest<3> (simple_knn.cu:132-146) with the distance in its compiled form
__device__ __forceinline__ void update3(const float4 &q, const float4 &p, float &b0, float &b1,
                                        float &b2) {
  const float dx = FS(p.x, q.x), dy = FS(p.y, q.y), dz = FS(p.z, q.z);
  float dist = FF(dz, dz, FF(dx, dx, FM(dy, dy)));
  if (b0 > dist) { float t = b0; b0 = dist; dist = t; }
  if (b1 > dist) { float t = b1; b1 = dist; dist = t; }
  if (b2 > dist) { b2 = dist; }
}

// One WARP = 32 Morton-consecutive points (32 / LEAF neighbouring leaves), walking the hierarchy
// TOGETHER: a node is entered when ANY lane's box distance is within that lane's current
// third-best, and a leaf's LEAF points are then offered to all 32 queries (uniform loads, every
// lane busy).  A
// per-thread walk of the same tree ran at 9 of 32 lanes (the lanes' prunings differ slightly, so
// they drift apart): 11.8 ms at 4 M points.  Offering a query more candidates than it needs
// cannot change its three smallest distances, so the result is the same set of floats.
__global__ void __launch_bounds__(128)
query_kernel(int P, const float4 *__restrict__ sorted, const float4 *__restrict__ boxes,
             KnnLevels lv, float *__restrict__ out) {
  const int i = blockIdx.x * 128 + threadIdx.x;
  const int wbase = i & ~31;                               // warp-uniform: first query of the warp
  if (wbase >= P) return;                                 // whole warp past the end
  const bool live = i < P;
  const float4 q = sorted[live ? i : P - 1];
  float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;
  // the warp's own points and their neighbours in Morton order first: they tighten the bound
  // before the descent
  const int plo = max(0, wbase - 32 * (int)NEAR), phi = min(P, wbase + 32 * ((int)NEAR + 1));
  for (int j = plo; j < phi; j++) {
    const float4 p = sorted[j];
    if (j != i) update3(q, p, b0, b1, b2);
  }
  const uint32_t near_lo = (uint32_t)plo / LEAF;          // leaves [near_lo, near_hi] are done
  const uint32_t near_hi = (uint32_t)(phi - 1) / LEAF;
  uint32_t cur[MAXL], end[MAXL];
  int lvl = lv.nlevel - 1;
  cur[lvl] = 0;
  end[lvl] = lv.cnt[lvl];
  const int top = lvl;
  while (lvl <= top) {
    if (cur[lvl] >= end[lvl]) {
      lvl++;
      continue;
    }
    const uint32_t n = cur[lvl]++;
    if (lvl == 0 && n >= near_lo && n <= near_hi) continue;
    const float4 lo = boxes[2 * (size_t)(lv.off[lvl] + n)];
    const float4 hi = boxes[2 * (size_t)(lv.off[lvl] + n) + 1];
    const float ex = fmaxf(0.f, fmaxf(lo.x - q.x, q.x - hi.x));
    const float ey = fmaxf(0.f, fmaxf(lo.y - q.y, q.y - hi.y));
    const float ez = fmaxf(0.f, fmaxf(lo.z - q.z, q.z - hi.z));
    const float d = ex * ex + ey * ey + ez * ez;
    // slack covers the rounding of d against the rounding of a candidate distance
    if (!__any_sync(0xffffffffu, live && !(d > b2 * 1.000001f))) continue;
    if (lvl == 0) {
      const int s = n * LEAF, e = min(P, s + LEAF);
      for (int j = s; j < e; j++) update3(q, sorted[j], b0, b1, b2);
    } else {
      lvl--;
      cur[lvl] = n * FAN;
      end[lvl] = min(n * FAN + FAN, lv.cnt[lvl]);
    }
  }
  if (live) out[__float_as_uint(q.w)] = (b0 + b1 + b2) / 3.0f;
}

}  // namespace

extern "C" {

size_t sknn_workspace_bytes(int P) { return KnnLayout(P < 0 ? 0 : P).total; }

int sknn_dist2(int P, const float *points, float *mean_dists, void *workspace,
               size_t workspace_bytes, void *stream) {
  if (P < 0) return splat_fail("sknn_dist2: P < 0");
  if (P == 0) return 0;
  if (!points || !mean_dists || !workspace) return splat_fail("sknn_dist2: NULL pointer");
  KnnLayout L(P);
  if (workspace_bytes < L.total)
    return splat_fail("sknn_dist2: workspace too small (%zu < %zu)", workspace_bytes, L.total);
  cudaStream_t st = (cudaStream_t)stream;
  char *w = align_ptr((const char *)workspace);
  uint32_t *num = (uint32_t *)(w + L.hdr);
  float *partial = (float *)(w + L.partial);
  unsigned long long *wa = (unsigned long long *)(w + L.words_a);
  unsigned long long *wb = (unsigned long long *)(w + L.words_b);
  float4 *sorted = (float4 *)(w + L.sorted);
  float4 *boxes = (float4 *)(w + L.boxes);
  uint32_t *hist = (uint32_t *)(w + L.hist);
  uint32_t *tickets = (uint32_t *)(w + L.tickets);
  uint32_t *status = (uint32_t *)(w + L.status);
  uint32_t *digit_base = (uint32_t *)(w + L.digit_base);

  if (cudaMemsetAsync(hist, 0, L.zero_bytes, st) != cudaSuccess)
    return splat_fail("sknn_dist2: memset failed: %s", cudaGetErrorString(cudaGetLastError()));
  int bb = (P + 255) / 256;
  if (bb > BB_BLOCKS) bb = BB_BLOCKS;
  bbox_partial_kernel<<<bb, 256, 0, LS(st)>>>(P, points, partial);
  bbox_final_kernel<<<1, 32, 0, LS(st)>>>(bb, partial, num, P);
  const KnnPlan plan(P);
  MortonParams mp;
  mp.idx_bits = plan.idx_bits;
  mp.axis_bits = plan.axis_bits;
  mp.npass = plan.npass;
  for (int i = 0; i < SORT_MAX_PASSES; i++) {
    mp.shift[i] = plan.shift[i];
    mp.bits[i] = plan.bits[i];
  }
  int mg = (P + 255) / 256;
  if (mg > splat_sm_count() * 4) mg = splat_sm_count() * 4;
  morton_kernel<<<mg, 256, 0, LS(st)>>>(P, points, partial + BB_BLOCKS * 6, wa, hist, mp);
  launch_digit_scan(hist, digit_base, plan.npass, nullptr, nullptr, 0, -1, HIST_REPL, st);
  unsigned long long *res = launch_sort_passes(wa, wb, nullptr, nullptr, num, (uint32_t)P, status,
                                               L.status_off, tickets, digit_base, plan.npass,
                                               plan.shift, plan.bits, 0, plan.idx_bits, (uint32_t)P, st);
  const int leaves = (P + LEAF - 1) / LEAF;
  gather_leaf_kernel<<<(leaves * LEAF + 255) / 256, 256, 0, LS(st)>>>(P, points, res, plan.idx_bits,
                                                                  sorted, boxes);
  KnnLevels lv;
  lv.nlevel = L.nlevel;
  for (int l = 0; l < L.nlevel; l++) {
    lv.cnt[l] = L.cnt[l];
    lv.off[l] = (uint32_t)L.box_off[l];
  }
  for (int l = 1; l < L.nlevel; l++)
    level_kernel<<<(L.cnt[l] + 255) / 256, 256, 0, LS(st)>>>(L.cnt[l], L.cnt[l - 1],
                                                         boxes + 2 * L.box_off[l - 1],
                                                         boxes + 2 * L.box_off[l]);
  query_kernel<<<(P + 127) / 128, 128, 0, LS(st)>>>(P, sorted, boxes, lv, mean_dists);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return splat_fail("sknn_dist2: launch failed: %s", cudaGetErrorString(e));
  return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------
// sknn_dist2_static — distCUDA2 for callers that must not allocate device memory.
//
// Who needs this: code that sits on top of this C-ABI but owns no allocator of its own.
// Typical cases are a sensor / tracking thread with a latency budget (cudaMalloc and
// cudaFree serialise the whole device and can stall a frame), a process in which another
// framework's caching allocator has already claimed the GPU's memory (a second allocator
// asking the driver for more can simply fail), or a plug-in whose host forbids allocating
// CUDA calls in its callbacks.  Such a caller passes only the points and the output array;
// the scratch that sknn_dist2 takes from its caller lives instead in the file-scope device
// array below, which every GPU that loads this library carries until the process exits.
//
// Limits and rules (also in include/splat_b200.h):
//   * at most SKNN_STATIC_MAX_POINTS points per call; larger clouds are refused with an
//     error, nothing is launched;
//   * the first call on a device looks up that device's address of the array
//     (cudaGetSymbolAddress with the device current) and caches it; make that first call
//     outside of any stream capture;
//   * all calls share the one array and it keeps whatever the previous call left there.
//     The tables that must start at zero (digit histograms, tickets, look-back status) are
//     cleared at the start of every call by sknn_dist2 itself (a cudaMemsetAsync on the
//     caller's stream, not an allocation); everything else is overwritten before it is read;
//   * because the array is shared, two calls whose GPU work overlaps (different streams, or
//     different host threads) race on it: serialise them;
//   * no cudaMalloc, cudaFree or allocating library routine is called on any path.
#include <mutex>

#define SKNN_STATIC_BYTES (4u << 20)   // >= KnnLayout(SKNN_STATIC_MAX_POINTS).total (3.4 MB)

static __device__ __align__(256) unsigned char g_sknn_static_scratch[SKNN_STATIC_BYTES];

namespace {
constexpr int kMaxDevices = 64;
void *g_static_scratch_ptr[kMaxDevices];   // device ordinal -> device address of the array
std::mutex g_static_scratch_mu;
}  // namespace

extern "C" int sknn_dist2_static(int P, const float *points, float *mean_dists, void *stream) {
  if (P < 0) return splat_fail("sknn_dist2_static: P < 0");
  if (P > SKNN_STATIC_MAX_POINTS)
    return splat_fail("sknn_dist2_static: %d points exceed the static scratch limit of %d; use "
                      "sknn_dist2 with a caller-owned workspace", P, SKNN_STATIC_MAX_POINTS);
  if (P == 0) return 0;
  if (KnnLayout(P).total > SKNN_STATIC_BYTES)
    return splat_fail("sknn_dist2_static: internal error, scratch layout larger than the array");
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices)
    return splat_fail("sknn_dist2_static: cudaGetDevice failed or device ordinal %d unsupported: %s",
                      dev, cudaGetErrorString(e));
  void *scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_static_scratch_mu);
    if (!g_static_scratch_ptr[dev]) {
      // In host code the array's name is only a handle; the address a kernel can use on
      // THIS device has to be asked for, with this device current.
      void *p = nullptr;
      e = cudaGetSymbolAddress(&p, g_sknn_static_scratch);
      if (e != cudaSuccess || !p)
        return splat_fail("sknn_dist2_static: cudaGetSymbolAddress failed: %s",
                          cudaGetErrorString(e));
      g_static_scratch_ptr[dev] = p;
    }
    scratch = g_static_scratch_ptr[dev];
  }
  return sknn_dist2(P, points, mean_dists, scratch, SKNN_STATIC_BYTES, stream);
}
