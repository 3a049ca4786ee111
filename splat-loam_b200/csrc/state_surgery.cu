// state_surgery.cu — pruning and growing the surfel parameter tensors and their Adam
// moments in one pass over all of them.
//
// Replaces the per-group tensor surgery of scene/gaussian_model.py:223-316
// (_prune_optimizer / prune_points: a boolean-mask index per tensor, each with its own
// nonzero + gather + host sync; cat_tensors_to_optimizer / densification_postfix: a torch.cat
// per parameter and two cat-with-zeros per moment), called once per keyframe from
// slam/mapper.py:50-138,216-233.  Up to 12 row-major float tensors that share the row index
// (4 parameters, 4 exp_avg, 4 exp_avg_sq) are compacted or extended by ONE kernel; the mask is
// scanned once, and the host learns the surviving row count through one 4-byte read.
#include "../../include/splat_b200.h"
#include "common.cuh"

namespace {

constexpr int SCAN_BLOCK = 1024;

struct RowSet {
  const float *src[SPLAT_SURGERY_MAX_TENSORS];
  const float *ext[SPLAT_SURGERY_MAX_TENSORS];
  float *dst[SPLAT_SURGERY_MAX_TENSORS];
  int width[SPLAT_SURGERY_MAX_TENSORS];
  int n;
};

// kept rows per block of 1024
__global__ void __launch_bounds__(SCAN_BLOCK)
keep_count_kernel(int P, const unsigned char *__restrict__ keep, uint32_t *__restrict__ block_count) {
  __shared__ uint32_t s_warp[32];
  const int i = blockIdx.x * SCAN_BLOCK + threadIdx.x;
  const bool k = i < P && keep[i] != 0;
  const uint32_t b = __ballot_sync(0xffffffffu, k);
  if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = __popc(b);
  __syncthreads();
  if (threadIdx.x < 32) {
    uint32_t v = s_warp[threadIdx.x];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0) block_count[blockIdx.x] = v;
  }
}

// exclusive scan of the block counts in place (one CTA, any number of blocks); the total goes
// to block_count[nblocks] and to the mapped host word
__global__ void __launch_bounds__(1024)
block_scan_kernel(int nblocks, uint32_t *block_count, uint32_t *host_total) {
  __shared__ uint32_t s_warp[32];
  __shared__ uint32_t s_carry;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) s_carry = 0u;
  __syncthreads();
  for (int base = 0; base < nblocks; base += 1024) {
    const int i = base + tid;
    const uint32_t v = i < nblocks ? block_count[i] : 0u;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += n;
    }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    uint32_t before = s_carry;
    for (int w = 0; w < wid; w++) before += s_warp[w];
    if (i < nblocks) block_count[i] = before + inc - v;
    __syncthreads();
    if (tid == 1023) s_carry = before + inc;
    __syncthreads();
  }
  if (tid == 0) {
    block_count[nblocks] = s_carry;
    if (host_total) *host_total = s_carry;
  }
}

// row i of every tensor moves to row (kept rows before i)
__global__ void __launch_bounds__(SCAN_BLOCK)
compact_rows_kernel(int P, const unsigned char *__restrict__ keep,
                    const uint32_t *__restrict__ block_offset, RowSet rs) {
  __shared__ uint32_t s_warp[32];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int i = blockIdx.x * SCAN_BLOCK + tid;
  const bool k = i < P && keep[i] != 0;
  const uint32_t b = __ballot_sync(0xffffffffu, k);
  if (lane == 0) s_warp[wid] = __popc(b);
  __syncthreads();
  if (!k) return;
  uint32_t before = block_offset[blockIdx.x] + __popc(b & ((1u << lane) - 1u));
  for (int w = 0; w < wid; w++) before += s_warp[w];
  for (int t = 0; t < rs.n; t++) {
    const int w = rs.width[t];
    const float *s = rs.src[t] + (size_t)i * w;
    float *d = rs.dst[t] + (size_t)before * w;
    for (int c = 0; c < w; c++) d[c] = s[c];
  }
}

// dst = [src (n_old rows) ; ext (n_new rows, or zeros when ext is NULL)] for every tensor; one
// thread per output float of the widest layout, grid-stride over rows
__global__ void __launch_bounds__(256)
append_rows_kernel(long long n_old, long long n_new, RowSet rs) {
  const long long total = n_old + n_new;
  for (int t = 0; t < rs.n; t++) {
    const long long w = rs.width[t];
    const long long cut = n_old * w, end = total * w;
    const float *s = rs.src[t], *e = rs.ext[t];
    float *d = rs.dst[t];
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < end;
         k += (long long)gridDim.x * blockDim.x)
      d[k] = k < cut ? s[k] : (e ? e[k - cut] : 0.0f);
  }
}

int fill_rowset(RowSet &rs, int ntensors, const float *const *src, const float *const *ext,
                float *const *dst, const int *width, bool need_src, const char *who) {
  if (ntensors < 1 || ntensors > SPLAT_SURGERY_MAX_TENSORS)
    return splat_fail("%s: ntensors must be 1..%d", who, SPLAT_SURGERY_MAX_TENSORS);
  if (!dst || !width || (need_src && !src)) return splat_fail("%s: NULL argument", who);
  rs.n = ntensors;
  for (int t = 0; t < ntensors; t++) {
    if (width[t] < 1) return splat_fail("%s: width[%d] = %d", who, t, width[t]);
    rs.src[t] = src ? src[t] : nullptr;
    rs.ext[t] = ext ? ext[t] : nullptr;
    rs.dst[t] = dst[t];
    rs.width[t] = width[t];
    if (!dst[t]) return splat_fail("%s: dst[%d] is NULL", who, t);
  }
  return 0;
}

thread_local uint32_t *g_total = nullptr, *g_total_dev = nullptr;

}  // namespace

extern "C" {

size_t splat_compact_workspace_bytes(int P) {
  const size_t nblocks = ((size_t)(P < 0 ? 0 : P) + SCAN_BLOCK - 1) / SCAN_BLOCK;
  return (nblocks + 1) * sizeof(uint32_t) + 256;
}

int splat_compact_plan(int P, const unsigned char *keep, void *workspace, int *count,
                       void *stream) {
  if (P < 0 || !count) return splat_fail("splat_compact_plan: bad arguments");
  *count = 0;
  if (P == 0) return 0;
  if (!keep || !workspace) return splat_fail("splat_compact_plan: NULL pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (!g_total) {
    if (cudaHostAlloc((void **)&g_total, 64, cudaHostAllocPortable | cudaHostAllocMapped) != cudaSuccess ||
        cudaHostGetDevicePointer((void **)&g_total_dev, g_total, 0) != cudaSuccess)
      return splat_fail("splat_compact_plan: cannot allocate the pinned count word");
  }
  uint32_t *bc = reinterpret_cast<uint32_t *>(align_ptr((const char *)workspace));
  const int nblocks = (P + SCAN_BLOCK - 1) / SCAN_BLOCK;
  keep_count_kernel<<<nblocks, SCAN_BLOCK, 0, LS(st)>>>(P, keep, bc);
  block_scan_kernel<<<1, 1024, 0, LS(st)>>>(nblocks, bc, g_total_dev);
  cudaError_t e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) return splat_fail("splat_compact_plan: %s", cudaGetErrorString(e));
  *count = (int)g_total[0];
  return 0;
}

int splat_compact_rows(int P, const unsigned char *keep, const void *workspace, int ntensors,
                       const float *const *src, float *const *dst, const int *width,
                       void *stream) {
  if (P < 0) return splat_fail("splat_compact_rows: P < 0");
  if (P == 0) return 0;
  if (!keep || !workspace) return splat_fail("splat_compact_rows: NULL pointer");
  RowSet rs;
  if (fill_rowset(rs, ntensors, src, nullptr, dst, width, true, "splat_compact_rows")) return 1;
  for (int t = 0; t < ntensors; t++)
    if (!src[t]) return splat_fail("splat_compact_rows: src[%d] is NULL", t);
  const uint32_t *bc = reinterpret_cast<const uint32_t *>(align_ptr((const char *)workspace));
  const int nblocks = (P + SCAN_BLOCK - 1) / SCAN_BLOCK;
  compact_rows_kernel<<<nblocks, SCAN_BLOCK, 0, LS((cudaStream_t)stream)>>>(P, keep, bc, rs);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : splat_fail("splat_compact_rows: %s", cudaGetErrorString(e));
}

int splat_append_rows(int n_old, int n_new, int ntensors, const float *const *src,
                      const float *const *ext, float *const *dst, const int *width,
                      void *stream) {
  if (n_old < 0 || n_new < 0) return splat_fail("splat_append_rows: negative row count");
  if (n_old + (long long)n_new == 0) return 0;
  RowSet rs;
  if (fill_rowset(rs, ntensors, src, ext, dst, width, n_old > 0, "splat_append_rows")) return 1;
  for (int t = 0; t < ntensors; t++)
    if (n_old > 0 && !rs.src[t]) return splat_fail("splat_append_rows: src[%d] is NULL", t);
  int wmax = 1;
  for (int t = 0; t < ntensors; t++) wmax = width[t] > wmax ? width[t] : wmax;
  const long long work = ((long long)n_old + n_new) * wmax;
  long long blocks = (work + 255) / 256;
  const long long cap = (long long)splat_sm_count() * 16;
  if (blocks > cap) blocks = cap;
  append_rows_kernel<<<(int)blocks, 256, 0, LS((cudaStream_t)stream)>>>(n_old, n_new, rs);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : splat_fail("splat_append_rows: %s", cudaGetErrorString(e));
}

}  // extern "C"
