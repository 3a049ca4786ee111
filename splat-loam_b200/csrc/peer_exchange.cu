// peer_exchange.cu — sum of the per-rank [P,10] gradient buffers of a keyframe-sharded window
// over NVLink peer memory (see peer_exchange.cuh for the protocol).  The push half is normally
// executed by the per-surfel VJP kernel itself (dssr_render_bwd.cu), so the transfer of a
// row starts the moment it has been computed.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../../include/splat_b200.h"
#include "common.cuh"
#include "peer_exchange.cuh"

struct splat_xchg {
  XchgView v;
  int P;
  int device;
  size_t arena_bytes;
  char *own;                         // cudaMalloc'ed arena of this rank
  bool opened[XCHG_MAX_WORLD];       // arenas mapped with cudaIpcOpenMemHandle
  uint32_t *fault_host;              // mapped pinned
  bool connected;
  bool local;                        // wired with splat_xchg_connect_local (same process)
  bool pushed;                      // a push for epoch v.epoch has been enqueued
};

namespace {

// Unfused push: copies a finished [P,10] array (or zeros, src == nullptr: a rank with no
// keyframe in this window) into the owners' staging slots, 256 rows (one owner) per CTA step.
__global__ void __launch_bounds__(256) xchg_push_kernel(XchgView x, const float *__restrict__ src,
                                                        int P) {
  __shared__ __align__(16) float s_rows[XCHG_SLICE_ALIGN * XCHG_ROW];
  for (int row0 = blockIdx.x * XCHG_SLICE_ALIGN; row0 < P; row0 += gridDim.x * XCHG_SLICE_ALIGN) {
    const int rows = min(XCHG_SLICE_ALIGN, P - row0);
    for (int k = threadIdx.x; k < XCHG_SLICE_ALIGN * XCHG_ROW; k += blockDim.x)
      s_rows[k] = (src && k < rows * XCHG_ROW) ? src[(size_t)row0 * XCHG_ROW + k] : 0.f;
    __syncthreads();
    xchg_store_block(x, row0, rows, s_rows);
    __syncthreads();
  }
  xchg_publish(x, 0);
}

// The owner sums its slice over the `world` staged copies (fixed rank order: every rank ends
// up with bit-identical sums) and stores the result into every rank's result array.
// DRAIN: the last CTA stays until every OTHER owner has delivered too, so the kernel's
// completion means "the local result is whole" and no separate wait kernel is needed.  (Not
// usable when several ranks share one stream: rank 0's kernel would wait for a later launch.)
template <bool DRAIN>
__global__ void __launch_bounds__(256) xchg_reduce_kernel(XchgView x) {
  __shared__ int s_last;
  xchg_await(x, 0);
  const size_t slice_vec = (size_t)x.slice_rows * XCHG_ROW / 4;       // float4 per slice
  const float4 *stage = reinterpret_cast<const float4 *>(x.stage(x.rank));
  // UNROLL independent vectors per thread and step: enough stores in flight to cover the
  // NVLink round trip (slice_vec is a multiple of 640, so of UNROLL * 32... see static_assert)
  constexpr int UNROLL = 4;
  static_assert(XCHG_SLICE_ALIGN * XCHG_ROW / 4 % (UNROLL * 32) == 0, "whole warps per step");
  const size_t lane = threadIdx.x & 31;
  const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const size_t nwarps = ((size_t)gridDim.x * blockDim.x) >> 5;
  for (size_t base = warp * (UNROLL * 32); base < slice_vec; base += nwarps * (UNROLL * 32)) {
    float4 acc[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; u++) acc[u] = stage[base + u * 32 + lane];
    for (int w = 1; w < x.world; w++) {
      float4 t[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; u++) t[u] = stage[(size_t)w * slice_vec + base + u * 32 + lane];
#pragma unroll
      for (int u = 0; u < UNROLL; u++) {
        acc[u].x += t[u].x; acc[u].y += t[u].y; acc[u].z += t[u].z; acc[u].w += t[u].w;
      }
    }
    const size_t at = (size_t)x.rank * slice_vec + base + lane;
    for (int r = 0; r < x.world; r++) {
      const int peer = (x.rank + r) % x.world;     // spread the ranks over the links
      float4 *dst = reinterpret_cast<float4 *>(x.result(peer)) + at;
#pragma unroll
      for (int u = 0; u < UNROLL; u++) dst[u * 32] = acc[u];
    }
  }
  const bool last = xchg_publish(x, 1);
  if (DRAIN) {
    if (threadIdx.x == 0) s_last = last;
    __syncthreads();
    if (s_last) {
      xchg_await(x, 1);
      if (threadIdx.x == 0) xchg_collect_fault(x);
    }
  }
}

__global__ void __launch_bounds__(32) xchg_wait_kernel(XchgView x) {
  xchg_await(x, 1);
  if (threadIdx.x == 0) xchg_collect_fault(x);
}

int xfail(const char *msg, cudaError_t e = cudaSuccess) {
  if (e != cudaSuccess) return splat_fail("%s: %s", msg, cudaGetErrorString(e));
  return splat_fail("%s", msg);
}

}  // namespace

void launch_xchg_push(const XchgView &x, const float *src, int P, cudaStream_t s) {
  const int blocks = (P + XCHG_SLICE_ALIGN - 1) / XCHG_SLICE_ALIGN;
  const int cap = splat_sm_count() * 8;
  xchg_push_kernel<<<blocks < cap ? blocks : cap, 256, 0, LS(s)>>>(x, src, P);
}
void launch_xchg_reduce(const XchgView &x, bool drain, cudaStream_t s) {
  const int sms = splat_sm_count();
  if (drain)
    xchg_reduce_kernel<true><<<sms, 256, 0, LS(s)>>>(x);
  else
    xchg_reduce_kernel<false><<<sms, 256, 0, LS(s)>>>(x);
}
void launch_xchg_wait(const XchgView &x, cudaStream_t s) {
  xchg_wait_kernel<<<1, 32, 0, LS(s)>>>(x);
}

static int xchg_ready(splat_xchg *x, const char *who) {
  if (!x) return splat_fail("%s: NULL exchange", who);
  if (!x->connected) return splat_fail("%s: exchange is not connected", who);
  if (*x->fault_host)
    return splat_fail("%s: an earlier exchange gave up waiting for rank %u (step %u) - its sums "
                      "are invalid on every rank; destroy and re-create the exchange", who,
                      (*x->fault_host - 1u) / 16u, (*x->fault_host - 1u) % 16u);
  return 0;
}

// dssr_backward's side of a fused push: opens a new epoch and hands out the device view
int xchg_begin_push(splat_xchg *x, int P, XchgView *view) {
  if (xchg_ready(x, "dssr_backward")) return 1;
  if (x->P != P) return splat_fail("dssr_backward: exchange was created for P=%d, not %d", x->P, P);
  if (x->pushed) return splat_fail("dssr_backward: the previous push has not been reduced yet");
  x->v.epoch++;
  x->pushed = true;
  *view = x->v;
  return 0;
}

// ------------------------------------------------------------------ C-ABI ------------
extern "C" {

int splat_xchg_create(int P, int world, int rank, splat_xchg **out) {
  if (!out) return xfail("splat_xchg_create: NULL out");
  *out = nullptr;
  if (P <= 0 || world < 1 || world > XCHG_MAX_WORLD || rank < 0 || rank >= world)
    return splat_fail("splat_xchg_create: bad arguments P=%d world=%d rank=%d (world <= %d)", P,
                      world, rank, XCHG_MAX_WORLD);
  splat_xchg *x = new splat_xchg();
  memset(x, 0, sizeof(*x));
  x->P = P;
  x->v.world = world;
  x->v.rank = rank;
  int rows = (P + world - 1) / world;
  x->v.slice_rows = (rows + XCHG_SLICE_ALIGN - 1) / XCHG_SLICE_ALIGN * XCHG_SLICE_ALIGN;
  x->v.epoch = 0;
  x->v.timeout_ns = XCHG_DEFAULT_TIMEOUT_NS;
  if (const char *t = getenv("SPLAT_B200_XCHG_TIMEOUT_MS")) {
    const double ms = atof(t);
    if (ms > 0) x->v.timeout_ns = (unsigned long long)(ms * 1e6);
  }
  x->arena_bytes = 256 + 2 * x->v.stage_bytes();
  cudaError_t e = cudaGetDevice(&x->device);
  if (e == cudaSuccess) e = cudaMalloc((void **)&x->own, x->arena_bytes);
  if (e == cudaSuccess) e = cudaMemset(x->own, 0, x->arena_bytes);
  if (e == cudaSuccess)
    e = cudaHostAlloc((void **)&x->fault_host, 64, cudaHostAllocPortable | cudaHostAllocMapped);
  if (e == cudaSuccess) {
    *x->fault_host = 0;
    e = cudaHostGetDevicePointer((void **)&x->v.fault, x->fault_host, 0);
  }
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    if (x->own) cudaFree(x->own);
    if (x->fault_host) cudaFreeHost(x->fault_host);
    delete x;
    return xfail("splat_xchg_create", e);
  }
  x->v.arena[rank] = x->own;
  x->connected = world == 1;
  *out = x;
  return 0;
}

int splat_xchg_export(splat_xchg *x, unsigned char *handle) {
  if (!x || !handle) return xfail("splat_xchg_export: NULL argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == SPLAT_XCHG_HANDLE_BYTES, "handle size");
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, x->own);
  if (e != cudaSuccess) return xfail("splat_xchg_export: cudaIpcGetMemHandle", e);
  memcpy(handle, &h, sizeof(h));
  return 0;
}

int splat_xchg_connect(splat_xchg *x, const unsigned char *handles) {
  if (!x || !handles) return xfail("splat_xchg_connect: NULL argument");
  for (int r = 0; r < x->v.world; r++) {
    if (r == x->v.rank || x->v.arena[r]) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, handles + (size_t)r * SPLAT_XCHG_HANDLE_BYTES, sizeof(h));
    void *p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess)
      return splat_fail("splat_xchg_connect: cudaIpcOpenMemHandle(rank %d): %s", r,
                        cudaGetErrorString(e));
    x->v.arena[r] = (char *)p;
    x->opened[r] = true;
  }
  x->connected = true;
  return 0;
}

int splat_xchg_connect_local(splat_xchg *x, splat_xchg *const *peers) {
  if (!x || !peers) return xfail("splat_xchg_connect_local: NULL argument");
  for (int r = 0; r < x->v.world; r++) {
    if (!peers[r] || peers[r]->v.world != x->v.world || peers[r]->v.rank != r ||
        peers[r]->P != x->P)
      return xfail("splat_xchg_connect_local: peers do not form one window");
    x->v.arena[r] = peers[r]->own;
    if (peers[r]->device != x->device) {   // kernels on our device store into the peer's arena
      int cur = 0, can = 0;
      cudaGetDevice(&cur);
      cudaDeviceCanAccessPeer(&can, x->device, peers[r]->device);
      if (!can)
        return splat_fail("splat_xchg_connect_local: device %d cannot access device %d",
                          x->device, peers[r]->device);
      cudaSetDevice(x->device);
      cudaError_t e = cudaDeviceEnablePeerAccess(peers[r]->device, 0);
      cudaSetDevice(cur);
      if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
      else if (e != cudaSuccess) return xfail("splat_xchg_connect_local: cudaDeviceEnablePeerAccess", e);
    }
  }
  x->local = x->v.world > 1;
  x->connected = true;
  return 0;
}

int splat_xchg_push(splat_xchg *x, const float *src, void *stream) {
  if (xchg_ready(x, "splat_xchg_push")) return 1;
  if (x->pushed) return xfail("splat_xchg_push: the previous push has not been reduced yet");
  if ((uintptr_t)src & 3) return xfail("splat_xchg_push: src must be 4-byte aligned");
  x->v.epoch++;
  x->pushed = true;
  launch_xchg_push(x->v, src, x->P, (cudaStream_t)stream);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : xfail("splat_xchg_push", e);
}

int splat_xchg_reduce(splat_xchg *x, void *stream) {
  if (xchg_ready(x, "splat_xchg_reduce")) return 1;
  if (!x->pushed) return xfail("splat_xchg_reduce: nothing has been pushed");
  launch_xchg_reduce(x->v, false, (cudaStream_t)stream);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : xfail("splat_xchg_reduce", e);
}

int splat_xchg_wait(splat_xchg *x, void *stream) {
  if (xchg_ready(x, "splat_xchg_wait")) return 1;
  if (!x->pushed) return xfail("splat_xchg_wait: nothing has been pushed");
  launch_xchg_wait(x->v, (cudaStream_t)stream);
  x->pushed = false;
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : xfail("splat_xchg_wait", e);
}

int splat_xchg_allreduce(splat_xchg *x, void *stream) {
  if (xchg_ready(x, "splat_xchg_allreduce")) return 1;
  if (x->local) {   // peers may share this stream: never wait inside the reduce kernel
    if (splat_xchg_reduce(x, stream)) return 1;
    return splat_xchg_wait(x, stream);
  }
  if (!x->pushed) return xfail("splat_xchg_allreduce: nothing has been pushed");
  launch_xchg_reduce(x->v, true, (cudaStream_t)stream);
  x->pushed = false;
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : xfail("splat_xchg_allreduce", e);
}

int splat_xchg_fault(splat_xchg *x) {
  if (!x || !x->fault_host) return 0;
  return (int)*x->fault_host;
}

float *splat_xchg_result(splat_xchg *x) {
  if (!x) return nullptr;
  return reinterpret_cast<float *>(x->own + 256 + x->v.stage_bytes());
}

int splat_xchg_destroy(splat_xchg *x) {
  if (!x) return 0;
  int cur = 0;
  cudaGetDevice(&cur);
  cudaSetDevice(x->device);
  cudaDeviceSynchronize();
  for (int r = 0; r < x->v.world; r++)
    if (x->opened[r]) cudaIpcCloseMemHandle(x->v.arena[r]);
  if (x->own) cudaFree(x->own);
  if (x->fault_host) cudaFreeHost(x->fault_host);
  cudaSetDevice(cur);
  delete x;
  return 0;
}

}  // extern "C"
