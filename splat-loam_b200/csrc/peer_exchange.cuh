// peer_exchange.cuh — device-side view of the keyframe-window gradient exchange.
//
// No counterpart in the reference (single GPU, slam/mapper.py:150-156).  Every rank owns a
// contiguous slice of the [P,10] gradient rows.  An exchange is
//   1. PUSH (fused into the per-surfel VJP kernel, or xchg_push_kernel): every rank stores its
//      row for surfel i straight into the OWNER's staging area, slot `rank` (NVLink stores);
//   2. REDUCE+GATHER (xchg_reduce_kernel): the owner sums the `world` staged copies of its
//      slice in rank order and stores the sum into every rank's result array;
//   3. WAIT (xchg_wait_kernel): until every owner has delivered.
// Steps are separated by flag barriers in peer memory: the last CTA of a step publishes
// `epoch` into flag word [step][rank] of every peer; the next step polls its own flags.
#pragma once
#include <cstdint>

#define XCHG_MAX_WORLD 16
#define XCHG_ROW 10             // floats per surfel: means3D 3 | opacity 1 | scales 2 | rotations 4
#define XCHG_SLICE_ALIGN 256    // slice_rows is a multiple of this: a 256-row block of the push
                                // kernels has ONE owner and a contiguous, 16-byte aligned target
#define XCHG_DEFAULT_TIMEOUT_NS (30ull * 1000 * 1000 * 1000)   // a wait gives up after this long
                                // (%globaltimer nanoseconds; SPLAT_B200_XCHG_TIMEOUT_MS overrides)

// Per-rank arena (one cudaMalloc, peer-mapped through CUDA IPC):
//   [ flags u32[2][XCHG_MAX_WORLD] | ticket u32[2] | pad to 256 B |
//     stage f32[world][slice_rows][10] | result f32[world*slice_rows][10] ]
struct XchgArena {
  uint32_t flags[2][XCHG_MAX_WORLD];
  uint32_t ticket[2];
  uint32_t fault;     // set (by ANY rank, through peer memory) when a wait gave up: the sums of
                      // this and every later epoch are not to be trusted
  uint32_t pad[64 - 2 * XCHG_MAX_WORLD - 3];
};
static_assert(sizeof(XchgArena) == 256, "arena header is 256 bytes");

struct XchgView {
  char *arena[XCHG_MAX_WORLD];   // every rank's arena as mapped in THIS process (own included)
  uint32_t *fault;               // mapped pinned host word: set when a wait gave up
  unsigned long long timeout_ns; // how long a wait spins before it gives up
  int world, rank, slice_rows;
  uint32_t epoch;

  __host__ __device__ size_t stage_bytes() const {
    return (size_t)world * slice_rows * XCHG_ROW * sizeof(float);
  }
  __device__ XchgArena *hdr(int r) const { return reinterpret_cast<XchgArena *>(arena[r]); }
  __device__ float *stage(int r) const { return reinterpret_cast<float *>(arena[r] + 256); }
  __device__ float *result(int r) const {
    return reinterpret_cast<float *>(arena[r] + 256 + stage_bytes());
  }
  // where rank `rank`'s copy of surfel row i lives: owner's staging, slot `rank`
  __device__ float *push_row(int i) const {
    const int owner = i / slice_rows, local = i - owner * slice_rows;
    return stage(owner) + ((size_t)rank * slice_rows + local) * XCHG_ROW;
  }
};

// Store a block's XCHG_SLICE_ALIGN rows (staged in shared memory, row-major, zero padded) to
// their owner as 16-byte vectors: remote stores want whole, contiguous sectors - a thread
// writing its own 40-byte row in 8-byte pieces uses a fraction of the link.
__device__ __forceinline__ void xchg_store_block(const XchgView &x, int row0, int rows,
                                                 const float *s_rows) {
  float4 *dst = reinterpret_cast<float4 *>(x.push_row(row0));
  const float4 *src = reinterpret_cast<const float4 *>(s_rows);
  const int nvec = (rows * XCHG_ROW + 3) / 4;
  for (int v = threadIdx.x; v < nvec; v += blockDim.x) dst[v] = src[v];
}

// Called by every thread of every CTA of a step after its last peer store: the last CTA to
// arrive publishes `epoch` to all peers.  Ordering is by scope cumulativity, as in a grid
// barrier: the CTA's stores -> bar.sync -> thread 0's device-scope fence -> ticket; the last
// CTA observes every ticket, then its SYSTEM-scope fence orders all of them before the flags.
// (A system fence in every thread would stall each CTA for an NVLink round trip.)
// Returns (to thread 0 only) whether this CTA was the last one.
__device__ __forceinline__ bool xchg_publish(const XchgView &x, int step) {
  bool last = false;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    uint32_t *ticket = &x.hdr(x.rank)->ticket[step];
    const uint32_t total = gridDim.x * gridDim.y;
    if (atomicAdd(ticket, 1u) == total - 1u) {
      last = true;
      *ticket = 0u;
      __threadfence_system();
      for (int r = 0; r < x.world; r++) {
        volatile uint32_t *f = &x.hdr(r)->flags[step][x.rank];
        *f = x.epoch;
      }
      __threadfence_system();
    }
  }
  return last;
}

__device__ __forceinline__ unsigned long long xchg_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Block-wide: wait until every rank has published `epoch` for `step` (flags live in OUR arena).
// Epochs only grow; wrap-around is handled by the signed difference.
// A wait that gives up POISONS the exchange for everybody: the fault word of every rank's arena
// is set through peer memory before this rank goes on (and publishes its own flags), so a peer
// that later sees this rank's arrival also sees the fault (xchg_collect_fault).  The sums of a
// poisoned epoch contain stale rows and must not be used.
__device__ __forceinline__ void xchg_await(const XchgView &x, int step) {
  if (threadIdx.x < x.world) {
    volatile uint32_t *f = &x.hdr(x.rank)->flags[step][threadIdx.x];
    const unsigned long long t0 = xchg_now_ns();
    while ((int32_t)(*f - x.epoch) < 0) {
      if (xchg_now_ns() - t0 > x.timeout_ns) {
        const uint32_t code = 1u + (uint32_t)step + 16u * (uint32_t)threadIdx.x;
        for (int r = 0; r < x.world; r++) {
          volatile uint32_t *pf = &x.hdr(r)->fault;
          *pf = code;
        }
        __threadfence_system();
        if (x.fault) *x.fault = code;
        break;
      }
      __nanosleep(100);
    }
  }
  __syncthreads();
  __threadfence_system();
}

// After the last wait of an exchange (one thread): hand a fault that any rank raised in our
// arena to the host (mapped pinned word), where the next splat_xchg_* call and
// splat_xchg_fault() find it.
__device__ __forceinline__ void xchg_collect_fault(const XchgView &x) {
  const uint32_t f = *(volatile uint32_t *)&x.hdr(x.rank)->fault;
  if (f && x.fault) *x.fault = f;
}

struct splat_xchg;
int xchg_begin_push(splat_xchg *x, int P, XchgView *view);
void launch_xchg_push(const XchgView &x, const float *src, int P, cudaStream_t s);
void launch_xchg_reduce(const XchgView &x, bool drain, cudaStream_t s);
void launch_xchg_wait(const XchgView &x, cudaStream_t s);
