// dssr_api.cu — the C-ABI of the rasterizer (include/splat_b200.h): argument checks,
// buffer carving and kernel sequencing.  Mirrors the host side of
// CudaRasterizer::Rasterizer::forward / backward / markVisible
// (submodules/diff-surfel-spherical-rasterization/cuda_rasterizer/rasterizer_impl.cu:
// 117-122, 170-262, 266-313).
#include <atomic>
#include <chrono>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/splat_b200.h"
#include "common.cuh"
#include "peer_exchange.cuh"

static thread_local std::string g_err;

namespace {

// Host-side state of dssr_forward, per host thread AND per device: a thread that drives
// several GPUs gets an event (events belong to the device they were created on), a count word
// and a size estimate for each of them.
constexpr int MAX_DEVICES = 64;
struct DeviceState {
  uint32_t *pinned = nullptr;      // 2 words: num_rendered, status (mapped pinned host memory)
  uint32_t *pinned_dev = nullptr;  // this device's view of `pinned`
  cudaEvent_t event = nullptr;     // "the count has landed"
  uint32_t last_count = 0;         // num_rendered of the previous forward on this device
};
thread_local DeviceState g_state[MAX_DEVICES];

#define fail splat_fail

#define CUDA_TRY(expr)                                                        \
  do {                                                                        \
    cudaError_t e_ = (expr);                                                  \
    if (e_ != cudaSuccess)                                                    \
      return fail("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_),     \
                  __FILE__, __LINE__);                                        \
  } while (0)

// CHECK_CUDA (auxiliary.h:368-377): with debug, synchronise and check after a launch.
#define LAUNCH_CHECK(debug, stream, what)                                     \
  do {                                                                        \
    cudaError_t e_ = cudaGetLastError();                                      \
    if (e_ == cudaSuccess && (debug)) e_ = cudaStreamSynchronize(stream);     \
    if (e_ != cudaSuccess)                                                    \
      return fail("[CUDA ERROR] in %s: %s", what, cudaGetErrorString(e_));    \
  } while (0)

// SPLAT_B200_TRACE=1: host-side time of each phase of dssr_forward on stderr
struct Trace {
  bool on;
  std::chrono::steady_clock::time_point t;
  Trace() : on(getenv("SPLAT_B200_TRACE") != nullptr), t(std::chrono::steady_clock::now()) {}
  void lap(const char *what) {
    if (!on) return;
    auto n = std::chrono::steady_clock::now();
    fprintf(stderr, "[splat trace] %-22s %8.1f us\n", what,
            std::chrono::duration<double, std::micro>(n - t).count());
    t = n;
  }
};

// Optional per-stage device timing (splat_profile_*): CUDA events recorded on the launch
// stream around each stage (one pair per stage per call, from a growing pool), summed per
// stage name when the caller reads them back after it has synchronised.
struct Prof {
  struct Rec {
    const char *name;
    cudaEvent_t beg, end;
    int dev;               // the device the two events were created on
  };
  bool on = false;
  std::string only;        // empty: every stage
  std::vector<Rec> pool;   // events are created once and reused
  size_t used = 0;
  Rec *take(const char *nm) {
    if (!on || (!only.empty() && only != nm)) return nullptr;
    int dev = 0;
    cudaGetDevice(&dev);
    if (used < pool.size() && pool[used].dev != dev) {   // pooled events of another device
      cudaEventDestroy(pool[used].beg);
      cudaEventDestroy(pool[used].end);
      pool[used].beg = pool[used].end = nullptr;
      if (cudaEventCreate(&pool[used].beg) != cudaSuccess ||
          cudaEventCreate(&pool[used].end) != cudaSuccess) {
        pool.resize(used);   // (later entries are recreated on demand)
        return nullptr;
      }
      pool[used].dev = dev;
    }
    if (used == pool.size()) {
      Rec r{nm, nullptr, nullptr, dev};
      if (cudaEventCreate(&r.beg) != cudaSuccess) return nullptr;
      if (cudaEventCreate(&r.end) != cudaSuccess) {
        cudaEventDestroy(r.beg);
        return nullptr;
      }
      pool.push_back(r);
    }
    Rec *r = &pool[used++];
    r->name = nm;
    return r;
  }
};
thread_local Prof g_prof;

struct ProfScope {
  Prof::Rec *r;
  cudaStream_t st;
  ProfScope(const char *nm, cudaStream_t st_) : r(g_prof.take(nm)), st(st_) {
    if (r) cudaEventRecord(r->beg, st);
  }
  ~ProfScope() {
    if (r) cudaEventRecord(r->end, st);
  }
};

int device_state(DeviceState **out) {
  int dev = 0;
  CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= MAX_DEVICES) return fail("device ordinal %d out of range", dev);
  DeviceState &d = g_state[dev];
  if (!d.pinned) {
    CUDA_TRY(cudaHostAlloc((void **)&d.pinned, 64, cudaHostAllocPortable | cudaHostAllocMapped));
    CUDA_TRY(cudaHostGetDevicePointer((void **)&d.pinned_dev, d.pinned, 0));
  }
  if (!d.event) CUDA_TRY(cudaEventCreateWithFlags(&d.event, cudaEventDisableTiming));
  *out = &d;
  return 0;
}

std::atomic<unsigned long long> g_launches{0};

int run_binning_and_render(const dssr_forward_args *a, const FwdParams &fp,
                           const GeomView &gv, const ImageView &iv, const SortPlan &sp,
                           uint32_t capacity, uint32_t n_estimate, cudaStream_t st) {
  const int keep = (a->debug & 2) ? 1 : 0;
  const bool dbg = a->debug & 1;
  BinLayout bl(capacity, keep, sp, iv.tiles);
  char *bptr = a->binning_alloc(a->binning_ctx, bl.total);
  if (!bptr) return fail("binning_alloc(%zu) returned NULL", bl.total);
  BinView bv = make_bin_view(bptr, capacity, keep, sp, iv.tiles);
  CUDA_TRY(cudaMemsetAsync(bv.hist, 0, bv.zero_bytes, st));
  {
    ProfScope ps("emit", st);
    launch_emit(fp, gv, iv, bv, sp, capacity, st);
  }
  LAUNCH_CHECK(dbg, st, "emit");
  {
    ProfScope ps("sort", st);
    launch_sort(fp, gv, iv, bv, sp, capacity, n_estimate, st);
  }
  LAUNCH_CHECK(dbg, st, "sort");
  {
    ProfScope ps("render_forward", st);
    launch_render_forward(fp, gv, iv, bv, capacity, st);
  }
  LAUNCH_CHECK(dbg, st, "render_forward");
  return 0;
}

}  // namespace

void splat_count_launch(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

bool splat_pdl_enabled() {
  static const bool on = [] {
    const char *e = getenv("SPLAT_B200_PDL");
    return !(e && e[0] == '0');
  }();
  return on;
}

int splat_sm_count() {
  static std::atomic<int> cache[MAX_DEVICES];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= MAX_DEVICES) dev = 0;
  int n = cache[dev].load(std::memory_order_relaxed);
  if (n <= 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n < 1)
      n = 1;
    cache[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

int splat_fail(const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return 1;
}

extern "C" {

const char *splat_last_error(void) { return g_err.c_str(); }
int splat_abi_version(void) { return 8; }
unsigned long long splat_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

size_t dssr_geometry_bytes(int P) { return GeomLayout(P < 0 ? 0 : P).total; }
size_t dssr_image_bytes(int W, int H) { return ImageLayout(W, H).total; }
size_t dssr_binning_bytes(int64_t capacity, int W, int H, int P, int keep_unsorted) {
  ImageLayout il(W, H);
  SortPlan sp(P, il.tiles);
  return BinLayout((size_t)(capacity < 0 ? 0 : capacity), keep_unsorted, sp, il.tiles).total;
}

int splat_profile_enable(int on) {
  g_prof.on = on != 0;
  g_prof.used = 0;
  return 0;
}

int splat_profile_select(const char *stage) {
  g_prof.only = stage ? stage : "";
  return 0;
}

int splat_profile_read(int max_stages, const char **names, float *ms) {
  int n = 0;
  for (size_t i = 0; i < g_prof.used; i++) {
    const Prof::Rec &r = g_prof.pool[i];
    float t = 0.f;
    if (cudaEventElapsedTime(&t, r.beg, r.end) != cudaSuccess) t = -1.f;
    int k = 0;
    while (k < n && names[k] != r.name) k++;   // stage names are string literals
    if (k == n) {
      if (n == max_stages) continue;
      names[n] = r.name;
      ms[n] = 0.f;
      n++;
    }
    if (t < 0.f || ms[k] < 0.f) ms[k] = -1.f; else ms[k] += t;
  }
  g_prof.used = 0;
  return n;
}

int dssr_forward(const dssr_forward_args *a, int *num_rendered) {
  if (!a || !num_rendered) return fail("dssr_forward: NULL argument");
  *num_rendered = 0;
  if (a->P < 0 || a->W <= 0 || a->H <= 0) return fail("dssr_forward: bad sizes P=%d W=%d H=%d", a->P, a->W, a->H);
  if (!a->out_maps) return fail("dssr_forward: out_maps is NULL");
  cudaStream_t st = (cudaStream_t)a->stream;
  const size_t HW = (size_t)a->W * a->H;
  if (a->P == 0) {  // rasterize_points.cu:73: nothing launched, zeros returned
    CUDA_TRY(cudaMemsetAsync(a->out_maps, 0, 7 * HW * sizeof(float), st));
    return 0;
  }
  if (!a->means3D || !a->opacities || !a->viewmatrix || !a->projmatrix || !a->radii)
    return fail("dssr_forward: NULL input pointer");
  const bool precomp = a->transMat_precomp != nullptr;
  if (!precomp && (!a->scales || !a->rotations))
    return fail("dssr_forward: need scales+rotations or transMat_precomp");
  if (!a->geom_alloc || !a->binning_alloc || !a->image_alloc)
    return fail("dssr_forward: NULL allocator callback");
  if (((uintptr_t)a->scales & 7) || ((uintptr_t)a->rotations & 15) ||
      ((uintptr_t)a->transMat_precomp & 15))
    return fail("dssr_forward: scales must be 8-byte and rotations / transMat_precomp 16-byte aligned");
  Trace tr;
  DeviceState *ds = nullptr;
  if (device_state(&ds)) return 1;
  const bool dbg = a->debug & 1;

  GeomLayout gl(a->P);
  ImageLayout il(a->W, a->H);
  SortPlan sp(a->P, il.tiles);
  if (!sp.ok)
    return fail("dssr_forward: P=%d with %d tiles does not fit the 64-bit sort word", a->P, il.tiles);
  char *gptr = a->geom_alloc(a->geom_ctx, gl.total);
  if (!gptr) return fail("geom_alloc(%zu) returned NULL", gl.total);
  char *iptr = a->image_alloc(a->image_ctx, il.total);
  if (!iptr) return fail("image_alloc(%zu) returned NULL", il.total);
  GeomView gv = make_geom_view(gptr, a->P);
  ImageView iv = make_image_view(iptr, a->W, a->H);
  tr.lap("alloc geom+image");

  FwdParams fp;
  fp.P = a->P; fp.W = a->W; fp.H = a->H;
  fp.scale_modifier = a->scale_modifier;
  fp.prefiltered = a->prefiltered;
  fp.means3D = a->means3D; fp.opacities = a->opacities;
  fp.scales = a->scales; fp.rotations = a->rotations;
  fp.transMat_precomp = a->transMat_precomp;
  fp.viewmatrix = a->viewmatrix; fp.projmatrix = a->projmatrix;
  fp.out_maps = a->out_maps; fp.radii = a->radii;

  {
    ProfScope ps("preprocess", st);
    launch_preprocess(fp, gv, iv, st);
    launch_chunk_scan(gv, ds->pinned_dev, st);   // count + status land in ds->pinned
  }
  LAUNCH_CHECK(dbg, st, "preprocess");
  CUDA_TRY(cudaEventRecord(ds->event, st));
  tr.lap("enqueue preprocess");

  // With a capacity hint the rest of the frame is enqueued BEFORE the host waits for
  // the instance count (the reference blocks here: rasterizer_impl.cu:216-219).
  int64_t hint = a->binning_hint;
  if (hint > 0x3fffffff) hint = 0x3fffffff;
  // (the sort sizes its CTAs for the number of instances it expects: the previous frame's
  // count of this thread when that fits the hint, else the hint itself)
  if (hint > 0)
    if (run_binning_and_render(a, fp, gv, iv, sp, (uint32_t)hint,
                               ds->last_count && ds->last_count <= hint ? ds->last_count
                                                                          : (uint32_t)hint, st))
      return 1;
  tr.lap("enqueue binning (hint)");
  CUDA_TRY(cudaEventSynchronize(ds->event));
  tr.lap("wait for count");
  const uint32_t N = ds->pinned[0], status = ds->pinned[1];
  if (status & 1u)
    return fail("Point is filtered although prefiltered is set. This shouldn't happen!");
  if (N > 0x3fffffffu) return fail("dssr_forward: %u instances exceed the 2^30 limit", N);
  ds->last_count = N;
  if (hint <= 0 || (int64_t)N > hint)
    if (run_binning_and_render(a, fp, gv, iv, sp, N, N, st)) return 1;
  tr.lap("enqueue binning (exact)");
  *num_rendered = (int)N;
  return 0;
}

int dssr_backward(const dssr_backward_args *a) {
  if (!a) return fail("dssr_backward: NULL argument");
  if (a->P < 0 || a->W <= 0 || a->H <= 0) return fail("dssr_backward: bad sizes");
  cudaStream_t st = (cudaStream_t)a->stream;
  const size_t P = (size_t)a->P;
  if (!a->dL_dmeans2D || !a->dL_dnormal || !a->dL_dopacity || !a->dL_dmeans3D ||
      !a->dL_dtransMat || !a->dL_dscales || !a->dL_drotations)
    return fail("dssr_backward: NULL gradient output");
  if (P == 0) return 0;
  if (!a->geom_buffer || !a->binning_buffer || !a->image_buffer || !a->dL_dout ||
      !a->radii || !a->viewmatrix || !a->projmatrix)
    return fail("dssr_backward: NULL input pointer");
  if (((uintptr_t)a->scales & 7) || ((uintptr_t)a->rotations & 15) ||
      ((uintptr_t)a->dL_dscales & 7) || ((uintptr_t)a->dL_drotations & 15) ||
      ((uintptr_t)a->dL_dtransMat & 15) || ((uintptr_t)a->packed_grads & 7))
    return fail("dssr_backward: scales / dL_dscales must be 8-byte and rotations / "
                "dL_drotations / dL_dtransMat 16-byte aligned");
  const bool dbg = a->debug & 1;
  // The four accumulation targets must start at zero: they are cleared - exactly these four
  // arrays, wherever the caller put them - by spare blocks of the work-order launch below.
  FillList fill;
  fill.ptr[0] = a->dL_dmeans2D;  fill.n[0] = 3 * P;
  fill.ptr[1] = a->dL_dnormal;   fill.n[1] = 3 * P;
  fill.ptr[2] = a->dL_dopacity;  fill.n[2] = P;
  fill.ptr[3] = a->dL_dtransMat; fill.n[3] = 12 * P;
  ImageLayout il(a->W, a->H);
  SortPlan sp(a->P, il.tiles);
  GeomView gv = make_geom_view(a->geom_buffer, a->P);
  ImageView iv = make_image_view(a->image_buffer, a->W, a->H);
  BinView bv = make_bin_view(a->binning_buffer, (size_t)a->R, 0, sp, il.tiles);  // only point_list and the header are used
  BwdParams bp;
  bp.P = a->P; bp.W = a->W; bp.H = a->H;
  bp.scales = a->scales; bp.rotations = a->rotations;
  bp.viewmatrix = a->viewmatrix; bp.projmatrix = a->projmatrix;
  bp.radii = a->radii; bp.dL_dout = a->dL_dout;
  bp.dL_dmeans2D = a->dL_dmeans2D; bp.dL_dnormal = a->dL_dnormal;
  bp.dL_dopacity = a->dL_dopacity; bp.dL_dmeans3D = a->dL_dmeans3D;
  bp.dL_dtransMat = a->dL_dtransMat; bp.dL_dscales = a->dL_dscales;
  bp.dL_drotations = a->dL_drotations;
  bp.packed = a->packed_grads;
  bp.packed_accumulate = a->packed_accumulate;
  XchgView xv;
  bp.push = nullptr;
  if (a->xchg) {
    if (xchg_begin_push(a->xchg, a->P, &xv)) return 1;
    bp.push = &xv;
  }
  {
    ProfScope ps("render_backward", st);
    // here, not in the forward: forward-only renders do not pay for it
    launch_bwd_order(iv, fill, st);
    launch_render_backward(bp, gv, iv, bv, st);
  }
  LAUNCH_CHECK(dbg, st, "render_backward");
  {
    ProfScope ps("backward_preprocess", st);
    launch_backward_preprocess(bp, st);
  }
  LAUNCH_CHECK(dbg, st, "backward_preprocess");
  return 0;
}

int dssr_mark_visible(int P, const float *means3D, const float *viewmatrix,
                      const float *projmatrix, unsigned char *present, void *stream) {
  (void)projmatrix;
  if (P < 0) return fail("dssr_mark_visible: P < 0");
  if (P == 0) return 0;
  if (!means3D || !viewmatrix || !present) return fail("dssr_mark_visible: NULL pointer");
  cudaStream_t st = (cudaStream_t)stream;
  launch_mark_visible(P, means3D, viewmatrix, present, st);
  LAUNCH_CHECK(false, st, "mark_visible");
  return 0;
}

int dssr_debug_state(const char *geom_buffer, const char *binning_buffer,
                     const char *image_buffer, const char *name, float *scratch,
                     void **ptr, size_t *count, void *stream) {
  if (!geom_buffer || !image_buffer || !name || !ptr || !count)
    return fail("dssr_debug_state: NULL argument");
  cudaStream_t st = (cudaStream_t)stream;
  GeomHeader gh;
  CUDA_TRY(cudaStreamSynchronize(st));
  CUDA_TRY(cudaMemcpy(&gh, align_ptr(geom_buffer), sizeof gh, cudaMemcpyDeviceToHost));
  if (gh.magic != 0x47454f31u) return fail("dssr_debug_state: not a geometry buffer");
  const int P = gh.P, W = gh.W, H = gh.H;
  const size_t R = gh.num_rendered, HW = (size_t)W * H;
  GeomView gv = make_geom_view(geom_buffer, P);
  ImageView iv = make_image_view(image_buffer, W, H);
  SortPlan sp(P, iv.tiles);
  BinView bv;
  memset(&bv, 0, sizeof bv);
  BinHeader bh;
  memset(&bh, 0, sizeof bh);
  if (binning_buffer) {
    CUDA_TRY(cudaMemcpy(&bh, align_ptr(binning_buffer), sizeof bh, cudaMemcpyDeviceToHost));
    if (bh.magic != 0x42494e31u) return fail("dssr_debug_state: not a binning buffer");
    bv = make_bin_view(binning_buffer, bh.capacity, bh.keep_unsorted, sp, iv.tiles);
  }
  const std::string n(name);
  if (n == "depths") { *ptr = gv.depths; *count = P; }
  else if (n == "means2D") { *ptr = gv.xy; *count = 2 * (size_t)P; }
  else if (n == "tiles_touched") { *ptr = gv.tiles; *count = P; }
  else if (n == "point_offsets") { *ptr = gv.offsets; *count = P; }
  else if (n == "transMat" || n == "normal_opacity") {
    if (!scratch) return fail("dssr_debug_state: %s needs scratch", name);
    launch_unpack_debug(gv, P, scratch, scratch + 12 * (size_t)P, st);
    CUDA_TRY(cudaStreamSynchronize(st));
    if (n == "transMat") { *ptr = scratch; *count = 12 * (size_t)P; }
    else { *ptr = scratch + 12 * (size_t)P; *count = 4 * (size_t)P; }
  }
  else if (n == "keys") {
    if (binning_buffer && bh.keys_deferred) {   // not materialised by the tile-first sort
      launch_rebuild_keys(iv, gv, bv, (uint32_t)P, bh.capacity, st);
      CUDA_TRY(cudaStreamSynchronize(st));
    }
    *ptr = bv.keys; *count = R;
  }
  else if (n == "point_list") { *ptr = bv.point_list; *count = R; }
  else if (n == "keys_unsorted") {
    if (!bv.keys_unsorted) return fail("dssr_debug_state: forward ran without debug bit1");
    *ptr = bv.keys_unsorted; *count = R;
  }
  else if (n == "ranges") { *ptr = iv.ranges; *count = 2 * (size_t)iv.tiles; }
  else if (n == "final_T") { *ptr = iv.final_T; *count = 3 * HW; }
  else if (n == "n_contrib") { *ptr = iv.n_contrib; *count = 2 * HW; }
  else return fail("dssr_debug_state: unknown array '%s'", name);
  if ((n == "keys" || n == "point_list") && !binning_buffer)
    return fail("dssr_debug_state: %s needs the binning buffer", name);
  return 0;
}

}  // extern "C"
