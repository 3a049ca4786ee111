/*
 * splat_b200.h — C-ABI of the B200-native (sm_100a) surfel spherical rasterizer and
 * distCUDA2 kNN scale initialiser.
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch / C++ types.
 * Every DEVICE pointer must live on the device that is current when the call is
 * made; every kernel is launched on the `stream` passed in (a cudaStream_t cast to
 * void*; NULL = legacy default stream, where the reference launches everything).
 *
 * Each entry point cites the reference interface it replaces (paths relative to the
 * reference repo, DSSR = submodules/diff-surfel-spherical-rasterization,
 * SKNN = submodules/simple-knn).
 *
 * Return value: 0 on success, non-zero on error; the message is available from
 * splat_last_error() (thread-local).  The Python binding turns that into
 * RuntimeError, like TORCH_CHECK / CHECK_CUDA do in the reference
 * (DSSR/rasterize_points.cu:21-22, DSSR/cuda_rasterizer/auxiliary.h:368-377).
 */
#ifndef SPLAT_B200_H_
#define SPLAT_B200_H_

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define SPLAT_API __attribute__((visibility("default")))
#else
#define SPLAT_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* Growable-buffer callback: returns a DEVICE pointer to at least `bytes` bytes that
 * stays valid until the matching backward has run.  Mirrors
 * std::function<char*(size_t)> (DSSR/cuda_rasterizer/rasterizer.h:10-12) and the
 * resize lambdas of DSSR/rasterize_points.cu:24-30. */
typedef char *(*splat_alloc_fn)(void *ctx, size_t bytes);

/* ---- rasterizer forward --------------------------------------------------------
 * Replaces CudaRasterizer::Rasterizer::forward
 * (DSSR/cuda_rasterizer/rasterizer.h:12-21, rasterizer_impl.cu:170-262) as called
 * by RasterizeGaussiansCUDA (DSSR/rasterize_points.cu:32-88). */
typedef struct {
  int P, W, H;                   /* surfels, image width, image height             */
  float scale_modifier;
  int prefiltered;               /* as reference: a culled surfel is then an error */
  int debug;                     /* bit0: synchronise + check after every launch
                                    (CHECK_CUDA); bit1: keep a copy of the unsorted
                                    keys for dssr_debug_state                      */
  const float *means3D;          /* [P,3]                                          */
  const float *opacities;        /* [P]                                            */
  const float *scales;           /* [P,2] or NULL when transMat_precomp is given   */
  const float *rotations;        /* [P,4] (w,x,y,z) or NULL                        */
  const float *transMat_precomp; /* [P,12] or NULL                                 */
  const float *viewmatrix;       /* [16] device, column-major world->sensor        */
  const float *projmatrix;       /* [16] device; fx=[0] fy=[5] cx=[8] cy=[9]       */
  float *out_maps;               /* [7,H,W] written in full                        */
  int *radii;                    /* [P] written in full                            */
  splat_alloc_fn geom_alloc;   void *geom_ctx;
  splat_alloc_fn binning_alloc; void *binning_ctx;
  splat_alloc_fn image_alloc;  void *image_ctx;
  /* Optional: instance capacity to pre-allocate for the binning buffer (0 = size it
   * exactly after the instance count is known, like the reference).  When the hint
   * is large enough the whole frame is enqueued before the host waits for the
   * count, so the GPU never idles on the host. */
  int64_t binning_hint;
  void *stream;
} dssr_forward_args;

/* *num_rendered receives the number of (surfel, tile) instances, the value the
 * reference returns (rasterizer_impl.cu:261). */
SPLAT_API int dssr_forward(const dssr_forward_args *args, int *num_rendered);

/* ---- rasterizer backward -------------------------------------------------------
 * Replaces CudaRasterizer::Rasterizer::backward
 * (DSSR/cuda_rasterizer/rasterizer.h:23-31, rasterizer_impl.cu:266-313) as called
 * by RasterizeGaussiansBackwardCUDA (DSSR/rasterize_points.cu:90-147).  All seven
 * gradient arrays are written in full (zeros where the reference leaves its
 * torch::zeros untouched). */
typedef struct {
  int P, W, H, R;                /* R = num_rendered of the forward                */
  float scale_modifier;
  int debug;
  const float *means3D, *scales, *rotations, *transMat_precomp;
  const float *viewmatrix, *projmatrix;
  const int *radii;              /* [P] from the forward                           */
  const char *geom_buffer, *binning_buffer, *image_buffer; /* from the forward     */
  const float *dL_dout;          /* [7,H,W]                                        */
  float *dL_dmeans2D;            /* [P,3] (always zero, as in the reference)       */
  float *dL_dnormal;             /* [P,3] scratch / output                         */
  float *dL_dopacity;            /* [P]                                            */
  float *dL_dmeans3D;            /* [P,3]                                          */
  float *dL_dtransMat;           /* [P,12]                                         */
  float *dL_dscales;             /* [P,2]                                          */
  float *dL_drotations;          /* [P,4]                                          */
  /* Optional (NULL = off; no counterpart in the reference): a contiguous [P,10] buffer
   * (dL_dmeans3D 3 | dL_dopacity 1 | dL_dscales 2 | dL_drotations 4 per surfel, 8-byte
   * aligned) that receives the same gradients, overwritten or - with packed_accumulate -
   * added to.  It is the message of the keyframe-window all-reduce: a rank accumulates
   * all of its keyframes into it without any extra kernel. */
  float *packed_grads;
  int packed_accumulate;
  /* Optional (NULL = off): a connected gradient exchange (splat_xchg_*, below).  The
   * per-surfel VJP kernel then also stores every [10] row - plus, with packed_accumulate,
   * what packed_grads held - into the staging slot of the rank that owns the row, over
   * NVLink, while it computes: the reduce-scatter half of the window all-reduce costs no
   * kernel and no extra pass over HBM.  Follow with splat_xchg_allreduce. */
  struct splat_xchg *xchg;
  void *stream;
} dssr_backward_args;

SPLAT_API int dssr_backward(const dssr_backward_args *args);

/* Replaces CudaRasterizer::Rasterizer::markVisible
 * (DSSR/cuda_rasterizer/rasterizer_impl.cu:117-122, checkFrustum :48-58). */
SPLAT_API int dssr_mark_visible(int P, const float *means3D, const float *viewmatrix,
                      const float *projmatrix, unsigned char *present,
                      void *stream);

/* ---- map post-processing of render() (the caller side) --------
 * Replaces what gaussian_renderer/__init__.py:51-92 and utils/graphic_utils.py:26-88
 * (depth_to_points, depth_to_normal) do to the rasterizer's [7,H,W] output with ~25 PyTorch
 * kernels forward and ~50 backward, by one kernel per direction:
 *   rend_normal [3,H,W] = world_view_transform[:3,:3] . normal map, / alpha where alpha > 0
 *   surf_depth  [1,H,W] = (depth / alpha where alpha > 0)(1 - depth_ratio) + median depth * depth_ratio
 *   surf_normal [3,H,W] = alpha * normalised cross product of the central differences of the
 *                         back-projected surf_depth (zero on the one-pixel border)
 * rend_alpha and rend_dist of render() are channels 1 and 6 of allmap itself.  viewmatrix /
 * projmatrix are the camera's world_view_transform / projection_matrix (device, 16 floats).
 * The backward takes the upstream gradients of the three maps (any may be NULL = zero) and
 * writes dL/dallmap [7,H,W] in full (channel 6 = 0; add the gradients that reach channels
 * 1 and 6 directly through rend_alpha / rend_dist yourself, as autograd does for views). */
typedef struct {
  int W, H;
  float depth_ratio;
  const float *allmap;
  const float *viewmatrix, *projmatrix;
  float *rend_normal, *surf_depth, *surf_normal;
  void *stream;
} dssr_post_forward_args;
typedef struct {
  int W, H;
  float depth_ratio;
  const float *allmap;
  const float *viewmatrix, *projmatrix;
  const float *dL_drend_normal, *dL_dsurf_depth, *dL_dsurf_normal;
  float *dL_dallmap;
  void *stream;
} dssr_post_backward_args;
SPLAT_API int dssr_post_forward(const dssr_post_forward_args *args);
SPLAT_API int dssr_post_backward(const dssr_post_backward_args *args);

/* The same two kernels with the mapper's per-keyframe image losses folded in
 * (slam/mapper.py:158-190; inline code there, no function to mirror):
 *   losses[0] geom_l1 = mean over H*W of |valid (surf_depth - gt_depth)|
 *   losses[1] normal  = lambda_normal * mean over valid of (1 - <rend_normal, surf_normal>)
 *   losses[2] alpha   = lambda_alpha * mean over valid of BCE(rend_alpha, 1)
 *   losses[3] total   = [0] + [2] + [1]        losses[4] = number of valid pixels
 * with valid <=> gt_valid == 1 (camera.image_valid, uint8).  The forward never writes the maps
 * unless asked to (rend_normal / surf_depth / surf_normal may be NULL); the backward derives
 * the maps' upstream gradients from the loss definition instead of reading them, and takes
 * dL_dlosses (device, 5 floats, or NULL = d total only) so nothing is synchronised with the
 * host.  `workspace`: DSSR_POST_LOSS_WORKSPACE_BYTES device bytes, 8-byte aligned, zeroed once
 * by the caller; every launch leaves it zeroed.  The scale regulariser of mapper.py:186-191
 * acts on the parameters, not on the image, and stays with the caller. */
#define DSSR_POST_LOSS_WORKSPACE_BYTES 64
typedef struct {
  int W, H;
  float depth_ratio;
  const float *allmap;
  const float *viewmatrix, *projmatrix;
  const float *gt_depth;             /* [H,W] camera.image_depth                     */
  const unsigned char *gt_valid;     /* [H,W] camera.image_valid                     */
  float lambda_normal, lambda_alpha;
  float *rend_normal, *surf_depth, *surf_normal;   /* optional map outputs           */
  float *losses;                     /* [5] device                                   */
  void *workspace;
  void *stream;
} dssr_post_loss_forward_args;
typedef struct {
  int W, H;
  float depth_ratio;
  const float *allmap;
  const float *viewmatrix, *projmatrix;
  const float *gt_depth;
  const unsigned char *gt_valid;
  float lambda_normal, lambda_alpha;
  const float *losses;               /* [5] from the forward                         */
  const float *dL_dlosses;           /* [5] device, or NULL                          */
  float *dL_dallmap;                 /* [7,H,W], written in full                     */
  void *stream;
} dssr_post_loss_backward_args;
SPLAT_API int dssr_post_loss_forward(const dssr_post_loss_forward_args *args);
SPLAT_API int dssr_post_loss_backward(const dssr_post_loss_backward_args *args);

/* ---- activation backward + Adam in one pass -----------------
 * Replaces, for the four parameter groups of scene/gaussian_model.py:97-121 (xyz, opacity,
 * scaling, rotation - arrays indexed in that order below): autograd's backward through the
 * activations (gaussian_model.py:58-74: identity, sigmoid, exp, F.normalize), the backward of
 * the scale regulariser (slam/mapper.py:186-191; scaling_max_penalty = 0 turns it off) and
 * torch.optim.Adam.step() (no weight decay, no amsgrad).  `packed_grads` is the [P,10] buffer
 * of dssr_backward / splat_xchg_result: gradients w.r.t. the ACTIVATED values.  param /
 * exp_avg / exp_avg_sq are the raw parameters and the optimizer's state tensors, updated in
 * place; step[k] is the step count AFTER this update (>= 1), as torch increments it first.
 * reg_out (optional, device float) is incremented by the regulariser's value. */
typedef struct {
  int P;
  const float *packed_grads;
  float *param[4], *exp_avg[4], *exp_avg_sq[4];
  double lr[4];
  long long step[4];
  double beta1, beta2, eps;
  double scaling_max, scaling_max_penalty;
  float *reg_out;
  void *stream;
} splat_adam_args;
SPLAT_API int splat_activation_adam_step(const splat_adam_args *args);

/* Buffer sizes the callbacks will be asked for (mirrors required<T>(),
 * DSSR/cuda_rasterizer/rasterizer_impl.h:59-64).  The binning buffer takes 126 bytes of
 * address space per instance of capacity (the reference: 28); most of it is the worst-case
 * room of the per-block lists, of which a frame touches about a fifth. */
SPLAT_API size_t dssr_geometry_bytes(int P);
SPLAT_API size_t dssr_image_bytes(int W, int H);
SPLAT_API size_t dssr_binning_bytes(int64_t capacity, int W, int H, int P, int keep_unsorted);

/* Debug accessor for parity tests: device pointer + element count of an internal
 * array of the LAST forward that used these buffers.  Names: "depths" f32[P],
 * "means2D" f32[2P], "transMat" f32[12P] (Tu.xyz0 Tv.xyz0 Tw.xyz1),
 * "normal_opacity" f32[4P], "tiles_touched" u32[P], "point_offsets" u32[P],
 * "keys_unsorted" u64[R] (debug bit1), "keys" u64[R], "point_list" u32[R],
 * "ranges" u32[2*tiles], "final_T" f32[3HW], "n_contrib" u32[2HW].
 * The reference has no such call; its tests re-derive the same arrays with
 * GeometryState/BinningState/ImageState::fromChunk (rasterizer_impl.h:18-56).
 * "transMat" and "normal_opacity" are materialised into `scratch` (device, at least
 * 16*P floats) because the product stores them in a packed record.  "keys" are
 * materialised on request too when the tile-first sort produced the frame (images of
 * up to 1024 tiles and 2 M surfels): that sort never writes point_list_keys, every key
 * is (tile << 32) | bits(depths[point_list[i]]) and is rebuilt from those arrays. */
SPLAT_API int dssr_debug_state(const char *geom_buffer, const char *binning_buffer,
                     const char *image_buffer, const char *name, float *scratch,
                     void **ptr, size_t *count, void *stream);

/* ---- distCUDA2 -----------------------------------------------------------------
 * Replaces SimpleKNN::knn (SKNN/simple_knn.cu:186-222) as called by distCUDA2
 * (SKNN/spatial.cu:15-26): mean of the squared distances to the 3 nearest
 * neighbours of every point.  `workspace` is caller-owned device scratch of at
 * least sknn_workspace_bytes(P) bytes. */
SPLAT_API size_t sknn_workspace_bytes(int P);
SPLAT_API int sknn_dist2(int P, const float *points, float *mean_dists, void *workspace,
               size_t workspace_bytes, void *stream);

/* distCUDA2 for callers that cannot allocate device memory (no counterpart in the
 * reference, which cudaMallocs inside every call: SKNN/simple_knn.cu:188-221).  Meant for
 * hosts without an allocator of their own: a latency-bound sensor or tracking thread
 * (cudaMalloc / cudaFree serialise the device), a process whose GPU memory already belongs
 * to another framework's caching allocator, a plug-in that may not make allocating CUDA
 * calls.  Inputs and output as sknn_dist2; the scratch is a file-scope device array inside
 * the library (4 MB on every GPU that loads it, held until exit), so:
 *   - P must not exceed SKNN_STATIC_MAX_POINTS, larger clouds are refused with an error;
 *   - the first call on a device resolves the array's address on that device; make it
 *     before any stream capture begins;
 *   - all calls share the array (it keeps the previous call's contents; what has to start
 *     at zero is cleared at the start of each call): calls whose GPU work overlaps race,
 *     serialise them;
 *   - no cudaMalloc, cudaFree or allocating library routine is called. */
#define SKNN_STATIC_MAX_POINTS 65536
SPLAT_API int sknn_dist2_static(int P, const float *points, float *mean_dists, void *stream);

/* ---- keyframe-window gradient exchange over NVLink peer memory -------------------
 * No counterpart in the reference (it maps on one GPU, slam/mapper.py:150-156).  One
 * process per GPU; every rank creates an exchange for the same P and world, the ranks swap
 * the 64-byte handles (any transport: torch.distributed, MPI, a file) and connect.  Then
 * per optimisation step:
 *     dssr_backward(... xchg ...)   or   splat_xchg_push(x, packed [P,10] or NULL = zeros)
 *     splat_xchg_allreduce(x, stream)          (= splat_xchg_reduce + splat_xchg_wait)
 * after which splat_xchg_result(x) - a [P,10] device array owned by the exchange - holds the
 * sum over ranks, bit-identical on every rank (each row is summed once, by its owner, in
 * rank order).  Work is ordered by `stream`; the kernels wait for peers through flags in
 * peer memory.  A wait gives up after 30 s (SPLAT_B200_XCHG_TIMEOUT_MS overrides, read by
 * splat_xchg_create); the rank that gave up marks EVERY rank's arena before it goes on, so
 * by the end of the same exchange every rank holds the fault: splat_xchg_fault() returns
 * non-zero once the stream has been synchronised, and every later splat_xchg_* call fails.
 * The sums of a faulted exchange contain stale rows - do not apply them; destroy and
 * re-create the exchange.  All ranks must make the same sequence of calls.  world <= 16;
 * peers must be NVLink/P2P reachable.
 * splat_xchg_connect_local wires exchanges that live in ONE process (tests, or a process
 * driving several GPUs; it enables peer access between the devices involved); with it, when
 * several ranks share a stream, issue every rank's reduce before any rank's wait. */
#define SPLAT_XCHG_HANDLE_BYTES 64
typedef struct splat_xchg splat_xchg;
SPLAT_API int splat_xchg_create(int P, int world, int rank, splat_xchg **out);
SPLAT_API int splat_xchg_export(splat_xchg *x, unsigned char *handle /* [64] */);
SPLAT_API int splat_xchg_connect(splat_xchg *x, const unsigned char *handles /* [world][64] */);
SPLAT_API int splat_xchg_connect_local(splat_xchg *x, splat_xchg *const *peers /* [world] */);
SPLAT_API int splat_xchg_push(splat_xchg *x, const float *packed, void *stream);
SPLAT_API int splat_xchg_reduce(splat_xchg *x, void *stream);
SPLAT_API int splat_xchg_wait(splat_xchg *x, void *stream);
SPLAT_API int splat_xchg_allreduce(splat_xchg *x, void *stream);
SPLAT_API float *splat_xchg_result(splat_xchg *x);
/* 0, or the fault code (1 + step + 16 * rank waited for) a wait of this or a peer rank left
 * behind.  Reads a host word: synchronise the stream of the last allreduce first. */
SPLAT_API int splat_xchg_fault(splat_xchg *x);
SPLAT_API int splat_xchg_destroy(splat_xchg *x);

/* ---- per-stage device timing ----------------------------------------------------
 * No counterpart in the reference (its only timing is wall-clock prints in
 * DSSR/tests/test_backward.py:54-85).  While enabled (per host thread), dssr_forward /
 * dssr_backward bracket each stage with CUDA events on the launch stream.  After the
 * caller has synchronised that stream, splat_profile_read returns the stages seen since
 * the last read: their static names and the elapsed milliseconds SUMMED over all calls
 * since then; the return value is the stage count.  Stages: preprocess, emit, sort,
 * render_forward, render_backward, backward_preprocess.  splat_profile_select restricts
 * the bracketing to one stage (NULL: all) - a pair of events costs a few microseconds of
 * launch overlap, so a throughput measurement instruments only the kernel it reports. */
SPLAT_API int splat_profile_enable(int on);
SPLAT_API int splat_profile_select(const char *stage);
SPLAT_API int splat_profile_read(int max_stages, const char **names, float *ms);

/* ---- pruning / growing the parameter tensors and their optimizer moments -----------
 * Replaces the tensor surgery of scene/gaussian_model.py:223-316 (_prune_optimizer,
 * prune_points, cat_tensors_to_optimizer, densification_postfix), which indexes / concatenates
 * every parameter and every Adam moment separately.  Up to SPLAT_SURGERY_MAX_TENSORS row-major
 * float tensors that share their row index (row width width[t] floats) go through one kernel.
 *   splat_compact_plan   scans keep[P] (non-zero = keep) into `workspace`
 *                        (splat_compact_workspace_bytes(P) bytes, device) and returns the
 *                        number of kept rows in *count (synchronises `stream`: the caller
 *                        sizes the new tensors with it - the one host round trip);
 *   splat_compact_rows   dst[t][r] = src[t][i] for the r-th kept row i (same keep, workspace);
 *   splat_append_rows    dst[t] = [src[t] (n_old rows); ext[t] (n_new rows)], ext == NULL or
 *                        ext[t] == NULL appends zeros (fresh moments).
 * dst must not alias src. */
#define SPLAT_SURGERY_MAX_TENSORS 12
SPLAT_API size_t splat_compact_workspace_bytes(int P);
SPLAT_API int splat_compact_plan(int P, const unsigned char *keep, void *workspace, int *count,
                                 void *stream);
SPLAT_API int splat_compact_rows(int P, const unsigned char *keep, const void *workspace,
                                 int ntensors, const float *const *src, float *const *dst,
                                 const int *width, void *stream);
SPLAT_API int splat_append_rows(int n_old, int n_new, int ntensors, const float *const *src,
                                const float *const *ext, float *const *dst, const int *width,
                                void *stream);

/* ---- photometric (range-image) registration -------------------------------------------
 * Replaces the photometric stage of the reference's gsaligner submodule as Splat-LOAM's
 * tracker drives it (slam/tracker.py:160-188 -> GSAligner::setReference / setQuery /
 * alignPhotometric, submodules/gsaligner/gsaligner/aligner.cu:181-258,329-342 ->
 * photo::PhotometricFac::compute, registration/photometric.cpp:55-77 and
 * registration/photometric_cuda.cu:8-395; image preparation tools/image_set_impl.cu:20-120,
 * 254-376): spherical camera, depth channel, identity sensor offset.
 *   gsa_photo_workspace_bytes  size of the caller-owned device workspace for an H x W image;
 *   gsa_photo_set_reference    reference range image (device, [H*W], > 1e-5 = measured):
 *                              mask, mask growing by `mask_grow_radius`, clamped central
 *                              differences (threshold `depth_derivative_threshold`);
 *   gsa_photo_set_query        query range image -> point cloud in the sensor frame;
 *   gsa_photo_align            `iterations` Gauss-Newton steps from `initial_guess` (HOST,
 *                              row-major 4x4 ref_T_query): per step one z-buffer projection and
 *                              one linearise + reduce + solve kernel, pose kept on the device,
 *                              no host round trip in between.  Outputs (HOST): transform
 *                              (row-major 4x4), stats = {residual, inlier ratio} as
 *                              PhotometricFac::compute returns them, and optionally `system`
 *                              = the sums of the LAST linearisation: upper triangle of H (21,
 *                              row by row) | b (6) | chi | good | inliers.  Synchronises
 *                              `stream` before it returns. */
typedef struct {
  int H, W;
  float fx, fy, cx, cy;               /* spherical camera matrix K (projmatrix[:3,:3]^T)  */
  float min_depth, max_depth;         /* photo_min_depth, photo_max_depth                 */
  float omega_depth;                  /* photo_omega_depth                                */
  float depth_rejection_threshold;    /* photo_depth_rejection_threshold                  */
  float kernel_chi_threshold;         /* photo_rho_kernel                                 */
  void *stream;
} gsa_photo_params;
SPLAT_API size_t gsa_photo_workspace_bytes(int H, int W);
SPLAT_API int gsa_photo_set_reference(void *workspace, int H, int W, const float *depth,
                                      float depth_derivative_threshold, int mask_grow_radius,
                                      void *stream);
SPLAT_API int gsa_photo_set_query(void *workspace, int H, int W, const float *depth, float fx,
                                  float fy, float cx, float cy, float min_depth, float max_depth,
                                  void *stream);
/* parity-test accessor: the prepared reference image as four [H*W] float planes on the
 * device: depth | d/dcol | d/drow | masked (1 / 0) */
SPLAT_API int gsa_photo_debug_reference(void *workspace, int H, int W, float *out, void *stream);
SPLAT_API int gsa_photo_align(void *workspace, const gsa_photo_params *p,
                              const float *initial_guess, int iterations, float *transform,
                              float *stats, double *system);

/* ---- range-image front end -------------------------------------------------------------
 * Replaces Preprocessor.__call__ (scene/preprocessing.py:24-83, numpy + the `pyprojections`
 * package on the CPU): a LiDAR cloud ([P,3] device) becomes the H x W spherical range image,
 * normal image ([3,H,W], -p/|p| of the pixel's point) and valid mask of a Camera.
 * PARITY UNPINNED: `pyprojections` (PyPI 0.0.3, pixi.toml:39) is not vendored with the reference
 * and not in this image; the conventions (documented in csrc/range_image.cu) are the ones
 * consistent with the camera matrix the reference ships in gsaligner/tests/data/K_ref.npy.
 *   rimg_spherical_intrinsics  calculate_spherical_intrinsics: K from the cloud's azimuth /
 *                              elevation extent; kept in the workspace and, if K_out != NULL,
 *                              written as a row-major 3x3 (device);
 *   rimg_project               Camera.project + the gathers: nearest return per pixel (ties to
 *                              the smaller index); K = NULL uses the workspace's intrinsics,
 *                              else a row-major 3x3 on the device; lut (int64 point index, -1
 *                              where empty) is optional. */
SPLAT_API size_t rimg_workspace_bytes(int H, int W);
SPLAT_API int rimg_spherical_intrinsics(int P, const float *cloud, int H, int W, void *workspace,
                                        float *K_out, void *stream);
SPLAT_API int rimg_project(int P, const float *cloud, int H, int W, const float *K,
                           float depth_min, float depth_max, void *workspace, long long *lut,
                           float *range_image, float *normal_image, unsigned char *valid,
                           void *stream);

/* Number of kernels this library has launched so far in this process (all threads, all
 * devices).  No counterpart in the reference.  A harness takes the difference around a timed
 * region to report how many of the library's kernels ran in it. */
SPLAT_API unsigned long long splat_launch_count(void);

SPLAT_API const char *splat_last_error(void);
SPLAT_API int splat_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SPLAT_B200_H_ */
