// dssr_render_fwd.cu — forward tile compositing.
//
// Replaces renderCUDA<3>, submodules/diff-surfel-spherical-rasterization/
// cuda_rasterizer/forward.cu:172-351 (launch :353-362).
//
// The reference runs one 256-thread CTA per 16x16 tile; every pixel walks the tile's
// whole depth-sorted list and the CTA leaves only when all 256 pixels are done.  On a
// 64x1024 range image that is 256 CTAs on 148 SMs.  Here the unit of work is a WARP
// that owns an 8x4 pixel sub-tile and walks its tile's list on its own:
//   * 32 list entries are fetched per step (one per lane, coalesced), each lane tests
//     its entry's conservative cull box against the sub-tile, a ballot compacts the
//     survivors, whose 64-byte records are gathered (128-bit loads, L2-resident) into
//     a warp-private shared-memory ring;
//   * the survivors are then composited in list order by all 32 pixels;
//   * the warp stops as soon as ITS 32 pixels are done.
// Skipping an entry whose cull box misses the sub-tile is exact: such an entry fails
// the reference's alpha >= 1/255 test at every pixel of the sub-tile, i.e. it is one
// the reference `continue`s over without touching any state.  `contributor` keeps
// counting list positions, so n_contrib / median_contributor are the reference's.
#include "raster_math.cuh"

namespace {

constexpr int FW_WARPS = 4;            // warps per CTA: a 16x8 pixel half tile
constexpr int SUB_W = 8, SUB_H = 4;    // pixels per warp

__global__ void __launch_bounds__(FW_WARPS * 32)
render_forward_kernel(const uint2 *__restrict__ ranges,
                      const uint32_t *__restrict__ point_list,
                      const float4 *__restrict__ rec, const float4 *__restrict__ cull,
                      const float *__restrict__ projmatrix, int W, int H, int tiles_x,
                      uint32_t capacity, uint32_t P, float *__restrict__ final_T,
                      uint32_t *__restrict__ n_contrib,
                      float *__restrict__ out) {
  __shared__ float4 s_rec[FW_WARPS][32][4];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int x0 = blockIdx.x * SPLAT_TILE + (wid & 1) * SUB_W;
  const int y0 = blockIdx.y * (2 * SUB_H) + (wid >> 1) * SUB_H;
  const int px = x0 + (lane & 7), py = y0 + (lane >> 3);
  const bool inside = px < W && py < H;
  const float fx = projmatrix[0], fy = projmatrix[5], cx = projmatrix[8], cy = projmatrix[9];
  const PixelRay ray = make_ray((float)px, (float)py, fx, fy, cx, cy);
  const float inv_period = 1.0f / (6.2831853f * fabsf(fx));
  // culling assumes |elevation| < pi/2 on every image row
  const float el_a = fabsf((0.0f - cy) / fy), el_b = fabsf(((float)(H - 1) - cy) / fy);
  const bool cull_on = fmaxf(el_a, el_b) < 1.55f;
  const float bx0 = (float)x0, bx1 = (float)(x0 + SUB_W - 1);
  const float by0 = (float)y0, by1 = (float)(y0 + SUB_H - 1);

  uint2 range = ranges[(y0 / SPLAT_TILE) * tiles_x + blockIdx.x];
  range.x = min(range.x, capacity);   // only bites when the capacity hint was too small
  range.y = min(range.y, capacity);
  const uint32_t lt = (1u << lane) - 1u;

  float T = 1.0f, D = 0.0f, N0 = 0.0f, N1 = 0.0f, N2 = 0.0f, M1 = 0.0f, M2 = 0.0f;
  float dist = 0.0f, median_depth = 0.0f;
  uint32_t last_contributor = 0, median_contributor = 0;
  bool done = !inside;

  // software prefetch of the next 32 entries
  uint32_t nid = 0;
  float4 ncb = make_float4(0.f, 0.f, 0.f, 0.f);
  // (id < P only fails on the garbage a too-small capacity hint leaves behind; that
  // frame is re-run with the exact size)
  bool nhave = range.x + lane < range.y;
  if (nhave) {
    nid = point_list[range.x + lane];
    nhave = nid < P;
    if (nhave) ncb = cull[nid];
  }
  for (uint32_t base = range.x; base < range.y; base += 32) {
    if (__all_sync(0xffffffffu, done)) break;
    const uint32_t id = nid;
    const float4 cb = ncb;
    const bool have = nhave;
    {
      uint32_t ni = base + 32 + lane;
      nhave = ni < range.y;
      if (nhave) {
        nid = point_list[ni];
        nhave = nid < P;
        if (nhave) ncb = cull[nid];
      }
    }
    const bool hit = have && (!cull_on || cull_overlap(cb, bx0, bx1, by0, by1, inv_period));
    uint32_t m = __ballot_sync(0xffffffffu, hit);
    if (m == 0) continue;
    if (hit) {
      const int slot = __popc(m & lt);
      const float4 *r = rec + 4 * (size_t)id;
      float4 a = __ldg(r), b = __ldg(r + 1), c = __ldg(r + 2), d = __ldg(r + 3);
      s_rec[wid][slot][0] = a;
      s_rec[wid][slot][1] = b;
      s_rec[wid][slot][2] = c;
      s_rec[wid][slot][3] = d;
    }
    __syncwarp();
    int slot = 0;
    while (m) {
      const int j = __ffs(m) - 1;
      m &= m - 1;
      const SurfelRec s = unpack_rec(s_rec[wid][slot][0], s_rec[wid][slot][1],
                                     s_rec[wid][slot][2], s_rec[wid][slot][3]);
      slot++;
      if (!done) {
        Hit h;
        if (intersect(ray, s, h)) {
          const uint32_t contributor = base - range.x + j + 1;
          const float test_T = FM(T, FS(1.0f, h.alpha));
          if (test_T < 0.0001f) {
            done = true;
          } else {
            const float w = FM(T, h.alpha);
            const float A = FS(1.0f, T);
            const float mm = FM(h.depth, h.depth), m2 = FA(h.depth, h.depth);
            float tmp = FF(A, mm, M2);
            tmp = FF(-M1, m2, tmp);
            dist = FF(w, tmp, dist);
            M2 = FF(w, mm, M2);
            M1 = FF(h.depth, w, M1);
            D = FF(h.depth, w, D);
            if (T > 0.5f) {
              median_depth = h.depth;
              median_contributor = contributor;
            }
            N0 = FF(s.nx, w, N0);
            N1 = FF(s.ny, w, N1);
            N2 = FF(s.nz, w, N2);
            T = test_T;
            last_contributor = contributor;
          }
        }
      }
      if (__all_sync(0xffffffffu, done)) break;
    }
    __syncwarp();
  }

  if (inside) {
    const size_t HW = (size_t)W * H;
    const size_t pix = (size_t)py * W + px;
    final_T[pix] = T;
    final_T[pix + HW] = M1;
    final_T[pix + 2 * HW] = M2;
    n_contrib[pix] = last_contributor;
    n_contrib[pix + HW] = median_contributor;
    out[pix] = D;                      // DEPTH_OFFSET
    out[pix + HW] = FS(1.0f, T);       // ALPHA_OFFSET
    out[pix + 2 * HW] = N0;            // NORMAL_OFFSET..+2
    out[pix + 3 * HW] = N1;
    out[pix + 4 * HW] = N2;
    out[pix + 5 * HW] = median_depth;  // MIDDEPTH_OFFSET
    out[pix + 6 * HW] = dist;          // DISTORTION_OFFSET
  }
}

}  // namespace

void launch_render_forward(const FwdParams &p, const GeomView &g, const ImageView &im,
                           const BinView &b, uint32_t capacity, cudaStream_t s) {
  dim3 grid(im.tiles_x, (p.H + 2 * SUB_H - 1) / (2 * SUB_H));
  render_forward_kernel<<<grid, FW_WARPS * 32, 0, s>>>(
      im.ranges, b.point_list, g.rec, g.cull, p.projmatrix, p.W, p.H, im.tiles_x,
      capacity, (uint32_t)p.P, im.final_T, im.n_contrib, p.out_maps);
}
