/*
 * oracle/oracle.c — TEST INFRASTRUCTURE.  A plain-C, CPU restatement of the reference's
 * algorithm for the hot path: the diff-surfel-spherical-rasterization forward and
 * backward and simple-knn's distCUDA2.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / reference arm may load this; the product never does.
 *
 * Parity status: the reference ships NO golden vectors or known-answer tests for this
 * path, so the oracle is pinned against outputs of the reference CUDA
 * build itself (oracle/_ref, run on the GPU box) stored under tests/golden/.  Because
 * the reference evaluates the pixel rays with the fast __sincosf intrinsics and
 * libdevice transcendentals, a CPU restatement agrees with it to ~1e-4 on the maps, not
 * bit for bit; bit-exactness is checked GPU-vs-GPU against oracle/_ref.
 *
 * Every function cites the reference source it restates; paths are relative to
 *   DSSR = submodules/diff-surfel-spherical-rasterization/cuda_rasterizer
 *   SKNN = submodules/simple-knn
 *
 * Build: gcc -O2 -fopenmp -shared -fPIC oracle.c -o _build/liboracle.so -lm
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define TILE 16
#define NEAR_N 0.1f
#define FAR_N 120.0f

typedef struct { float x, y, z; } v3;

static v3 v3make(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static float dot3(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static v3 cross3(v3 a, v3 b) {
  return v3make(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

/* getHigherMsb, DSSR/rasterizer_impl.cu:31-44 */
int orc_higher_msb(uint32_t n) {
  uint32_t msb = sizeof(n) * 4, step = msb;
  while (step > 1) {
    step /= 2;
    if (n >> msb) msb += step; else msb -= step;
  }
  if (n >> msb) msb++;
  return (int)msb;
}

/* quat_to_rotmat, DSSR/auxiliary.h:283-298; R[c][r] column-major */
static void quat_to_rot(const float *q, float R[3][3]) {
  float s = 1.0f / sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  float w = q[0] * s, x = q[1] * s, y = q[2] * s, z = q[3] * s;
  R[0][0] = 1.f - 2.f * (y * y + z * z); R[0][1] = 2.f * (x * y + w * z); R[0][2] = 2.f * (x * z - w * y);
  R[1][0] = 2.f * (x * y - w * z); R[1][1] = 1.f - 2.f * (x * x + z * z); R[1][2] = 2.f * (y * z + w * x);
  R[2][0] = 2.f * (x * z + w * y); R[2][1] = 2.f * (y * z - w * x); R[2][2] = 1.f - 2.f * (x * x + y * y);
}

/* getRectSpherical, DSSR/auxiliary.h:57-67 */
static void rect_spherical(float px, float py, int radius, int grid_y, int *x0, int *x1,
                           int *y0, int *y1) {
  int a;
  *x0 = ((int)(px - radius - TILE - 1)) / TILE;
  *x1 = ((int)(px + radius + TILE - 1)) / TILE;
  a = (int)((py - radius) / TILE); if (a < 0) a = 0; if (a > grid_y) a = grid_y; *y0 = a;
  a = (int)((py + radius + TILE - 1) / TILE); if (a < 0) a = 0; if (a > grid_y) a = grid_y; *y1 = a;
}

static void sph_project(v3 p, const float *proj, float *px, float *py) {
  /* xyz_to_sph, DSSR/auxiliary.h:350-353; projection DSSR/forward.cu:65-66 */
  float az = atan2f(p.y, p.x), el = atan2f(p.z, sqrtf(p.x * p.x + p.y * p.y));
  *px = az * proj[0] + proj[8];
  *py = el * proj[5] + proj[9];
}

/* preprocessCUDA (forward), DSSR/forward.cu:100-170 with in_frustum (auxiliary.h:262-279),
 * compute_transmat (forward.cu:19-43), compute_aabb (forward.cu:46-97).
 * T is written as the reference stores it: [P][12] = Tu.xyz 0 | Tv.xyz 0 | Tw.xyz 1. */
void orc_preprocess(int P, const float *means3D, const float *scales, const float *rots,
                    const float *opac, const float *precomp, const float *V, const float *proj,
                    int W, int H, float scale_mod, float *depths, int *radii, float *xy, float *T,
                    float *nrm_opa, uint32_t *tiles_touched) {
  const int grid_y = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; i++) {
    radii[i] = 0; tiles_touched[i] = 0; depths[i] = 0; xy[2 * i] = xy[2 * i + 1] = 0;
    memset(T + 12 * i, 0, 48); memset(nrm_opa + 4 * i, 0, 16);
    const float *p = means3D + 3 * i;
    /* transformPoint4x3, auxiliary.h:69-77 (V is column-major) */
    v3 pv = v3make(V[0] * p[0] + V[4] * p[1] + V[8] * p[2] + V[12],
                   V[1] * p[0] + V[5] * p[1] + V[9] * p[2] + V[13],
                   V[2] * p[0] + V[6] * p[1] + V[10] * p[2] + V[14]);
    float depth = sqrtf(dot3(pv, pv));
    if (depth <= NEAR_N || depth > FAR_N) continue;
    v3 Tu, Tv, Tw, nrm;
    if (!precomp) {
      float R[3][3];
      quat_to_rot(rots + 4 * i, R);
      float sx = scale_mod * scales[2 * i], sy = scale_mod * scales[2 * i + 1];
      v3 L0 = v3make(R[0][0] * sx, R[0][1] * sx, R[0][2] * sx);
      v3 L1 = v3make(R[1][0] * sy, R[1][1] * sy, R[1][2] * sy);
      v3 L2 = v3make(R[2][0], R[2][1], R[2][2]);
      Tu = v3make(V[0] * L0.x + V[4] * L0.y + V[8] * L0.z, V[1] * L0.x + V[5] * L0.y + V[9] * L0.z,
                  V[2] * L0.x + V[6] * L0.y + V[10] * L0.z);
      Tv = v3make(V[0] * L1.x + V[4] * L1.y + V[8] * L1.z, V[1] * L1.x + V[5] * L1.y + V[9] * L1.z,
                  V[2] * L1.x + V[6] * L1.y + V[10] * L1.z);
      Tw = pv;
      nrm = v3make(V[0] * L2.x + V[4] * L2.y + V[8] * L2.z, V[1] * L2.x + V[5] * L2.y + V[9] * L2.z,
                   V[2] * L2.x + V[6] * L2.y + V[10] * L2.z);
    } else {
      const float *t = precomp + 12 * i;
      Tu = v3make(t[0], t[1], t[2]); Tv = v3make(t[4], t[5], t[6]); Tw = v3make(t[8], t[9], t[10]);
      nrm = v3make(0, 0, 1);
    }
    if (-dot3(pv, nrm) <= 0.0f) continue;   /* BACKFACE_CULL, forward.cu:143-146 */
    float cx, cy;
    sph_project(Tw, proj, &cx, &cy);
    float delta_x = (float)(0.5 * W - (double)cx);
    float cnx = cx + delta_x, ext = 0;
    static const float KU[4] = {3, 3, -3, -3}, KV[4] = {3, -3, -3, 3};
    for (int k = 0; k < 4; k++) {
      v3 q = v3make(Tu.x * KU[k] + Tv.x * KV[k] + Tw.x, Tu.y * KU[k] + Tv.y * KV[k] + Tw.y,
                    Tu.z * KU[k] + Tv.z * KV[k] + Tw.z);
      float qx, qy;
      sph_project(q, proj, &qx, &qy);
      float qn = fmodf(fmodf(qx + delta_x, (float)W) + (float)W, (float)W);
      float dx = cnx - qn, dy = cy - qy, d = sqrtf(dx * dx + dy * dy);
      if (k == 0 || d > ext) ext = d;
    }
    float radius = ceilf(0.5f * ext) * 2;
    int x0, x1, y0, y1;
    rect_spherical(cx, cy, (int)radius, grid_y, &x0, &x1, &y0, &y1);
    if ((x1 - x0) * (y1 - y0) == 0) continue;
    depths[i] = depth; radii[i] = (int)radius; xy[2 * i] = cx; xy[2 * i + 1] = cy;
    float *t = T + 12 * i;
    t[0] = Tu.x; t[1] = Tu.y; t[2] = Tu.z; t[3] = 0;
    t[4] = Tv.x; t[5] = Tv.y; t[6] = Tv.z; t[7] = 0;
    t[8] = Tw.x; t[9] = Tw.y; t[10] = Tw.z; t[11] = 1;
    nrm_opa[4 * i] = nrm.x; nrm_opa[4 * i + 1] = nrm.y; nrm_opa[4 * i + 2] = nrm.z;
    nrm_opa[4 * i + 3] = opac[i];
    tiles_touched[i] = (uint32_t)((x1 - x0) * (y1 - y0));
  }
}

/* duplicateWithKeys, DSSR/rasterizer_impl.cu:62-89: emission order (idx, y, x).
 * Returns the number of instances written (== sum tiles_touched). */
long orc_build_keys(int P, const float *xy, const float *depths, const int *radii, int W, int H,
                    uint64_t *keys, uint32_t *vals) {
  const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
  long off = 0;
  for (int i = 0; i < P; i++) {
    if (radii[i] <= 0) continue;
    int x0, x1, y0, y1;
    rect_spherical(xy[2 * i], xy[2 * i + 1], radii[i], gy, &x0, &x1, &y0, &y1);
    uint32_t dbits;
    memcpy(&dbits, depths + i, 4);
    for (int y = y0; y < y1; y++)
      for (int x = x0; x < x1; x++) {
        int xw = ((x % gx) + gx) % gx;
        keys[off] = ((uint64_t)(y * gx + xw) << 32) | dbits;
        vals[off] = (uint32_t)i;
        off++;
      }
  }
  return off;
}

/* cub::DeviceRadixSort::SortPairs on bits [0, nbits): stable LSD radix, 8 bits a pass
 * (DSSR/rasterizer_impl.cu:232-238).  tmp arrays are caller-provided. */
void orc_sort_pairs(long N, uint64_t *keys, uint32_t *vals, uint64_t *ktmp, uint32_t *vtmp,
                    int nbits) {
  for (int sh = 0; sh < nbits; sh += 8) {
    long cnt[257] = {0};
    int bits = nbits - sh < 8 ? nbits - sh : 8;
    uint64_t mask = (1u << bits) - 1u;
    for (long i = 0; i < N; i++) cnt[((keys[i] >> sh) & mask) + 1]++;
    for (int d = 0; d < 256; d++) cnt[d + 1] += cnt[d];
    for (long i = 0; i < N; i++) {
      long p = cnt[(keys[i] >> sh) & mask]++;
      ktmp[p] = keys[i]; vtmp[p] = vals[i];
    }
    memcpy(keys, ktmp, (size_t)N * 8); memcpy(vals, vtmp, (size_t)N * 4);
  }
}

/* cudaMemset + identifyTileRanges, DSSR/rasterizer_impl.cu:94-114, 240-248 */
void orc_ranges(long N, const uint64_t *keys, int tiles, uint32_t *ranges) {
  memset(ranges, 0, (size_t)tiles * 8);
  for (long i = 0; i < N; i++) {
    uint32_t cur = (uint32_t)(keys[i] >> 32);
    if (i == 0) ranges[2 * cur] = 0;
    else {
      uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
      if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)i; ranges[2 * cur] = (uint32_t)i; }
    }
    if (i == N - 1) ranges[2 * cur + 1] = (uint32_t)N;
  }
}

typedef struct {
  v3 xyz, pi_az, pi_el;
} ray_t;

/* per-pixel ray set-up, DSSR/forward.cu:259-264 (sph_to_xyz: auxiliary.h:360-366; the
 * reference uses __sincosf / rsqrtf here, the oracle the precise libm functions) */
static ray_t make_ray(float px, float py, const float *proj) {
  ray_t r;
  float az = (1.f / proj[0]) * (px - proj[8]), el = (1.f / proj[5]) * (py - proj[9]);
  r.xyz = v3make(cosf(az) * cosf(el), sinf(az) * cosf(el), sinf(el));
  float inorm = 1.0f / sqrtf(r.xyz.y * r.xyz.y + r.xyz.x * r.xyz.x);
  r.pi_az = v3make(r.xyz.y * inorm, -r.xyz.x * inorm, 0.0f);
  r.pi_el = cross3(r.pi_az, r.xyz);
  return r;
}

typedef struct {
  v3 k, l, p, beta;
  float sx, sy, rho3d, rho2d, dx, dy, depth, depth2d, G, alpha;
} hit_t;

/* the per-(pixel, surfel) evaluation shared by forward.cu:254-301 and backward.cu:136-182;
 * returns 0 when the reference `continue`s */
static int intersect(const ray_t *r, float px, float py, const float *T, const float *cxy,
                     float opa, hit_t *h) {
  v3 Tu = v3make(T[0], T[1], T[2]), Tv = v3make(T[4], T[5], T[6]), Tw = v3make(T[8], T[9], T[10]);
  h->k = v3make(dot3(Tu, r->pi_az), dot3(Tv, r->pi_az), dot3(Tw, r->pi_az));
  h->l = v3make(dot3(Tu, r->pi_el), dot3(Tv, r->pi_el), dot3(Tw, r->pi_el));
  h->p = cross3(h->k, h->l);
  if (h->p.z == 0.0f) return 0;
  h->sx = h->p.x / h->p.z; h->sy = h->p.y / h->p.z;
  h->rho3d = h->sx * h->sx + h->sy * h->sy;
  h->dx = cxy[0] - px; h->dy = cxy[1] - py;
  h->rho2d = 2.0f * (h->dx * h->dx + h->dy * h->dy);   /* FilterInvSquare */
  h->beta = v3make(h->sx * Tu.x + h->sy * Tv.x + Tw.x, h->sx * Tu.y + h->sy * Tv.y + Tw.y,
                   h->sx * Tu.z + h->sy * Tv.z + Tw.z);
  if (dot3(h->beta, r->xyz) < 0.0f) return 0;
  h->depth2d = sqrtf(dot3(Tw, Tw));
  h->depth = (h->rho3d <= h->rho2d) ? sqrtf(dot3(h->beta, h->beta)) : h->depth2d;
  if (h->depth < NEAR_N) return 0;
  float rho = fminf(h->rho3d, h->rho2d);
  float power = -0.5f * rho;
  if (power > 0.0f) return 0;
  h->G = expf(power);
  h->alpha = fminf(0.99f, opa * h->G);
  if (h->alpha < 1.0f / 255.0f) return 0;
  return 1;
}

/* renderCUDA (forward), DSSR/forward.cu:172-351.  out = [7][H][W]; final_T = [3][H][W]
 * (T, M1, M2); n_contrib = [2][H][W] (last, median; "no median" is stored as 0, which is
 * what the float -1 -> uint32 conversion of forward.cu:221,348 yields on NVIDIA GPUs). */
void orc_forward(int W, int H, const uint32_t *ranges, const uint32_t *point_list,
                 const float *xy, const float *T, const float *nrm_opa, const float *proj,
                 float *final_T, uint32_t *n_contrib, float *out) {
  const int gx = (W + TILE - 1) / TILE;
  const size_t HW = (size_t)W * H;
#pragma omp parallel for schedule(dynamic, 16)
  for (long pix = 0; pix < (long)HW; pix++) {
    int py = (int)(pix / W), px = (int)(pix % W);
    int tile = (py / TILE) * gx + px / TILE;
    uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
    ray_t ray = make_ray((float)px, (float)py, proj);
    float Tr = 1.0f, D = 0, N[3] = {0, 0, 0}, M1 = 0, M2 = 0, dist = 0, med = 0;
    uint32_t contributor = 0, last = 0, medc = 0;
    for (uint32_t e = r0; e < r1; e++) {
      contributor++;
      uint32_t id = point_list[e];
      hit_t h;
      if (!intersect(&ray, (float)px, (float)py, T + 12 * (size_t)id, xy + 2 * (size_t)id,
                     nrm_opa[4 * (size_t)id + 3], &h))
        continue;
      float test_T = Tr * (1 - h.alpha);
      if (test_T < 0.0001f) break;   /* done = true: nothing further is blended */
      float w = h.alpha * Tr, A = 1 - Tr, m = h.depth;
      dist += (m * m * A + M2 - 2 * m * M1) * w;
      M1 += m * w; M2 += m * m * w; D += h.depth * w;
      if (Tr > 0.5f) { med = h.depth; medc = contributor; }
      for (int c = 0; c < 3; c++) N[c] += nrm_opa[4 * (size_t)id + c] * w;
      Tr = test_T; last = contributor;
    }
    final_T[pix] = Tr; final_T[pix + HW] = M1; final_T[pix + 2 * HW] = M2;
    n_contrib[pix] = last; n_contrib[pix + HW] = medc;
    out[pix] = D; out[pix + HW] = 1 - Tr;
    out[pix + 2 * HW] = N[0]; out[pix + 3 * HW] = N[1]; out[pix + 4 * HW] = N[2];
    out[pix + 5 * HW] = med; out[pix + 6 * HW] = dist;
  }
}

static void atomic_add_d(double *p, double v) {
#pragma omp atomic
  *p += v;
}

/* renderCUDA (backward), DSSR/backward.cu:19-301, quirks kept as written (SURVEY.md §8
 * a10, Appendix A).  Accumulators are double (the reference's are float atomics whose
 * order is not deterministic). dT = [P][12], dN = [P][3], dO = [P]. */
void orc_backward_render(int W, int H, const uint32_t *ranges, const uint32_t *point_list,
                         const float *xy, const float *T, const float *nrm_opa, const float *proj,
                         const float *final_T, const uint32_t *n_contrib, const float *dL_dout,
                         double *dT, double *dN, double *dO) {
  const int gx = (W + TILE - 1) / TILE;
  const size_t HW = (size_t)W * H;
  const float fx = proj[0], fy = proj[5];
#pragma omp parallel for schedule(dynamic, 16)
  for (long pix = 0; pix < (long)HW; pix++) {
    int py = (int)(pix / W), px = (int)(pix % W);
    int tile = (py / TILE) * gx + px / TILE;
    uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
    ray_t ray = make_ray((float)px, (float)py, proj);
    const float T_final = final_T[pix];
    float Tr = T_final;
    const uint32_t last_contributor = n_contrib[pix], median_contributor = n_contrib[pix + HW];
    const float gD = dL_dout[pix], gA = dL_dout[pix + HW], gMed = dL_dout[pix + 5 * HW],
                gReg = dL_dout[pix + 6 * HW];
    const float gN[3] = {dL_dout[pix + 2 * HW], dL_dout[pix + 3 * HW], dL_dout[pix + 4 * HW]};
    const float final_D = final_T[pix + HW], final_D2 = final_T[pix + 2 * HW], final_A = 1 - T_final;
    float accum_depth_rec = 0, accum_alpha_rec = 0, accum_normal_rec[3] = {0, 0, 0};
    float last_alpha = 0, last_depth = 0, last_normal[3] = {0, 0, 0}, last_dL_dT = 0;
    uint32_t contributor = r1 - r0;
    for (uint32_t e = r1; e-- > r0;) {
      contributor--;
      if (contributor >= last_contributor) continue;
      uint32_t id = point_list[e];
      const float *Ts = T + 12 * (size_t)id;
      const float *no = nrm_opa + 4 * (size_t)id;
      hit_t h;
      if (!intersect(&ray, (float)px, (float)py, Ts, xy + 2 * (size_t)id, no[3], &h)) continue;
      const float alpha = h.alpha, depth = h.depth;
      Tr = Tr / (1.f - alpha);
      const float dchannel_dcolor = alpha * Tr;
      float dL_dalpha = 0, dLd_depth = 0;
      const float m_d = depth;
      float dL_dweight = (final_D2 + m_d * m_d * final_A - 2 * m_d * final_D) * gReg;
      dL_dalpha += dL_dweight - last_dL_dT;
      last_dL_dT = dL_dweight * alpha + (1 - alpha) * last_dL_dT;
      dLd_depth += 2.0f * (Tr * alpha) * (m_d * final_A - final_D) * gReg;
      if (contributor == median_contributor - 1u) dLd_depth += gMed;
      dLd_depth += dchannel_dcolor * gD;
      accum_depth_rec = last_alpha * last_depth + (1.f - last_alpha) * accum_depth_rec;
      last_depth = depth;
      dL_dalpha += (depth - accum_depth_rec) * gD;
      for (int c = 0; c < 3; c++) {
        accum_normal_rec[c] = last_alpha * last_normal[c] + (1.f - last_alpha) * accum_normal_rec[c];
        last_normal[c] = no[c];
        dL_dalpha += (no[c] - accum_normal_rec[c]) * gN[c];
        atomic_add_d(&dN[3 * (size_t)id + c], (double)(alpha * Tr * gN[c]));
      }
      accum_alpha_rec = last_alpha * 1.0f + (1.f - last_alpha) * accum_alpha_rec;
      dL_dalpha += (1 - accum_alpha_rec) * gA;
      dL_dalpha *= Tr;
      last_alpha = alpha;
      atomic_add_d(&dO[id], (double)(h.G * dL_dalpha));          /* Q1 */
      v3 Tu = v3make(Ts[0], Ts[1], Ts[2]), Tv = v3make(Ts[4], Ts[5], Ts[6]),
         Tw = v3make(Ts[8], Ts[9], Ts[10]);
      if (h.rho3d <= h.rho2d) {
        float bTu = dot3(Tu, h.beta), bTv = dot3(Tv, h.beta);
        float dsx = -alpha * dL_dalpha * h.sx + bTu * alpha * Tr * dLd_depth / depth;   /* Q2 */
        float dsy = -alpha * dL_dalpha * h.sy + bTv * alpha * Tr * dLd_depth / depth;
        v3 dp = v3make(dsx / h.p.z, dsy / h.p.z, -(dsx * h.p.x + dsy * h.p.y) / (h.p.z * h.p.z));
        v3 k1 = cross3(dp, h.k), k2 = cross3(dp, h.l);
        float cu = dLd_depth * h.sx / depth, cv = dLd_depth * h.sy / depth, cw = dLd_depth / depth;
        double *d = dT + 12 * (size_t)id;
        atomic_add_d(d + 0, k1.x * ray.pi_el.x - k2.x * ray.pi_az.x + cu * h.beta.x);
        atomic_add_d(d + 1, k1.x * ray.pi_el.y - k2.x * ray.pi_az.y + cu * h.beta.y);
        atomic_add_d(d + 2, k1.x * ray.pi_el.z - k2.x * ray.pi_az.z + cu * h.beta.z);
        atomic_add_d(d + 4, k1.y * ray.pi_el.x - k2.y * ray.pi_az.x + cv * h.beta.x);
        atomic_add_d(d + 5, k1.y * ray.pi_el.y - k2.y * ray.pi_az.y + cv * h.beta.y);
        atomic_add_d(d + 6, k1.y * ray.pi_el.z - k2.y * ray.pi_az.z + cv * h.beta.z);
        atomic_add_d(d + 8, k1.z * ray.pi_el.x - k2.z * ray.pi_az.x + cw * h.beta.x);
        atomic_add_d(d + 9, k1.z * ray.pi_el.y - k2.z * ray.pi_az.y + cw * h.beta.y);
        atomic_add_d(d + 10, k1.z * ray.pi_el.z - k2.z * ray.pi_az.z + cw * h.beta.z);
      } else {
        float ddx = -2 * alpha * dL_dalpha * 2.0f * h.dx, ddy = -2 * alpha * dL_dalpha * 2.0f * h.dy; /* Q7 */
        float ss = Tw.x * Tw.x + Tw.y * Tw.y, sn = sqrtf(ss), d2 = h.depth2d;
        float gx_ = fx * ddx * (-Tw.y / ss) + fy * ddy * (-Tw.x * Tw.z / (sn * d2));  /* Q3 */
        float gy_ = fx * ddx * (Tw.x / ss) + fy * ddy * (-Tw.y * Tw.z / (sn * d2));
        float gz_ = fy * ddy * sn / d2;
        float cw = dLd_depth / d2;
        double *d = dT + 12 * (size_t)id;
        atomic_add_d(d + 8, gx_ + cw * Tw.x);
        atomic_add_d(d + 9, gy_ + cw * Tw.y);
        atomic_add_d(d + 10, gz_ + cw * Tw.z);
      }
    }
  }
}

/* preprocessCUDA (backward) / compute_transmat_aabb, DSSR/backward.cu:303-396 and
 * quat_to_rotmat_vjp, DSSR/auxiliary.h:300-339 */
void orc_backward_preprocess(int P, const int *radii, const float *scales, const float *rots,
                             const float *V, const float *dT, const float *dNrm, float *dmeans,
                             float *dscales, float *drots) {
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; i++) {
    memset(dmeans + 3 * i, 0, 12); memset(dscales + 2 * i, 0, 8); memset(drots + 4 * i, 0, 16);
    if (!(radii[i] > 0)) continue;
    const float *d = dT + 12 * (size_t)i;
    v3 du = v3make(d[0], d[1], d[2]), dv = v3make(d[4], d[5], d[6]), dw = v3make(d[8], d[9], d[10]);
    dmeans[3 * i] = V[0] * dw.x + V[1] * dw.y + V[2] * dw.z;
    dmeans[3 * i + 1] = V[4] * dw.x + V[5] * dw.y + V[6] * dw.z;
    dmeans[3 * i + 2] = V[8] * dw.x + V[9] * dw.y + V[10] * dw.z;
    if (!scales || !rots) continue;
    float R[3][3];
    const float *q = rots + 4 * i;
    quat_to_rot(q, R);
    float sx = scales[2 * i], sy = scales[2 * i + 1];   /* Q5: no scale_modifier */
    float c[3][3] = {{sx * du.x, sx * du.y, sx * du.z}, {sy * dv.x, sy * dv.y, sy * dv.z},
                     {dNrm[3 * i], dNrm[3 * i + 1], dNrm[3 * i + 2]}};
    float vR[3][3];
    for (int col = 0; col < 3; col++)
      for (int r = 0; r < 3; r++)
        vR[col][r] = V[4 * r] * c[col][0] + V[4 * r + 1] * c[col][1] + V[4 * r + 2] * c[col][2];
    v3 a = v3make(V[0] * R[0][0] + V[4] * R[0][1] + V[8] * R[0][2], V[1] * R[0][0] + V[5] * R[0][1] + V[9] * R[0][2],
                  V[2] * R[0][0] + V[6] * R[0][1] + V[10] * R[0][2]);
    v3 b = v3make(V[0] * R[1][0] + V[4] * R[1][1] + V[8] * R[1][2], V[1] * R[1][0] + V[5] * R[1][1] + V[9] * R[1][2],
                  V[2] * R[1][0] + V[6] * R[1][1] + V[10] * R[1][2]);
    dscales[2 * i] = dot3(du, a);
    dscales[2 * i + 1] = dot3(dv, b);
    float s = 1.0f / sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    float w = q[0] * s, x = q[1] * s, y = q[2] * s, z = q[3] * s;   /* Q6 */
    drots[4 * i] = 2.f * (x * (vR[1][2] - vR[2][1]) + y * (vR[2][0] - vR[0][2]) + z * (vR[0][1] - vR[1][0]));
    drots[4 * i + 1] = 2.f * (-2.f * x * (vR[1][1] + vR[2][2]) + y * (vR[0][1] + vR[1][0]) +
                              z * (vR[0][2] + vR[2][0]) + w * (vR[1][2] - vR[2][1]));
    drots[4 * i + 2] = 2.f * (x * (vR[0][1] + vR[1][0]) - 2.f * y * (vR[0][0] + vR[2][2]) +
                              z * (vR[1][2] + vR[2][1]) + w * (vR[2][0] - vR[0][2]));
    drots[4 * i + 3] = 2.f * (x * (vR[0][2] + vR[2][0]) + y * (vR[1][2] + vR[2][1]) -
                              2.f * z * (vR[0][0] + vR[1][1]) + w * (vR[0][1] - vR[1][0]));
  }
}

/* checkFrustum, DSSR/rasterizer_impl.cu:48-58 */
void orc_mark_visible(int P, const float *means3D, const float *V, uint8_t *present) {
  for (int i = 0; i < P; i++) {
    const float *p = means3D + 3 * i;
    v3 pv = v3make(V[0] * p[0] + V[4] * p[1] + V[8] * p[2] + V[12], V[1] * p[0] + V[5] * p[1] + V[9] * p[2] + V[13],
                   V[2] * p[0] + V[6] * p[1] + V[10] * p[2] + V[14]);
    float d = sqrtf(dot3(pv, pv));
    present[i] = !(d <= NEAR_N || d > FAR_N);
  }
}

/* distCUDA2 / SimpleKNN::knn, SKNN/simple_knn.cu:186-222.  The reference's Morton boxes
 * only prune (conservatively), so its result is the exact 3-NN: mean of the three
 * smallest squared distances (updateKBest<3>, :132-146), FLT_MAX-initialised
 * (:155), self excluded by position (:164-166).  Brute force over points [q0, q1). */
void orc_knn_range(int P, const float *pts, int q0, int q1, float *out) {
#pragma omp parallel for schedule(dynamic, 64)
  for (int i = q0; i < q1; i++) {
    float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;
    const float *q = pts + 3 * (size_t)i;
    for (int j = 0; j < P; j++) {
      if (j == i) continue;
      const float *p = pts + 3 * (size_t)j;
      float dx = p[0] - q[0], dy = p[1] - q[1], dz = p[2] - q[2];
      float dist = fmaf(dz, dz, fmaf(dx, dx, dy * dy));   /* the compiled form of :135 */
      if (b0 > dist) { float t = b0; b0 = dist; dist = t; }
      if (b1 > dist) { float t = b1; b1 = dist; dist = t; }
      if (b2 > dist) b2 = dist;
    }
    out[i] = (b0 + b1 + b2) / 3.0f;
  }
}

/* Developer statistics (not a reference restatement): for every 8x4 pixel sub-tile, over the
 * list prefix its pixels traverse, count entries / entries with at least one pixel passing
 * the alpha test / passing (pixel, entry) pairs.  stats[0..3] = traversed, true_hit, pairs,
 * entries whose box (xmin,xmax,ymin,ymax per surfel, may be NULL) overlaps the sub-tile. */
void orc_pair_stats(int W, int H, const uint32_t *ranges, const uint32_t *point_list,
                    const float *xy, const float *T, const float *nrm_opa, const float *proj,
                    const uint32_t *n_contrib, const float *boxes, double *stats) {
  const int gx = (W + TILE - 1) / TILE;
  double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : s0, s1, s2, s3)
  for (int st = 0; st < ((W + 7) / 8) * ((H + 3) / 4); st++) {
    int sx = (st % ((W + 7) / 8)) * 8, sy = (st / ((W + 7) / 8)) * 4;
    int tile = (sy / TILE) * gx + sx / TILE;
    uint32_t r0 = ranges[2 * tile];
    uint32_t trav = 0;
    ray_t rays[32];
    for (int l = 0; l < 32; l++) {
      int px = sx + (l & 7), py = sy + (l >> 3);
      rays[l] = make_ray((float)px, (float)py, proj);
      if (px < W && py < H && n_contrib[(size_t)py * W + px] > trav) trav = n_contrib[(size_t)py * W + px];
    }
    for (uint32_t e = 0; e < trav; e++) {
      uint32_t id = point_list[r0 + e];
      int any = 0;
      for (int l = 0; l < 32; l++) {
        int px = sx + (l & 7), py = sy + (l >> 3);
        if (px >= W || py >= H) continue;
        hit_t h;
        if (intersect(&rays[l], (float)px, (float)py, T + 12 * (size_t)id, xy + 2 * (size_t)id,
                      nrm_opa[4 * (size_t)id + 3], &h)) { any = 1; s2 += 1; }
      }
      s0 += 1; s1 += any;
      if (boxes) {
        const float *b = boxes + 4 * (size_t)id;
        if (!(b[2] > sy + 3) && !(b[3] < sy) && !(b[0] > sx + 7) && !(b[1] < sx)) s3 += 1;
      }
    }
  }
  stats[0] = s0; stats[1] = s1; stats[2] = s2; stats[3] = s3;
}

/* Developer statistics per 4x4 sub-tile: out[st*7+0..6] = traversed entries, entries whose
 * box overlaps the sub-tile, entries with at least one passing pixel, sum over the 32-entry
 * steps of ceil(visited in step / 2) (the two-at-a-time evaluation steps), steps with a visit, (pixel, visited
 * entry) pairs whose pixel lies inside the entry's box, pairs that pass the alpha test. */
void orc_subtile_stats(int W, int H, const uint32_t *ranges, const uint32_t *point_list,
                       const float *xy, const float *T, const float *nrm_opa, const float *proj,
                       const uint32_t *n_contrib, const float *boxes, uint32_t *out) {
  const int gx = (W + TILE - 1) / TILE;
  const int nsx = (W + 3) / 4, nsy = (H + 3) / 4;
#pragma omp parallel for schedule(dynamic, 4)
  for (int st = 0; st < nsx * nsy; st++) {
    int sx = (st % nsx) * 4, sy = (st / nsx) * 4;
    int tile = (sy / TILE) * gx + sx / TILE;
    uint32_t r0 = ranges[2 * tile], trav = 0, nb = 0, nt = 0, in_step = 0, pairs = 0, busy = 0, inbox = 0, okpix = 0;
    ray_t rays[16];
    for (int l = 0; l < 16; l++) {
      int px = sx + (l & 3), py = sy + (l >> 2);
      rays[l] = make_ray((float)px, (float)py, proj);
      if (px < W && py < H && n_contrib[(size_t)py * W + px] > trav) trav = n_contrib[(size_t)py * W + px];
    }
    for (uint32_t e = 0; e < trav; e++) {
      if ((e & 31u) == 0u) { pairs += (in_step + 1) / 2; busy += in_step > 0; in_step = 0; }
      uint32_t id = point_list[r0 + e];
      const float *b = boxes + 4 * (size_t)id;
      /* x is periodic with the azimuth period 2*pi*|fx| (as raster_math.cuh:cull_overlap) */
      const double ip = 1.0 / (6.283185307179586 * fabs((double)proj[0]));
      if (b[2] > sy + 3 || b[3] < sy ||
          floor(((double)(sx + 3) - b[0]) * ip) < ceil(((double)sx - b[1]) * ip)) continue;
      nb++;
      in_step++;
      int any = 0;
      for (int l = 0; l < 16; l++) {
        int px = sx + (l & 3), py = sy + (l >> 2);
        hit_t h;
        if (!(b[2] > py || b[3] < py || floor(((double)px - b[0]) * ip) < ceil(((double)px - b[1]) * ip)))
          inbox++;
        if (px < W && py < H && intersect(&rays[l], (float)px, (float)py, T + 12 * (size_t)id,
                                          xy + 2 * (size_t)id, nrm_opa[4 * (size_t)id + 3], &h)) {
          any = 1;
          okpix++;
        }
      }
      nt += any;
    }
    pairs += (in_step + 1) / 2; busy += in_step > 0;
    out[7 * st] = trav; out[7 * st + 1] = nb; out[7 * st + 2] = nt; out[7 * st + 3] = pairs; out[7 * st + 4] = busy;
    out[7 * st + 5] = inbox; out[7 * st + 6] = okpix;
  }
}
