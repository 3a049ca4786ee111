// raster_math.cuh — the per-pixel ray set-up and per-(pixel, surfel) intersection maths
// shared by the forward and backward compositing kernels.
//
// Follows renderCUDA, submodules/diff-surfel-spherical-rasterization/cuda_rasterizer/
// forward.cu:249-331 (and the identical re-computation in backward.cu:135-182).  Every
// value that feeds a discrete decision (skip tests, early termination, median, branch
// between the 3D and the 2D-filter depth) is computed with explicitly rounded single
// operations in the order nvcc compiles the reference's expressions to, so those
// decisions — and with them the forward maps — are bit-identical to the reference.
#pragma once
#include "common.cuh"

struct PixelRay {
  float X, Y, Z;      // ray direction (cos az cos el, sin az cos el, sin el)
  float ax, ay;       // pi_az = (ax, ay, 0)
  float ex, ey, ez;   // pi_el = pi_az x dir
  float pxf, pyf;     // integer pixel coordinates as floats (no +0.5, forward.cu:190)
};

// forward.cu:225-226, 259-264
__device__ __forceinline__ PixelRay make_ray(float pxf, float pyf, float fx, float fy,
                                             float cx, float cy) {
  PixelRay r;
  r.pxf = pxf;
  r.pyf = pyf;
  const float i_fx = 1.0f / fx, i_fy = 1.0f / fy;
  float az = FM(FS(pxf, cx), i_fx);
  float el = FM(FS(pyf, cy), i_fy);
  float sa, ca, se, ce;
  __sincosf(az, &sa, &ca);   // sph_to_xyz(float2) uses the fast intrinsics (auxiliary.h:360-366)
  __sincosf(el, &se, &ce);
  r.X = FM(ca, ce);
  r.Y = FM(sa, ce);
  r.Z = se;
  float inorm = rsqrtf(FF(r.Y, r.Y, FM(r.X, r.X)));
  r.ax = FM(r.Y, inorm);
  r.ay = FM(r.X, -inorm);
  r.ex = FF(r.Z, r.ay, -FM(0.0f, r.Y));
  r.ey = FF(0.0f, r.X, -FM(r.Z, r.ax));
  r.ez = FF(r.Y, r.ax, -FM(r.X, r.ay));
  return r;
}

struct SurfelRec {   // the 64-byte packed record written by the preprocess kernel
  float tux, tuy, tuz, tvx, tvy, tvz, twx, twy, twz, nx, ny, nz, opa, cx, cy, d2d;
};

__device__ __forceinline__ SurfelRec unpack_rec(float4 a, float4 b, float4 c, float4 d) {
  SurfelRec s;
  s.tux = a.x; s.tuy = a.y; s.tuz = a.z; s.tvx = a.w;
  s.tvy = b.x; s.tvz = b.y; s.twx = b.z; s.twy = b.w;
  s.twz = c.x; s.nx = c.y; s.ny = c.z; s.nz = c.w;
  s.opa = d.x; s.cx = d.y; s.cy = d.z; s.d2d = d.w;
  return s;
}

struct Hit {
  float kx, ky, kz, lx, ly, lz;   // transformed planes
  float px, py, pz;               // k x l
  float sx, sy;                   // intersection in splat coordinates
  float rho3d, rho2d, dx, dy;
  float bx, by, bz;               // intersection point in the sensor frame
  float depth, G, alpha;
};

// First half of the evaluation: the transformed planes, their cross product and the
// pixel-space distance (forward.cu:265-277).  Returns false when this lane provably ends in
// one of the reference's `continue`s: p.z == 0, or both rho3d = |p.xy|^2 / p.z^2 and rho2d
// exceed the surfel's cutoff, i.e. alpha < 1/255.  The test uses the very floats the exact
// evaluation goes on with; NaNs compare false and fall through to the exact path.
__device__ __forceinline__ bool intersect_begin(const PixelRay &r, const SurfelRec &s, float cut,
                                                Hit &h) {
  h.kx = FF(0.0f, s.tuz, FF(s.tux, r.ax, FM(s.tuy, r.ay)));
  h.ky = FF(0.0f, s.tvz, FF(s.tvx, r.ax, FM(s.tvy, r.ay)));
  h.kz = FF(0.0f, s.twz, FF(s.twx, r.ax, FM(s.twy, r.ay)));
  h.lx = FF(s.tuz, r.ez, FF(s.tux, r.ex, FM(s.tuy, r.ey)));
  h.ly = FF(s.tvz, r.ez, FF(s.tvx, r.ex, FM(s.tvy, r.ey)));
  h.lz = FF(s.twz, r.ez, FF(s.twx, r.ex, FM(s.twy, r.ey)));
  h.pz = FF(h.kx, h.ly, -FM(h.ky, h.lx));
  h.px = FF(h.ky, h.lz, -FM(h.kz, h.ly));
  h.py = FF(h.kz, h.lx, -FM(h.kx, h.lz));
  h.dx = FS(s.cx, r.pxf);
  h.dy = FS(s.cy, r.pyf);
  float t = FF(h.dy, h.dy, FM(h.dx, h.dx));
  h.rho2d = FA(t, t);   // FilterInvSquare = 2
  if (h.pz == 0.0f) return false;
  const bool far3d = h.px * h.px + h.py * h.py > cut * (h.pz * h.pz);
  return !(far3d && h.rho2d > cut);
}

// Second half (forward.cu:273-301).  Returns false when the reference would `continue`.
__device__ __forceinline__ bool intersect_finish(const PixelRay &r, const SurfelRec &s, Hit &h) {
  h.sx = __fdiv_rn(h.px, h.pz);
  h.sy = __fdiv_rn(h.py, h.pz);
  h.rho3d = FF(h.sx, h.sx, FM(h.sy, h.sy));
  h.bx = FA(s.twx, FF(s.tux, h.sx, FM(s.tvx, h.sy)));
  h.by = FA(s.twy, FF(s.tuy, h.sx, FM(s.tvy, h.sy)));
  h.bz = FA(s.twz, FF(s.tuz, h.sx, FM(s.tvz, h.sy)));
  float valid = FF(r.Z, h.bz, FF(r.X, h.bx, FM(r.Y, h.by)));
  if (valid < 0.0f) return false;
  h.depth = (h.rho3d <= h.rho2d)
                ? __fsqrt_rn(FF(h.bz, h.bz, FF(h.bx, h.bx, FM(h.by, h.by))))
                : s.d2d;
  if (h.depth < SPLAT_NEAR) return false;
  float power = FM(fminf(h.rho3d, h.rho2d), -0.5f);
  if (power > 0.0f) return false;
  h.G = expf(power);
  h.alpha = fminf(0.99f, FM(s.opa, h.G));
  // forward: (double)alpha < 1.0/255.0; backward: alpha < 1.0f/255.0f.  No float lies
  // between the two constants, so both are "alpha < 0x3B808081".
  if (h.alpha < __uint_as_float(SPLAT_ALPHA_MIN)) return false;
  return true;
}

// Both halves without a branch, for callers that know the pair passes (the backward composites
// exactly the pairs the forward blended): the same operations on the same values in the same
// order - only the early exits are gone - so every field equals what intersect_begin +
// intersect_finish leave behind.  Returns what they would have returned.
__device__ __forceinline__ bool intersect_all(const PixelRay &r, const SurfelRec &s, Hit &h) {
  h.kx = FF(0.0f, s.tuz, FF(s.tux, r.ax, FM(s.tuy, r.ay)));
  h.ky = FF(0.0f, s.tvz, FF(s.tvx, r.ax, FM(s.tvy, r.ay)));
  h.kz = FF(0.0f, s.twz, FF(s.twx, r.ax, FM(s.twy, r.ay)));
  h.lx = FF(s.tuz, r.ez, FF(s.tux, r.ex, FM(s.tuy, r.ey)));
  h.ly = FF(s.tvz, r.ez, FF(s.tvx, r.ex, FM(s.tvy, r.ey)));
  h.lz = FF(s.twz, r.ez, FF(s.twx, r.ex, FM(s.twy, r.ey)));
  h.pz = FF(h.kx, h.ly, -FM(h.ky, h.lx));
  h.px = FF(h.ky, h.lz, -FM(h.kz, h.ly));
  h.py = FF(h.kz, h.lx, -FM(h.kx, h.lz));
  h.dx = FS(s.cx, r.pxf);
  h.dy = FS(s.cy, r.pyf);
  const float t = FF(h.dy, h.dy, FM(h.dx, h.dx));
  h.rho2d = FA(t, t);
  h.sx = __fdiv_rn(h.px, h.pz);
  h.sy = __fdiv_rn(h.py, h.pz);
  h.rho3d = FF(h.sx, h.sx, FM(h.sy, h.sy));
  h.bx = FA(s.twx, FF(s.tux, h.sx, FM(s.tvx, h.sy)));
  h.by = FA(s.twy, FF(s.tuy, h.sx, FM(s.tvy, h.sy)));
  h.bz = FA(s.twz, FF(s.tuz, h.sx, FM(s.tvz, h.sy)));
  const float valid = FF(r.Z, h.bz, FF(r.X, h.bx, FM(r.Y, h.by)));
  const float d3 = __fsqrt_rn(FF(h.bz, h.bz, FF(h.bx, h.bx, FM(h.by, h.by))));
  h.depth = (h.rho3d <= h.rho2d) ? d3 : s.d2d;
  const float power = FM(fminf(h.rho3d, h.rho2d), -0.5f);
  h.G = expf(power);
  h.alpha = fminf(0.99f, FM(s.opa, h.G));
  return !(h.pz == 0.0f) && !(valid < 0.0f) && !(h.depth < SPLAT_NEAR) && !(power > 0.0f) &&
         !(h.alpha < __uint_as_float(SPLAT_ALPHA_MIN));
}

// Does the surfel's cull box (see dssr_preprocess.cu:cull_box) touch the pixel
// rectangle [x0,x1] x [y0,y1]?  x is periodic with the azimuth period.  Written so
// that any NaN answers "yes".
__device__ __forceinline__ bool cull_overlap(float4 cb, float x0, float x1, float y0,
                                             float y1, float inv_period) {
  return !(cb.z > y1) && !(cb.w < y0) &&
         !(floorf((x1 - cb.x) * inv_period) < ceilf((x0 - cb.y) * inv_period));
}

// Which of the 16 4x4-pixel sub-tiles of the 16x16 tile at (tx0, ty0) can the surfel's cull
// box touch?  bit (sy*4 + sx).  Same NaN-means-yes convention as cull_overlap.
// (rows and columns separately: a caller that walks a row of tiles computes the rows once)
__device__ __forceinline__ uint32_t subtile_rows(float4 cb, float ty0) {
  uint32_t rm = 0;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const float y0 = ty0 + 4.0f * i;
    if (!(cb.z > y0 + 3.0f) && !(cb.w < y0)) rm |= 1u << i;
  }
  return rm;
}
__device__ __forceinline__ uint32_t subtile_cols(float4 cb, float tx0, float inv_period) {
  uint32_t cm = 0;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const float x0 = tx0 + 4.0f * i;
    if (!(floorf((x0 + 3.0f - cb.x) * inv_period) < ceilf((x0 - cb.y) * inv_period))) cm |= 1u << i;
  }
  return cm;
}
__device__ __forceinline__ uint32_t subtile_combine(uint32_t rm, uint32_t cm) {
  return (cm * 0x1111u) & (((rm & 1u) * 0xfu) | ((rm & 2u) * 0x78u) | ((rm & 4u) * 0x3c0u) |
                           ((rm & 8u) * 0x1e00u));
}
__device__ __forceinline__ uint32_t subtile_mask(float4 cb, float tx0, float ty0,
                                                 float inv_period) {
  return subtile_combine(subtile_rows(cb, ty0), subtile_cols(cb, tx0, inv_period));
}
