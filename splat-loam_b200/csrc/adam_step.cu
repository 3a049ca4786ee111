// adam_step.cu — one kernel for what follows the gradient exchange in a mapping iteration:
// the backward of the parameter activations and the Adam update of all four parameter
// groups, consuming the [P,10] gradient buffer directly.
//
// Replaces, per iteration:
//   * autograd's backward through scene/gaussian_model.py:58-74 (get_opacity = sigmoid,
//     get_scaling = exp, get_rotation = F.normalize, get_xyz = identity),
//   * the scale regulariser's backward (slam/mapper.py:186-191),
//   * torch.optim.Adam(eps=1e-15).step() over the four groups of
//     scene/gaussian_model.py:97-121
// - some 30 launches - by ONE pass: 70 floats read and 30 written per surfel.
// The optimizer state stays where torch keeps it (exp_avg / exp_avg_sq tensors of the
// caller's torch.optim.Adam), so densification and pruning code keeps working.
#include <cmath>

#include "../../include/splat_b200.h"
#include "common.cuh"

namespace {

struct Group {
  float *p, *m, *v;
  float step_size;       // lr / (1 - beta1^t)
  float bc2_sqrt;        // sqrt(1 - beta2^t)       (denom = sqrt(v) / bc2_sqrt + eps)
};

struct AdamParams {
  int P;
  const float *grads;    // [P,10] w.r.t. the ACTIVATED values: xyz 3 | opacity 1 | scaling 2 | rotation 4
  Group g[4];            // xyz, opacity, scaling, rotation
  float omb1, beta2, omb2, eps;   // 1 - beta1, beta2, 1 - beta2 (formed in double, as torch does)
  float scaling_max, scaling_penalty;
  float *reg_out;       // optional: += penalty * sum(max(scale) - scaling_max) over the offenders
};

// torch.optim.Adam (single-tensor and foreach forms agree on this arithmetic):
//   m.lerp_(g, 1 - beta1);  v = beta2 v + (1 - beta2) g g;
//   p -= step_size * m / (sqrt(v) / sqrt(bc2) + eps)
__device__ __forceinline__ float adam(float g, float &m, float &v, float p, const Group &G,
                                      const AdamParams &a) {
  m = m + a.omb1 * (g - m);
  v = v * a.beta2 + a.omb2 * g * g;
  const float denom = sqrtf(v) / G.bc2_sqrt + a.eps;
  return p - G.step_size * (m / denom);
}

__global__ void __launch_bounds__(256) activation_adam_kernel(AdamParams a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  float reg = 0.0f;
  if (i < a.P) {
    const float2 *row = reinterpret_cast<const float2 *>(a.grads + 10 * (size_t)i);
    const float2 r0 = row[0], r1 = row[1], r2 = row[2], r3 = row[3], r4 = row[4];
    // ---- xyz: identity activation ----
    {
      const Group &G = a.g[0];
      const float gx[3] = {r0.x, r0.y, r1.x};
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const size_t j = 3 * (size_t)i + k;
        float m = G.m[j], v = G.v[j];
        G.p[j] = adam(gx[k], m, v, G.p[j], G, a);
        G.m[j] = m;
        G.v[j] = v;
      }
    }
    // ---- opacity: sigmoid ----
    {
      const Group &G = a.g[1];
      const float raw = G.p[i];
      const float s = 1.0f / (1.0f + expf(-raw));
      const float g = r1.y * ((1.0f - s) * s);          // torch: grad * (1 - y) * y
      float m = G.m[i], v = G.v[i];
      G.p[i] = adam(g, m, v, raw, G, a);
      G.m[i] = m;
      G.v[i] = v;
    }
    // ---- scaling: exp, plus the regulariser on max(scale) ----
    {
      const Group &G = a.g[2];
      float2 raw = reinterpret_cast<float2 *>(G.p)[i];
      const float s0 = expf(raw.x), s1 = expf(raw.y);
      float g0 = r2.x, g1 = r2.y;
      if (a.scaling_penalty != 0.0f) {
        const bool first = s0 >= s1;                     // torch.max(dim): first maximal index
        const float mx = first ? s0 : s1;
        if (mx >= a.scaling_max) {
          (first ? g0 : g1) += a.scaling_penalty;
          reg = a.scaling_penalty * (mx - a.scaling_max);
        }
      }
      g0 *= s0;                                           // torch: grad * exp(x)
      g1 *= s1;
      float2 m = reinterpret_cast<float2 *>(G.m)[i], v = reinterpret_cast<float2 *>(G.v)[i];
      raw.x = adam(g0, m.x, v.x, raw.x, G, a);
      raw.y = adam(g1, m.y, v.y, raw.y, G, a);
      reinterpret_cast<float2 *>(G.p)[i] = raw;
      reinterpret_cast<float2 *>(G.m)[i] = m;
      reinterpret_cast<float2 *>(G.v)[i] = v;
    }
    // ---- rotation: F.normalize (eps 1e-12) ----
    {
      const Group &G = a.g[3];
      float4 q = reinterpret_cast<float4 *>(G.p)[i];
      const float gq[4] = {r3.x, r3.y, r4.x, r4.y};
      const float norm = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
      const float d = fmaxf(norm, 1e-12f);
      float g[4];
      if (norm >= 1e-12f) {           // clamp_min passes the gradient where norm >= eps
        const float u[4] = {q.x / d, q.y / d, q.z / d, q.w / d};
        const float dot = u[0] * gq[0] + u[1] * gq[1] + u[2] * gq[2] + u[3] * gq[3];
#pragma unroll
        for (int k = 0; k < 4; k++) g[k] = (gq[k] - u[k] * dot) / d;
      } else {
#pragma unroll
        for (int k = 0; k < 4; k++) g[k] = gq[k] / d;     // the clamp is active: y = x / eps
      }
      float4 m = reinterpret_cast<float4 *>(G.m)[i], v = reinterpret_cast<float4 *>(G.v)[i];
      q.x = adam(g[0], m.x, v.x, q.x, G, a);
      q.y = adam(g[1], m.y, v.y, q.y, G, a);
      q.z = adam(g[2], m.z, v.z, q.z, G, a);
      q.w = adam(g[3], m.w, v.w, q.w, G, a);
      reinterpret_cast<float4 *>(G.p)[i] = q;
      reinterpret_cast<float4 *>(G.m)[i] = m;
      reinterpret_cast<float4 *>(G.v)[i] = v;
    }
  }
  if (a.reg_out) {   // block sum, one atomic per CTA
    __shared__ float s_red[8];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) reg += __shfl_xor_sync(0xffffffffu, reg, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = reg;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int w = 0; w < 8; w++) t += s_red[w];
      if (t != 0.0f) atomicAdd(a.reg_out, t);
    }
  }
}

}  // namespace

extern "C" int splat_activation_adam_step(const splat_adam_args *x) {
  if (!x) return splat_fail("splat_activation_adam_step: NULL argument");
  if (x->P < 0) return splat_fail("splat_activation_adam_step: bad P=%d", x->P);
  if (x->P == 0) return 0;
  if (!x->packed_grads) return splat_fail("splat_activation_adam_step: packed_grads is NULL");
  AdamParams a;
  a.P = x->P;
  a.grads = x->packed_grads;
  for (int k = 0; k < 4; k++) {
    if (!x->param[k] || !x->exp_avg[k] || !x->exp_avg_sq[k])
      return splat_fail("splat_activation_adam_step: NULL tensor in group %d", k);
    if (x->step[k] < 1) return splat_fail("splat_activation_adam_step: step[%d] must be >= 1", k);
    a.g[k].p = x->param[k];
    a.g[k].m = x->exp_avg[k];
    a.g[k].v = x->exp_avg_sq[k];
    // bias corrections in double, as torch does in Python floats
    const double bc1 = 1.0 - std::pow(x->beta1, (double)x->step[k]);
    const double bc2 = 1.0 - std::pow(x->beta2, (double)x->step[k]);
    a.g[k].step_size = (float)(x->lr[k] / bc1);
    a.g[k].bc2_sqrt = (float)std::sqrt(bc2);
  }
  if (((uintptr_t)x->packed_grads & 7) || ((uintptr_t)x->param[2] & 7) || ((uintptr_t)x->exp_avg[2] & 7) ||
      ((uintptr_t)x->exp_avg_sq[2] & 7) || ((uintptr_t)x->param[3] & 15) ||
      ((uintptr_t)x->exp_avg[3] & 15) || ((uintptr_t)x->exp_avg_sq[3] & 15))
    return splat_fail("splat_activation_adam_step: packed_grads / scaling tensors must be 8-byte, "
                      "rotation tensors 16-byte aligned");
  a.omb1 = (float)(1.0 - x->beta1);
  a.beta2 = (float)x->beta2;
  a.omb2 = (float)(1.0 - x->beta2);
  a.eps = (float)x->eps;
  a.scaling_max = (float)x->scaling_max;
  a.scaling_penalty = (float)x->scaling_max_penalty;
  a.reg_out = x->reg_out;
  activation_adam_kernel<<<(x->P + 255) / 256, 256, 0, LS((cudaStream_t)x->stream)>>>(a);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : splat_fail("splat_activation_adam_step: %s", cudaGetErrorString(e));
}
