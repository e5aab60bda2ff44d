// K10 device code shared by the optimizer kernel (optim_kernels.cu) and the parameter-server kernel that fuses the
// optimizer step with the push into peer memory (ps_kernels.cu): update rules of Keras 2.10 per optimizer kind.
// Formulas match elephas_b200/models/optimizers.py (the CPU path and numerical oracle).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../include/elephas_b200.h"
#include "ptx.cuh"

namespace eb {

struct OptimDev {
  int kind;
  long n;
  float* w32;
  const float* g32;
  float *s0, *s1, *s2;
  __nv_bfloat16* w16;
  __nv_bfloat16* delta16;
  float* delta32;
  int apply;
  float lr, decay, momentum, rho, eps, beta1, beta2;
  int nesterov, centered, amsgrad;
  const long long* step;
};

__device__ __forceinline__ float step_value(const OptimDev& p, long i, float g, float lr_t, float lr_adam) {
  switch (p.kind) {
    case EB_OPT_SGD: {
      if (p.momentum == 0.f) return lr_t * g;
      float v = p.momentum * p.s0[i] - lr_t * g;
      p.s0[i] = v;
      return p.nesterov ? -(p.momentum * v - lr_t * g) : -v;
    }
    case EB_OPT_RMSPROP: {
      float rms = p.rho * p.s0[i] + (1.f - p.rho) * g * g;
      p.s0[i] = rms;
      float denom = rms;
      if (p.centered) {
        float mg = p.rho * p.s2[i] + (1.f - p.rho) * g;
        p.s2[i] = mg;
        denom = rms - mg * mg;
      }
      float st = g / (sqrtf(denom) + p.eps);
      if (p.momentum > 0.f) {
        float m = p.momentum * p.s1[i] + lr_t * st;
        p.s1[i] = m;
        return m;
      }
      return lr_t * st;
    }
    case EB_OPT_ADAM: {
      float m = p.beta1 * p.s0[i] + (1.f - p.beta1) * g;
      float v = p.beta2 * p.s1[i] + (1.f - p.beta2) * g * g;
      p.s0[i] = m;
      p.s1[i] = v;
      if (p.amsgrad) {
        float vh = fmaxf(p.s2[i], v);
        p.s2[i] = vh;
        v = vh;
      }
      return lr_adam * m / (sqrtf(v) + p.eps);
    }
    case EB_OPT_ADAGRAD: {
      float acc = p.s0[i] + g * g;
      p.s0[i] = acc;
      return lr_t * g / (sqrtf(acc) + p.eps);
    }
    case EB_OPT_ADADELTA: {
      float ag = p.rho * p.s0[i] + (1.f - p.rho) * g * g;
      float upd = g * sqrtf(p.s1[i] + p.eps) / sqrtf(ag + p.eps);
      p.s0[i] = ag;
      p.s1[i] = p.rho * p.s1[i] + (1.f - p.rho) * upd * upd;
      return lr_t * upd;
    }
    case EB_OPT_ADAMAX: {
      float m = p.beta1 * p.s0[i] + (1.f - p.beta1) * g;
      float u = fmaxf(p.beta2 * p.s1[i], fabsf(g));
      p.s0[i] = m;
      p.s1[i] = u;
      return lr_adam * m / (u + p.eps);
    }
    default: return 0.f;
  }
}

__device__ __forceinline__ float4 ld4(const float* p, long q) { return reinterpret_cast<const float4*>(p)[q]; }
__device__ __forceinline__ void st4(float* p, long q, const float4& v) { reinterpret_cast<float4*>(p)[q] = v; }

// The step of four consecutive elements (quad q) for one optimizer kind: slots travel as 16-byte vectors and the
// kind is a template parameter, so the inner loop has no switch and every stream is a full-width access.
template <int KIND>
__device__ __forceinline__ void step_quad(const OptimDev& p, long q, const float4& g4, float lr_t, float lr_adam,
                                          float (&st)[4]) {
  const float g[4] = {g4.x, g4.y, g4.z, g4.w};
  if (KIND == EB_OPT_SGD) {
    if (p.momentum == 0.f) {
#pragma unroll
      for (int j = 0; j < 4; ++j) st[j] = lr_t * g[j];
      return;
    }
    const float4 v4 = ld4(p.s0, q);
    float v[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      v[j] = p.momentum * v[j] - lr_t * g[j];
      st[j] = p.nesterov ? -(p.momentum * v[j] - lr_t * g[j]) : -v[j];
    }
    st4(p.s0, q, make_float4(v[0], v[1], v[2], v[3]));
  } else if (KIND == EB_OPT_RMSPROP) {
    const float4 r4 = ld4(p.s0, q);
    float rms[4] = {r4.x, r4.y, r4.z, r4.w}, denom[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) denom[j] = rms[j] = p.rho * rms[j] + (1.f - p.rho) * g[j] * g[j];
    st4(p.s0, q, make_float4(rms[0], rms[1], rms[2], rms[3]));
    if (p.centered) {
      const float4 m4 = ld4(p.s2, q);
      float mg[4] = {m4.x, m4.y, m4.z, m4.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        mg[j] = p.rho * mg[j] + (1.f - p.rho) * g[j];
        denom[j] = rms[j] - mg[j] * mg[j];
      }
      st4(p.s2, q, make_float4(mg[0], mg[1], mg[2], mg[3]));
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) st[j] = lr_t * (g[j] / (sqrtf(denom[j]) + p.eps));
    if (p.momentum > 0.f) {
      const float4 m4 = ld4(p.s1, q);
      float m[4] = {m4.x, m4.y, m4.z, m4.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) st[j] = m[j] = p.momentum * m[j] + st[j];
      st4(p.s1, q, make_float4(m[0], m[1], m[2], m[3]));
    }
  } else if (KIND == EB_OPT_ADAM) {
    const float4 m4 = ld4(p.s0, q), v4 = ld4(p.s1, q);
    float m[4] = {m4.x, m4.y, m4.z, m4.w}, v[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      m[j] = p.beta1 * m[j] + (1.f - p.beta1) * g[j];
      v[j] = p.beta2 * v[j] + (1.f - p.beta2) * g[j] * g[j];
    }
    st4(p.s0, q, make_float4(m[0], m[1], m[2], m[3]));
    st4(p.s1, q, make_float4(v[0], v[1], v[2], v[3]));
    if (p.amsgrad) {
      const float4 h4 = ld4(p.s2, q);
      const float h[4] = {h4.x, h4.y, h4.z, h4.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = fmaxf(h[j], v[j]);
      st4(p.s2, q, make_float4(v[0], v[1], v[2], v[3]));
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) st[j] = lr_adam * m[j] / (sqrtf(v[j]) + p.eps);
  } else if (KIND == EB_OPT_ADAGRAD) {
    const float4 a4 = ld4(p.s0, q);
    float acc[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      acc[j] += g[j] * g[j];
      st[j] = lr_t * g[j] / (sqrtf(acc[j]) + p.eps);
    }
    st4(p.s0, q, make_float4(acc[0], acc[1], acc[2], acc[3]));
  } else if (KIND == EB_OPT_ADADELTA) {
    const float4 a4 = ld4(p.s0, q), d4 = ld4(p.s1, q);
    float ag[4] = {a4.x, a4.y, a4.z, a4.w}, ad[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      ag[j] = p.rho * ag[j] + (1.f - p.rho) * g[j] * g[j];
      const float upd = g[j] * sqrtf(ad[j] + p.eps) / sqrtf(ag[j] + p.eps);
      ad[j] = p.rho * ad[j] + (1.f - p.rho) * upd * upd;
      st[j] = lr_t * upd;
    }
    st4(p.s0, q, make_float4(ag[0], ag[1], ag[2], ag[3]));
    st4(p.s1, q, make_float4(ad[0], ad[1], ad[2], ad[3]));
  } else {  // EB_OPT_ADAMAX
    const float4 m4 = ld4(p.s0, q), u4 = ld4(p.s1, q);
    float m[4] = {m4.x, m4.y, m4.z, m4.w}, u[4] = {u4.x, u4.y, u4.z, u4.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      m[j] = p.beta1 * m[j] + (1.f - p.beta1) * g[j];
      u[j] = fmaxf(p.beta2 * u[j], fabsf(g[j]));
      st[j] = lr_adam * m[j] / (u[j] + p.eps);
    }
    st4(p.s0, q, make_float4(m[0], m[1], m[2], m[3]));
    st4(p.s1, q, make_float4(u[0], u[1], u[2], u[3]));
  }
}

// Learning rates of this step from the device-resident iteration counter: lr_t (time decay) and the bias-corrected
// rate of Adam / Adamax.
template <int KIND>
__device__ __forceinline__ void step_rates(const OptimDev& p, float& lr_t, float& lr_adam) {
  const float it = static_cast<float>(*p.step);
  lr_t = p.decay != 0.f ? p.lr / (1.f + p.decay * it) : p.lr;
  lr_adam = lr_t;
  if (KIND == EB_OPT_ADAM) lr_adam = lr_t * sqrtf(1.f - powf(p.beta2, it + 1.f)) / (1.f - powf(p.beta1, it + 1.f));
  if (KIND == EB_OPT_ADAMAX) lr_adam = lr_t / (1.f - powf(p.beta1, it + 1.f));
}

// Argument checks + EbOptimArgs -> OptimDev (host side, shared by both entry points). Returns non-zero on error.
int fill_optim(const EbOptimArgs* a, OptimDev& p, const char* who);
void launch_advance_step(void* step, cudaStream_t s);

}  // namespace eb
