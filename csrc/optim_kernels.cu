// K10: fused optimizer step over the flat parameter buffer (Keras 2.10 update rules).
//
// One kernel reads the fp32 gradient, updates the optimizer slots and either
//   * applies the step to the fp32 master weights and refreshes the bf16 compute copy, or
//   * ("delta mode") leaves the weights alone and writes the step itself as a bf16 / fp32 delta
//     into the communication buffer that the fused all-reduce (K6) or the PS push (K7) consumes
//     - the synchronous path of the reference averages exactly these deltas
//       (elephas/spark_model.py:221-227, elephas/worker.py:44-45).
// The iteration counter lives in device memory and is advanced by a one-thread kernel, so a step
// captured in a CUDA graph replays with correct learning-rate decay and Adam bias correction
// without any host-side parameter update.
//
// Formulas match elephas_b200/models/optimizers.py (the CPU path and numerical oracle).
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../include/elephas_b200.h"
#include "host.h"
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

template <int KIND>
__global__ void __launch_bounds__(256) optimizer_kernel(const OptimDev p) {
  __shared__ float sh_lr[2];
  if (threadIdx.x == 0) {
    const float it = static_cast<float>(*p.step);
    const float lr_t = p.decay != 0.f ? p.lr / (1.f + p.decay * it) : p.lr;
    float lr_adam = lr_t;
    if (KIND == EB_OPT_ADAM) lr_adam = lr_t * sqrtf(1.f - powf(p.beta2, it + 1.f)) / (1.f - powf(p.beta1, it + 1.f));
    if (KIND == EB_OPT_ADAMAX) lr_adam = lr_t / (1.f - powf(p.beta1, it + 1.f));
    sh_lr[0] = lr_t;
    sh_lr[1] = lr_adam;
  }
  __syncthreads();
  const float lr_t = sh_lr[0], lr_adam = sh_lr[1];
  const long n4 = p.n / 4;
#pragma unroll 2
  for (long q = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; q < n4;
       q += static_cast<long>(gridDim.x) * blockDim.x) {
    const float4 g = ld4(p.g32, q);
    float st[4];
    step_quad<KIND>(p, q, g, lr_t, lr_adam, st);
    if (p.apply) {
      float4 w = ld4(p.w32, q);
      w.x -= st[0]; w.y -= st[1]; w.z -= st[2]; w.w -= st[3];
      st4(p.w32, q, w);
      if (p.w16) {
        uint2 o;
        o.x = pack_bf16x2(w.x, w.y);
        o.y = pack_bf16x2(w.z, w.w);
        reinterpret_cast<uint2*>(p.w16)[q] = o;
      }
    }
    if (p.delta16) {
      uint2 o;
      o.x = pack_bf16x2(st[0], st[1]);
      o.y = pack_bf16x2(st[2], st[3]);
      reinterpret_cast<uint2*>(p.delta16)[q] = o;
    }
    if (p.delta32) st4(p.delta32, q, make_float4(st[0], st[1], st[2], st[3]));
  }
  // tail (n % 4 elements): the scalar form
  for (long i = n4 * 4 + blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < p.n;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const float st = step_value(p, i, p.g32[i], lr_t, lr_adam);
    if (p.apply) {
      const float w = p.w32[i] - st;
      p.w32[i] = w;
      if (p.w16) p.w16[i] = __float2bfloat16_rn(w);
    }
    if (p.delta16) p.delta16[i] = __float2bfloat16_rn(st);
    if (p.delta32) p.delta32[i] = st;
  }
}

__global__ void advance_step_kernel(long long* step) { *step += 1; }

}  // namespace eb

extern "C" int eb_optimizer_step(const EbOptimArgs* a, void* stream) {
  using namespace eb;
  if (a == nullptr) return fail("eb_optimizer_step: null args");
  if (a->n <= 0) return 0;
  if (a->step == nullptr) return fail("eb_optimizer_step: the device step counter is required");
  const int need[7][3] = {{0, 0, 0}, {1, 0, 0}, {1, 1, 1}, {1, 1, 1}, {1, 0, 0}, {1, 1, 0}, {1, 1, 0}};
  if (a->kind < EB_OPT_SGD || a->kind > EB_OPT_ADAMAX) return fail("eb_optimizer_step: unknown optimizer kind %d", a->kind);
  const bool sgd_plain = a->kind == EB_OPT_SGD && a->momentum == 0.f;
  if (!sgd_plain && need[a->kind][0] && a->slot0 == nullptr) return fail("eb_optimizer_step: slot0 missing");
  if (a->kind == EB_OPT_RMSPROP && a->momentum > 0.f && a->slot1 == nullptr) return fail("eb_optimizer_step: slot1 missing");
  if (a->kind == EB_OPT_RMSPROP && a->centered && a->slot2 == nullptr) return fail("eb_optimizer_step: slot2 missing");
  if ((a->kind == EB_OPT_ADAM || a->kind == EB_OPT_ADADELTA || a->kind == EB_OPT_ADAMAX) && a->slot1 == nullptr)
    return fail("eb_optimizer_step: slot1 missing");
  if (a->kind == EB_OPT_ADAM && a->amsgrad && a->slot2 == nullptr) return fail("eb_optimizer_step: slot2 missing");
  if (a->apply && a->w32 == nullptr) return fail("eb_optimizer_step: apply needs the fp32 weights");
  OptimDev p;
  p.kind = a->kind; p.n = a->n; p.w32 = a->w32; p.g32 = a->g32;
  p.s0 = a->slot0; p.s1 = a->slot1; p.s2 = a->slot2;
  p.w16 = static_cast<__nv_bfloat16*>(a->w16);
  p.delta16 = static_cast<__nv_bfloat16*>(a->delta_bf16); p.delta32 = a->delta_f32;
  p.apply = a->apply;
  p.lr = a->lr; p.decay = a->decay; p.momentum = a->momentum; p.rho = a->rho; p.eps = a->eps;
  p.beta1 = a->beta1; p.beta2 = a->beta2;
  p.nesterov = a->nesterov; p.centered = a->centered; p.amsgrad = a->amsgrad;
  p.step = reinterpret_cast<const long long*>(a->step);
  if ((reinterpret_cast<uintptr_t>(a->g32) & 15) || (a->w32 && (reinterpret_cast<uintptr_t>(a->w32) & 15)) ||
      (a->w16 && (reinterpret_cast<uintptr_t>(a->w16) & 7)) || (a->delta_bf16 && (reinterpret_cast<uintptr_t>(a->delta_bf16) & 7)) ||
      (a->delta_f32 && (reinterpret_cast<uintptr_t>(a->delta_f32) & 15)) ||
      (reinterpret_cast<uintptr_t>(a->slot0) & 15) || (reinterpret_cast<uintptr_t>(a->slot1) & 15) ||
      (reinterpret_cast<uintptr_t>(a->slot2) & 15))
    return fail("eb_optimizer_step: buffers must be 16-byte aligned (bf16 buffers 8-byte)");
  long blocks = (a->n / 4 + 255) / 256 + 1;
  const long cap = static_cast<long>(num_sms()) * 8;
  if (blocks > cap) blocks = cap;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int grid = static_cast<int>(blocks);
  switch (a->kind) {
    case EB_OPT_SGD: optimizer_kernel<EB_OPT_SGD><<<grid, 256, 0, s>>>(p); break;
    case EB_OPT_RMSPROP: optimizer_kernel<EB_OPT_RMSPROP><<<grid, 256, 0, s>>>(p); break;
    case EB_OPT_ADAM: optimizer_kernel<EB_OPT_ADAM><<<grid, 256, 0, s>>>(p); break;
    case EB_OPT_ADAGRAD: optimizer_kernel<EB_OPT_ADAGRAD><<<grid, 256, 0, s>>>(p); break;
    case EB_OPT_ADADELTA: optimizer_kernel<EB_OPT_ADADELTA><<<grid, 256, 0, s>>>(p); break;
    default: optimizer_kernel<EB_OPT_ADAMAX><<<grid, 256, 0, s>>>(p); break;
  }
  EB_LAUNCH_OK("optimizer_kernel");
  if (a->advance) {
    advance_step_kernel<<<1, 1, 0, s>>>(reinterpret_cast<long long*>(a->step));
    EB_LAUNCH_OK("advance_step_kernel");
  }
  return 0;
}
