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

__global__ void __launch_bounds__(256) optimizer_kernel(const OptimDev p) {
  __shared__ float sh_lr[2];
  if (threadIdx.x == 0) {
    const float it = static_cast<float>(*p.step);
    const float lr_t = p.decay != 0.f ? p.lr / (1.f + p.decay * it) : p.lr;
    float lr_adam = lr_t;
    if (p.kind == EB_OPT_ADAM) lr_adam = lr_t * sqrtf(1.f - powf(p.beta2, it + 1.f)) / (1.f - powf(p.beta1, it + 1.f));
    if (p.kind == EB_OPT_ADAMAX) lr_adam = lr_t / (1.f - powf(p.beta1, it + 1.f));
    sh_lr[0] = lr_t;
    sh_lr[1] = lr_adam;
  }
  __syncthreads();
  const float lr_t = sh_lr[0], lr_adam = sh_lr[1];
  // four consecutive elements per thread per iteration: 16-byte accesses on every stream
  const long n4 = p.n / 4;
  for (long q = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; q < n4;
       q += static_cast<long>(gridDim.x) * blockDim.x) {
    const float4 g = reinterpret_cast<const float4*>(p.g32)[q];
    float st[4];
    st[0] = step_value(p, 4 * q + 0, g.x, lr_t, lr_adam);
    st[1] = step_value(p, 4 * q + 1, g.y, lr_t, lr_adam);
    st[2] = step_value(p, 4 * q + 2, g.z, lr_t, lr_adam);
    st[3] = step_value(p, 4 * q + 3, g.w, lr_t, lr_adam);
    if (p.apply) {
      float4 w = reinterpret_cast<float4*>(p.w32)[q];
      w.x -= st[0]; w.y -= st[1]; w.z -= st[2]; w.w -= st[3];
      reinterpret_cast<float4*>(p.w32)[q] = w;
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
    if (p.delta32) reinterpret_cast<float4*>(p.delta32)[q] = make_float4(st[0], st[1], st[2], st[3]);
  }
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
      (a->delta_f32 && (reinterpret_cast<uintptr_t>(a->delta_f32) & 15)))
    return fail("eb_optimizer_step: buffers must be 16-byte aligned (bf16 buffers 8-byte)");
  long blocks = (a->n / 4 + 255) / 256 + 1;
  const long cap = static_cast<long>(num_sms()) * 8;
  if (blocks > cap) blocks = cap;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  optimizer_kernel<<<static_cast<int>(blocks), 256, 0, s>>>(p);
  EB_LAUNCH_OK("optimizer_kernel");
  if (a->advance) {
    advance_step_kernel<<<1, 1, 0, s>>>(reinterpret_cast<long long*>(a->step));
    EB_LAUNCH_OK("advance_step_kernel");
  }
  return 0;
}
