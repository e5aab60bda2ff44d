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
#include "optim.cuh"
#include "ptx.cuh"

namespace eb {

template <int KIND>
__global__ void __launch_bounds__(256) optimizer_kernel(const OptimDev p) {
  __shared__ float sh_lr[2];
  if (threadIdx.x == 0) step_rates<KIND>(p, sh_lr[0], sh_lr[1]);
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

namespace eb {
int fill_optim(const EbOptimArgs* a, OptimDev& p, const char* who) {
  if (a == nullptr) return fail("%s: null optimizer args", who);
  if (a->step == nullptr) return fail("%s: the device step counter is required", who);
  const int need[7][3] = {{0, 0, 0}, {1, 0, 0}, {1, 1, 1}, {1, 1, 1}, {1, 0, 0}, {1, 1, 0}, {1, 1, 0}};
  if (a->kind < EB_OPT_SGD || a->kind > EB_OPT_ADAMAX) return fail("%s: unknown optimizer kind %d", who, a->kind);
  const bool sgd_plain = a->kind == EB_OPT_SGD && a->momentum == 0.f;
  if (!sgd_plain && need[a->kind][0] && a->slot0 == nullptr) return fail("%s: slot0 missing", who);
  if (a->kind == EB_OPT_RMSPROP && a->momentum > 0.f && a->slot1 == nullptr) return fail("%s: slot1 missing", who);
  if (a->kind == EB_OPT_RMSPROP && a->centered && a->slot2 == nullptr) return fail("%s: slot2 missing", who);
  if ((a->kind == EB_OPT_ADAM || a->kind == EB_OPT_ADADELTA || a->kind == EB_OPT_ADAMAX) && a->slot1 == nullptr)
    return fail("%s: slot1 missing", who);
  if (a->kind == EB_OPT_ADAM && a->amsgrad && a->slot2 == nullptr) return fail("%s: slot2 missing", who);
  if (a->apply && a->w32 == nullptr) return fail("%s: apply needs the fp32 weights", who);
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
    return fail("%s: buffers must be 16-byte aligned (bf16 buffers 8-byte)", who);
  return 0;
}

// advances the device-resident iteration counter (stream ordered: the step that follows sees it)
void launch_advance_step(void* step, cudaStream_t s) { advance_step_kernel<<<1, 1, 0, s>>>(reinterpret_cast<long long*>(step)); }
}  // namespace eb

extern "C" int eb_optimizer_step(const EbOptimArgs* a, void* stream) {
  using namespace eb;
  if (a == nullptr) return fail("eb_optimizer_step: null args");
  if (a->n <= 0) return 0;
  OptimDev p;
  if (fill_optim(a, p, "eb_optimizer_step")) return 1;
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
