// Activation functions (Keras names and semantics), their derivatives expressed in
// terms of the *output* value, and a counter-based Philox4x32-10 generator used for
// dropout masks (the backward pass regenerates the mask from (seed, element index),
// nothing is stored). Shared by the GEMM epilogues and the elementwise kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace eb {

// Activation ids: keep in sync with elephas_b200/ops/constants.py and include/elephas_b200.h.
enum Act : int {
  ACT_LINEAR = 0,
  ACT_RELU = 1,
  ACT_SIGMOID = 2,
  ACT_TANH = 3,
  ACT_SOFTPLUS = 4,
  ACT_SOFTSIGN = 5,
  ACT_ELU = 6,
  ACT_SELU = 7,
  ACT_EXPONENTIAL = 8,
  ACT_HARD_SIGMOID = 9,
};

__host__ __device__ __forceinline__ float act_apply(int act, float x) {
  switch (act) {
    case ACT_RELU: return x > 0.f ? x : 0.f;
    case ACT_SIGMOID: return 1.f / (1.f + expf(-x));
    case ACT_TANH: return tanhf(x);
    case ACT_SOFTPLUS: return x > 20.f ? x : log1pf(expf(x));
    case ACT_SOFTSIGN: return x / (1.f + fabsf(x));
    case ACT_ELU: return x > 0.f ? x : expm1f(x);
    case ACT_SELU: {
      const float a = 1.6732632423543772f, l = 1.0507009873554805f;
      return x > 0.f ? l * x : l * a * expm1f(x);
    }
    case ACT_EXPONENTIAL: return expf(x);
    case ACT_HARD_SIGMOID: {
      float y = 0.2f * x + 0.5f;
      return y < 0.f ? 0.f : (y > 1.f ? 1.f : y);
    }
    default: return x;
  }
}

// d act(z) / dz written as a function of y = act(z).
__host__ __device__ __forceinline__ float act_grad_from_out(int act, float y) {
  switch (act) {
    case ACT_RELU: return y > 0.f ? 1.f : 0.f;
    case ACT_SIGMOID: return y * (1.f - y);
    case ACT_TANH: return 1.f - y * y;
    case ACT_SOFTPLUS: return 1.f - expf(-y);
    case ACT_SOFTSIGN: {
      float t = 1.f - fabsf(y);
      return t * t;
    }
    case ACT_ELU: return y > 0.f ? 1.f : y + 1.f;
    case ACT_SELU: {
      const float a = 1.6732632423543772f, l = 1.0507009873554805f;
      return y > 0.f ? l : y + l * a;
    }
    case ACT_EXPONENTIAL: return y;
    case ACT_HARD_SIGMOID: return (y > 0.f && y < 1.f) ? 0.2f : 0.f;
    default: return 1.f;
  }
}


// Whole-chunk forms with the activation switch OUTSIDE the element loop (one uniform branch per chunk instead
// of a divergence-managed switch per element).
template <int N>
__device__ __forceinline__ void act_apply_chunk(int act, float (&x)[N]) {
#define EB_ACT_CASE(ID)                                   \
  case ID:                                                \
    _Pragma("unroll") for (int j = 0; j < N; ++j) x[j] = act_apply(ID, x[j]); \
    break;
  switch (act) {
    EB_ACT_CASE(ACT_RELU) EB_ACT_CASE(ACT_SIGMOID) EB_ACT_CASE(ACT_TANH) EB_ACT_CASE(ACT_SOFTPLUS)
    EB_ACT_CASE(ACT_SOFTSIGN) EB_ACT_CASE(ACT_ELU) EB_ACT_CASE(ACT_SELU) EB_ACT_CASE(ACT_EXPONENTIAL)
    EB_ACT_CASE(ACT_HARD_SIGMOID)
    default: break;
  }
#undef EB_ACT_CASE
}

// g[j] = act'(y[j] * y_scale) for a whole chunk.
template <int N>
__device__ __forceinline__ void act_grad_chunk(int act, const float (&y)[N], float y_scale, float (&g)[N]) {
#define EB_ACT_CASE(ID)                                   \
  case ID:                                                \
    _Pragma("unroll") for (int j = 0; j < N; ++j) g[j] = act_grad_from_out(ID, y[j] * y_scale); \
    break;
  switch (act) {
    EB_ACT_CASE(ACT_RELU) EB_ACT_CASE(ACT_SIGMOID) EB_ACT_CASE(ACT_TANH) EB_ACT_CASE(ACT_SOFTPLUS)
    EB_ACT_CASE(ACT_SOFTSIGN) EB_ACT_CASE(ACT_ELU) EB_ACT_CASE(ACT_SELU) EB_ACT_CASE(ACT_EXPONENTIAL)
    EB_ACT_CASE(ACT_HARD_SIGMOID)
    default:
      _Pragma("unroll") for (int j = 0; j < N; ++j) g[j] = 1.f;
      break;
  }
#undef EB_ACT_CASE
}

// ------------------------------------------------------------------ Philox
struct Philox4 {
  uint32_t x, y, z, w;
};

__host__ __device__ __forceinline__ uint32_t mulhi32(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
  return __umulhi(a, b);
#else
  return static_cast<uint32_t>((static_cast<uint64_t>(a) * b) >> 32);
#endif
}

// Philox4x32-10. counter = (c0, c1, 0, 0), key = (k0, k1).
__host__ __device__ __forceinline__ Philox4 philox4x32(uint32_t c0, uint32_t c1, uint32_t k0, uint32_t k1) {
  uint32_t c2 = 0, c3 = 0;
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    uint32_t hi0 = mulhi32(M0, c0), lo0 = M0 * c0;
    uint32_t hi1 = mulhi32(M1, c2), lo1 = M1 * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0;
    uint32_t n1 = lo1;
    uint32_t n2 = hi0 ^ c3 ^ k1;
    uint32_t n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return Philox4{c0, c1, c2, c3};
}

// Keep-mask for 4 consecutive elements whose first flat index is idx (idx % 4 == 0).
// Element e is kept iff its 32-bit draw >= p * 2^32. `stream` separates layers/steps.
__host__ __device__ __forceinline__ void dropout_keep4(uint64_t seed, uint64_t stream, uint64_t idx, uint32_t thresh,
                                                       bool (&keep)[4]) {
  uint64_t ctr = (idx >> 2) + stream * 0x9E3779B97F4A7C15ull;
  Philox4 r = philox4x32(static_cast<uint32_t>(ctr), static_cast<uint32_t>(ctr >> 32), static_cast<uint32_t>(seed),
                         static_cast<uint32_t>(seed >> 32));
  keep[0] = r.x >= thresh;
  keep[1] = r.y >= thresh;
  keep[2] = r.z >= thresh;
  keep[3] = r.w >= thresh;
}

__host__ __device__ __forceinline__ uint32_t dropout_threshold(float p) {
  double t = static_cast<double>(p) * 4294967296.0;
  if (t <= 0.0) return 0u;
  if (t >= 4294967295.0) return 0xFFFFFFFFu;
  return static_cast<uint32_t>(t);
}

}  // namespace eb
