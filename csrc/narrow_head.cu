// Dense layers with a handful of output units (regression heads: Dense(1), the Boston-housing and wide-MLP models;
// small classifiers up to 8 units). A 128-wide tensor-core tile is almost empty for them and the work is one pass over
// the activations, so these three kernels run at HBM speed instead:
//
//   head_fwd    z[b][o]      = bias[o] + sum_k A[b][k] W[o][k]                       (warp per row, W in shared memory)
//   head_wgrad  dW[o][k]    += sum_b dZ[b][o] A[b][k]                                (column sums weighted by dZ)
//   head_dgrad  dZp[b][k]    = (sum_o dZ[b][o] W[o][k]) * act'(Aprev[b][k])          (outer product + activation gradient)
//
// bf16 operands, fp32 accumulation, the same layouts as the tcgen05 GEMM (K1) they stand in for. Reference call sites:
// the last Dense layer of examples/ml_mlp_regression.py / tests (Keras fit / predict, elephas/worker.py:42,108,123).
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <mutex>

#include "../include/elephas_b200.h"
#include "host.h"
#include "math.cuh"
#include "ptx.cuh"

namespace eb {

constexpr int kHeadMaxOut = 8;

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}

// One warp per row of A. K % 8 == 0, rows 16-byte aligned.
__global__ void __launch_bounds__(256) head_fwd_kernel(const __nv_bfloat16* __restrict__ A, long lda,
                                                       const __nv_bfloat16* __restrict__ W, long ldw,
                                                       const float* __restrict__ bias, float* __restrict__ z, long ldz,
                                                       int rows, int K, int out) {
  extern __shared__ uint4 w_sm[];  // [out][K / 8]
  const int k8 = K / 8;
  for (int i = threadIdx.x; i < out * k8; i += blockDim.x) {
    const int o = i / k8, c = i - o * k8;
    w_sm[i] = *reinterpret_cast<const uint4*>(W + o * ldw + c * 8);
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int r = blockIdx.x * 8 + warp; r < rows; r += gridDim.x * 8) {
    float acc[kHeadMaxOut] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    const __nv_bfloat16* a = A + static_cast<long>(r) * lda;
#pragma unroll 4
    for (int c = lane; c < k8; c += 32) {
      float x[8];
      unpack8(ld_nc_u4(a + c * 8), x);
#pragma unroll
      for (int o = 0; o < kHeadMaxOut; ++o) {
        if (o < out) {
          float w[8];
          unpack8(w_sm[o * k8 + c], w);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[o] += x[j] * w[j];
        }
      }
    }
#pragma unroll
    for (int o = 0; o < kHeadMaxOut; ++o) {
      if (o < out) {
        float v = acc[o];
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
        if (lane == 0) z[static_cast<long>(r) * ldz + o] = v + (bias ? bias[o] : 0.f);
      }
    }
  }
}

// Block = 32 x 8 threads: 256 columns (8 per thread) x a slab of rows; partials meet in shared memory, one atomicAdd
// per (unit, column) and block. dw must be zeroed by the caller.
__global__ void __launch_bounds__(256) head_wgrad_kernel(const __nv_bfloat16* __restrict__ dz, long lddz,
                                                         const __nv_bfloat16* __restrict__ A, long lda,
                                                         float* __restrict__ dw, long lddw, int rows, int K, int out) {
  __shared__ float red[8][256];
  const int c = blockIdx.x * 256 + threadIdx.x * 8;
  const int rows_per_block = (rows + gridDim.y - 1) / gridDim.y;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(rows, r0 + rows_per_block);
  float acc[kHeadMaxOut][8];
#pragma unroll
  for (int o = 0; o < kHeadMaxOut; ++o)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[o][j] = 0.f;
  if (c < K) {
#pragma unroll 2
    for (int r = r0 + threadIdx.y; r < r1; r += 8) {
      float x[8];
      unpack8(ld_nc_u4(A + static_cast<long>(r) * lda + c), x);
#pragma unroll
      for (int o = 0; o < kHeadMaxOut; ++o) {
        if (o < out) {
          const float g = __bfloat162float(dz[static_cast<long>(r) * lddz + o]);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[o][j] += g * x[j];
        }
      }
    }
  }
  const int t = threadIdx.y * 32 + threadIdx.x;
  const int col = blockIdx.x * 256 + t;
  for (int o = 0; o < out; ++o) {
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 8; ++j) red[threadIdx.y][threadIdx.x * 8 + j] = acc[o][j];
    __syncthreads();
    if (col < K) {
      float s = red[0][t];
#pragma unroll
      for (int j = 1; j < 8; ++j) s += red[j][t];
      atomicAdd(dw + o * lddw + col, s);
    }
  }
}

// x indexes groups of 8 columns, y strides over rows (like the vectorised activation kernels).
__global__ void __launch_bounds__(128) head_dgrad_kernel(const __nv_bfloat16* __restrict__ dz, long lddz,
                                                         const __nv_bfloat16* __restrict__ W, long ldw,
                                                         const __nv_bfloat16* __restrict__ aux, long ldaux,
                                                         __nv_bfloat16* __restrict__ dzp, long lddzp, int rows, int K,
                                                         int out, int act) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (c >= K) return;
  float w[kHeadMaxOut][8];
#pragma unroll
  for (int o = 0; o < kHeadMaxOut; ++o) {
    if (o < out) unpack8(*reinterpret_cast<const uint4*>(W + o * ldw + c), w[o]);
  }
#pragma unroll 2
  for (int r = blockIdx.y; r < rows; r += gridDim.y) {
    float g[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int o = 0; o < kHeadMaxOut; ++o) {
      if (o < out) {
        const float d = __bfloat162float(dz[static_cast<long>(r) * lddz + o]);
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] += d * w[o][j];
      }
    }
    float y[8], da[8];
    unpack8(ld_nc_u4(aux + static_cast<long>(r) * ldaux + c), y);
    act_grad_chunk<8>(act, y, 1.f, da);
    uint4 o4;
    o4.x = pack_bf16x2(g[0] * da[0], g[1] * da[1]); o4.y = pack_bf16x2(g[2] * da[2], g[3] * da[3]);
    o4.z = pack_bf16x2(g[4] * da[4], g[5] * da[5]); o4.w = pack_bf16x2(g[6] * da[6], g[7] * da[7]);
    *reinterpret_cast<uint4*>(dzp + static_cast<long>(r) * lddzp + c) = o4;
  }
}

static int check_head(const EbHeadArgs& a, const char* what) {
  if (a.rows <= 0 || a.K <= 0 || a.out <= 0) return fail("%s: empty problem", what);
  if (a.out > kHeadMaxOut) return fail("%s: at most %d output units (got %d)", what, kHeadMaxOut, a.out);
  if (a.K % 8 || a.lda % 8 || a.ldw % 8) return fail("%s: K, lda and ldw must be multiples of 8", what);
  if ((reinterpret_cast<uintptr_t>(a.A) & 15) || (reinterpret_cast<uintptr_t>(a.W) & 15))
    return fail("%s: A and W must be 16-byte aligned", what);
  return 0;
}

}  // namespace eb

using namespace eb;

extern "C" int eb_head_fwd(const EbHeadArgs* a, void* stream) {
  if (a == nullptr) return fail("eb_head_fwd: null args");
  if (check_head(*a, "eb_head_fwd")) return 1;
  if (a->z == nullptr) return fail("eb_head_fwd: output missing");
  const size_t smem = static_cast<size_t>(a->out) * a->K * 2;
  if (smem > 96 * 1024) return fail("eb_head_fwd: weights (%zu bytes) do not fit in shared memory", smem);
  {
    // once per device; several worker threads may get here at the same time
    static std::mutex init_mutex;
    static bool attr_done[64] = {false};
    int dev = 0;
    EB_CUDA_OK(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> guard(init_mutex);
    if (!attr_done[dev & 63]) {
      EB_CUDA_OK(cudaFuncSetAttribute(head_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
      attr_done[dev & 63] = true;
    }
  }
  int blocks = (a->rows + 7) / 8;
  const int cap = num_sms() * 4;
  if (blocks > cap) blocks = cap;
  head_fwd_kernel<<<blocks, 256, smem, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(a->A), a->lda, static_cast<const __nv_bfloat16*>(a->W), a->ldw, a->bias, a->z,
      a->ldz, a->rows, a->K, a->out);
  EB_LAUNCH_OK("head_fwd_kernel");
  return 0;
}

extern "C" int eb_head_wgrad(const EbHeadArgs* a, void* stream) {
  if (a == nullptr) return fail("eb_head_wgrad: null args");
  if (check_head(*a, "eb_head_wgrad")) return 1;
  if (a->dz == nullptr || a->dw == nullptr) return fail("eb_head_wgrad: dz and dw are required");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  EB_CUDA_OK(cudaMemsetAsync(a->dw, 0, sizeof(float) * (static_cast<size_t>(a->out - 1) * a->lddw + a->K), s));
  const int cb = (a->K + 255) / 256;
  int rb = (num_sms() * 4 + cb - 1) / cb;
  if (rb > (a->rows + 31) / 32) rb = (a->rows + 31) / 32;
  if (rb < 1) rb = 1;
  head_wgrad_kernel<<<dim3(cb, rb), dim3(32, 8), 0, s>>>(static_cast<const __nv_bfloat16*>(a->dz), a->lddz,
                                                         static_cast<const __nv_bfloat16*>(a->A), a->lda, a->dw, a->lddw,
                                                         a->rows, a->K, a->out);
  EB_LAUNCH_OK("head_wgrad_kernel");
  return 0;
}

extern "C" int eb_head_dgrad(const EbHeadArgs* a, void* stream) {
  if (a == nullptr) return fail("eb_head_dgrad: null args");
  if (check_head(*a, "eb_head_dgrad")) return 1;
  if (a->dz == nullptr || a->dzp == nullptr || a->aux == nullptr) return fail("eb_head_dgrad: dz, aux and dzp are required");
  if (a->ldaux % 8 || a->lddzp % 8 || (reinterpret_cast<uintptr_t>(a->aux) & 15) || (reinterpret_cast<uintptr_t>(a->dzp) & 15))
    return fail("eb_head_dgrad: aux / dzp must be 16-byte aligned with leading dimensions that are multiples of 8");
  const int gx = (a->K / 8 + 127) / 128;
  long gy = (static_cast<long>(num_sms()) * 8 + gx - 1) / gx;
  if (gy > a->rows) gy = a->rows;
  if (gy > 65535) gy = 65535;
  head_dgrad_kernel<<<dim3(gx, static_cast<unsigned>(gy)), 128, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(a->dz), a->lddz, static_cast<const __nv_bfloat16*>(a->W), a->ldw,
      static_cast<const __nv_bfloat16*>(a->aux), a->ldaux, static_cast<__nv_bfloat16*>(a->dzp), a->lddzp, a->rows, a->K,
      a->out, a->act);
  EB_LAUNCH_OK("head_dgrad_kernel");
  return 0;
}
