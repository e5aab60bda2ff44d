// Memory-bound helper kernels of the native engine: dtype staging, bias-gradient column sums,
// standalone activation / dropout (for layers that are not fused into a GEMM epilogue) and the
// delta kernel (K5). All are vectorised (16-byte accesses), grid-stride, sized to the SM count.
//
// Reference sites absorbed: Keras Activation / Dropout layers (tests/conftest.py:12-18),
// subtract_params (elephas/utils/functional_utils.py:16-23 via elephas/worker.py:44-45).
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../include/elephas_b200.h"
#include "host.h"
#include "math.cuh"
#include "ptx.cuh"

namespace eb {

static inline int grid_for(long n_items, int threads, int per_thread) {
  long blocks = (n_items + static_cast<long>(threads) * per_thread - 1) / (static_cast<long>(threads) * per_thread);
  long cap = static_cast<long>(num_sms()) * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

// dst[r][0..cols) = bf16(src[r][0..cols)), dst[r][cols..ld_dst) = 0. Stages fp32 inputs into the padded
// bf16 layout the GEMM's TMA descriptors need.
__global__ void stage_f32_to_bf16_kernel(const float* __restrict__ src, long ld_src, __nv_bfloat16* __restrict__ dst,
                                         long ld_dst, long rows, int cols) {
  const long total = rows * ld_dst;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const long r = i / ld_dst;
    const int c = static_cast<int>(i - r * ld_dst);
    dst[i] = __float2bfloat16_rn(c < cols ? src[r * ld_src + c] : 0.f);
  }
}

// Flat fp32 -> bf16 (same indexing): refreshes the bf16 compute copy of the weights.
__global__ void cast_f32_to_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long n) {
  const long n8 = n / 8;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n8;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const float4 a = reinterpret_cast<const float4*>(src)[2 * i];
    const float4 b = reinterpret_cast<const float4*>(src)[2 * i + 1];
    uint4 o;
    o.x = pack_bf16x2(a.x, a.y); o.y = pack_bf16x2(a.z, a.w);
    o.z = pack_bf16x2(b.x, b.y); o.w = pack_bf16x2(b.z, b.w);
    reinterpret_cast<uint4*>(dst)[i] = o;
  }
  for (long i = n8 * 8 + blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long>(gridDim.x) * blockDim.x)
    dst[i] = __float2bfloat16_rn(src[i]);
}

// out[n] (+)= sum_m X[m][n] for a bf16 matrix: the bias gradient db = colsum(dZ).
// Block = 32 x 8 threads: each block owns 64 columns (2 per thread, bf16x2 loads) and a slab of rows;
// partial sums are combined through shared memory and one atomicAdd per column per block.
__global__ void colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x, long ld, int rows, int cols,
                                   float* __restrict__ out) {
  __shared__ float2 red[8][32];
  const int c = blockIdx.x * 64 + threadIdx.x * 2;
  const int rows_per_block = (rows + gridDim.y - 1) / gridDim.y;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(rows, r0 + rows_per_block);
  float2 acc = make_float2(0.f, 0.f);
  if (c < cols) {
    for (int r = r0 + threadIdx.y; r < r1; r += 8) {
      const uint32_t v = *reinterpret_cast<const uint32_t*>(x + static_cast<long>(r) * ld + c);
      acc.x += bf16_lo(v);
      acc.y += bf16_hi(v);
    }
  }
  red[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && c < cols) {
    float2 s = red[0][threadIdx.x];
#pragma unroll
    for (int j = 1; j < 8; ++j) { s.x += red[j][threadIdx.x].x; s.y += red[j][threadIdx.x].y; }
    atomicAdd(out + c, s.x);
    if (c + 1 < cols) atomicAdd(out + c + 1, s.y);
  }
}

// y = dropout(act(x)) on a padded bf16 matrix (standalone Activation / Dropout layers).
__global__ void act_dropout_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, long ld,
                                       int rows, int cols, int act, float p, uint32_t thresh, float scale,
                                       uint64_t seed, uint64_t stream, const long long* step_ptr) {
  if (step_ptr) stream += static_cast<uint64_t>(*step_ptr) * 1024ull;
  const long groups_per_row = (cols + 3) / 4;
  const long total = static_cast<long>(rows) * groups_per_row;
  for (long g = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; g < total;
       g += static_cast<long>(gridDim.x) * blockDim.x) {
    const long r = g / groups_per_row;
    const int c0 = static_cast<int>(g - r * groups_per_row) * 4;
    bool keep[4] = {true, true, true, true};
    if (p > 0.f) dropout_keep4(seed, stream, static_cast<uint64_t>(r) * cols + c0, thresh, keep);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      if (c0 + e < cols) {
        float v = act_apply(act, __bfloat162float(x[r * ld + c0 + e]));
        y[r * ld + c0 + e] = __float2bfloat16_rn(keep[e] ? v * scale : 0.f);
      }
    }
  }
}

// dx = dy * act'(y_pre_dropout) * mask * scale, y being the stored forward output.
__global__ void act_dropout_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y,
                                       __nv_bfloat16* __restrict__ dx, long ld, int rows, int cols, int act, float p,
                                       uint32_t thresh, float scale, uint64_t seed, uint64_t stream,
                                       const long long* step_ptr) {
  if (step_ptr) stream += static_cast<uint64_t>(*step_ptr) * 1024ull;
  const long groups_per_row = (cols + 3) / 4;
  const long total = static_cast<long>(rows) * groups_per_row;
  for (long g = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; g < total;
       g += static_cast<long>(gridDim.x) * blockDim.x) {
    const long r = g / groups_per_row;
    const int c0 = static_cast<int>(g - r * groups_per_row) * 4;
    bool keep[4] = {true, true, true, true};
    if (p > 0.f) dropout_keep4(seed, stream, static_cast<uint64_t>(r) * cols + c0, thresh, keep);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      if (c0 + e < cols) {
        const long i = r * ld + c0 + e;
        const float yy = __bfloat162float(y[i]) * (p > 0.f ? 1.f - p : 1.f);
        const float g_in = __bfloat162float(dy[i]);
        dx[i] = __float2bfloat16_rn(keep[e] ? g_in * scale * act_grad_from_out(act, yy) : 0.f);
      }
    }
  }
}

// K5: delta = before - after over a flat fp32 buffer; written as fp32 and/or bf16.
__global__ void delta_kernel(const float* __restrict__ before, const float* __restrict__ after,
                             float* __restrict__ out32, __nv_bfloat16* __restrict__ out16, long n) {
  const long n4 = n / 4;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const float4 b = reinterpret_cast<const float4*>(before)[i];
    const float4 a = reinterpret_cast<const float4*>(after)[i];
    const float4 d = make_float4(b.x - a.x, b.y - a.y, b.z - a.z, b.w - a.w);
    if (out32) reinterpret_cast<float4*>(out32)[i] = d;
    if (out16) {
      uint2 o;
      o.x = pack_bf16x2(d.x, d.y);
      o.y = pack_bf16x2(d.z, d.w);
      reinterpret_cast<uint2*>(out16)[i] = o;
    }
  }
  for (long i = n4 * 4 + blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const float d = before[i] - after[i];
    if (out32) out32[i] = d;
    if (out16) out16[i] = __float2bfloat16_rn(d);
  }
}

}  // namespace eb

using namespace eb;

extern "C" int eb_stage_f32_to_bf16(const float* src, int64_t ld_src, void* dst, int64_t ld_dst, int64_t rows,
                                    int32_t cols, void* stream) {
  if (rows <= 0) return 0;
  if (ld_dst < cols || ld_src < cols) return fail("eb_stage_f32_to_bf16: leading dimensions smaller than cols");
  stage_f32_to_bf16_kernel<<<grid_for(rows * ld_dst, 256, 4), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      src, ld_src, static_cast<__nv_bfloat16*>(dst), ld_dst, rows, cols);
  EB_LAUNCH_OK("stage_f32_to_bf16_kernel");
  return 0;
}

extern "C" int eb_cast_f32_to_bf16(const float* src, void* dst, int64_t n, void* stream) {
  if (n <= 0) return 0;
  if ((reinterpret_cast<uintptr_t>(src) & 15) || (reinterpret_cast<uintptr_t>(dst) & 15))
    return fail("eb_cast_f32_to_bf16: buffers must be 16-byte aligned");
  cast_f32_to_bf16_kernel<<<grid_for(n / 8 + 1, 256, 2), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      src, static_cast<__nv_bfloat16*>(dst), n);
  EB_LAUNCH_OK("cast_f32_to_bf16_kernel");
  return 0;
}

extern "C" int eb_colsum_bf16(const void* x, int64_t ld, int32_t rows, int32_t cols, float* out, int32_t accumulate,
                              void* stream) {
  if (rows <= 0 || cols <= 0) return 0;
  if (ld % 2) return fail("eb_colsum_bf16: ld must be even");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (!accumulate) EB_CUDA_OK(cudaMemsetAsync(out, 0, sizeof(float) * cols, s));
  dim3 block(32, 8);
  int col_blocks = (cols + 63) / 64;
  int row_blocks = (num_sms() * 4 + col_blocks - 1) / col_blocks;
  if (row_blocks > (rows + 63) / 64) row_blocks = (rows + 63) / 64;
  if (row_blocks < 1) row_blocks = 1;
  colsum_bf16_kernel<<<dim3(col_blocks, row_blocks), block, 0, s>>>(static_cast<const __nv_bfloat16*>(x), ld, rows, cols,
                                                                    out);
  EB_LAUNCH_OK("colsum_bf16_kernel");
  return 0;
}

extern "C" int eb_act_dropout_fwd(const void* x, void* y, int64_t ld, int32_t rows, int32_t cols, int32_t act,
                                  float drop_p, uint64_t seed, uint64_t drop_stream, const void* step_ptr, void* stream) {
  if (rows <= 0 || cols <= 0) return 0;
  const float scale = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  act_dropout_fwd_kernel<<<grid_for(static_cast<long>(rows) * ((cols + 3) / 4), 256, 2), 256, 0,
                           static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), ld, rows, cols, act, drop_p,
      dropout_threshold(drop_p), scale, seed, drop_stream, static_cast<const long long*>(step_ptr));
  EB_LAUNCH_OK("act_dropout_fwd_kernel");
  return 0;
}

extern "C" int eb_act_dropout_bwd(const void* dy, const void* y, void* dx, int64_t ld, int32_t rows, int32_t cols,
                                  int32_t act, float drop_p, uint64_t seed, uint64_t drop_stream, const void* step_ptr,
                                  void* stream) {
  if (rows <= 0 || cols <= 0) return 0;
  const float scale = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  act_dropout_bwd_kernel<<<grid_for(static_cast<long>(rows) * ((cols + 3) / 4), 256, 2), 256, 0,
                           static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(y), static_cast<__nv_bfloat16*>(dx), ld,
      rows, cols, act, drop_p, dropout_threshold(drop_p), scale, seed, drop_stream,
      static_cast<const long long*>(step_ptr));
  EB_LAUNCH_OK("act_dropout_bwd_kernel");
  return 0;
}

extern "C" int eb_delta(const float* before, const float* after, float* out_f32, void* out_bf16, int64_t n,
                        void* stream) {
  if (n <= 0) return 0;
  delta_kernel<<<grid_for(n / 4 + 1, 256, 2), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      before, after, out_f32, static_cast<__nv_bfloat16*>(out_bf16), n);
  EB_LAUNCH_OK("delta_kernel");
  return 0;
}
