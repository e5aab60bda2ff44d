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

// Grid of the row-strided vector kernels: gx blocks across the columns, enough row lanes for ~8 blocks per SM.
static inline dim3 rows_grid(int gx, long rows) {
  long gy = (static_cast<long>(num_sms()) * 8 + gx - 1) / gx;
  if (gy > rows) gy = rows;
  if (gy > 65535) gy = 65535;
  if (gy < 1) gy = 1;
  return dim3(gx, static_cast<unsigned>(gy));
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

// Vector path of the above (cols, ld_dst multiples of 8, ld_src of 4, 16-byte aligned bases): a thread converts 8
// consecutive elements (two 16-byte loads, one 16-byte store); x indexes column groups, y strides over rows, so
// there is no integer division and four rows are in flight per thread.
__global__ void __launch_bounds__(128) stage_f32_to_bf16_vec_kernel(const float* __restrict__ src, long ld_src,
                                                                    __nv_bfloat16* __restrict__ dst, long ld_dst,
                                                                    int rows, int cols) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (c >= ld_dst) return;
  const bool live = c < cols;
#pragma unroll 4
  for (int r = blockIdx.y; r < rows; r += gridDim.y) {
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    if (live) {
      const float4 a = ld_nc_f4(src + r * ld_src + c);
      const float4 b = ld_nc_f4(src + r * ld_src + c + 4);
      o.x = pack_bf16x2(a.x, a.y); o.y = pack_bf16x2(a.z, a.w);
      o.z = pack_bf16x2(b.x, b.y); o.w = pack_bf16x2(b.z, b.w);
    }
    *reinterpret_cast<uint4*>(dst + r * ld_dst + c) = o;
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

// Vector path (cols, ld multiples of 8, 16-byte aligned): block = 32 x 8 threads owns 256 columns (16-byte loads,
// 8 columns per thread) and a slab of rows, four rows in flight per thread; the eight row-partials per column meet
// in shared memory and every thread publishes one column with a single atomicAdd.
__global__ void __launch_bounds__(256) colsum_bf16_vec_kernel(const __nv_bfloat16* __restrict__ x, long ld, int rows,
                                                              int cols, float* __restrict__ out) {
  __shared__ float red[8][256];
  const int c = blockIdx.x * 256 + threadIdx.x * 8;
  const int rows_per_block = (rows + gridDim.y - 1) / gridDim.y;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(rows, r0 + rows_per_block);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (c < cols) {
#pragma unroll 4
    for (int r = r0 + threadIdx.y; r < r1; r += 8) {
      const uint4 v = ld_nc_u4(x + static_cast<long>(r) * ld + c);
      acc[0] += bf16_lo(v.x); acc[1] += bf16_hi(v.x); acc[2] += bf16_lo(v.y); acc[3] += bf16_hi(v.y);
      acc[4] += bf16_lo(v.z); acc[5] += bf16_hi(v.z); acc[6] += bf16_lo(v.w); acc[7] += bf16_hi(v.w);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[threadIdx.y][threadIdx.x * 8 + j] = acc[j];
  __syncthreads();
  const int t = threadIdx.y * 32 + threadIdx.x;
  const int col = blockIdx.x * 256 + t;
  if (col < cols) {
    float s = red[0][t];
#pragma unroll
    for (int j = 1; j < 8; ++j) s += red[j][t];
    atomicAdd(out + col, s);
  }
}

// Vector paths of the two kernels below (cols, ld multiples of 8): 8 elements per thread, 16-byte accesses, the
// activation switch hoisted out of the element loop, the same Philox counters (two groups of four per thread).
__global__ void __launch_bounds__(128) act_dropout_fwd_vec_kernel(const __nv_bfloat16* __restrict__ x,
                                                                  __nv_bfloat16* __restrict__ y, long ld, int rows,
                                                                  int cols, int act, float p, uint32_t thresh, float scale,
                                                                  uint64_t seed, uint64_t stream,
                                                                  const long long* step_ptr) {
  if (step_ptr) stream += static_cast<uint64_t>(*step_ptr) * 1024ull;
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (c >= cols) return;
#pragma unroll 4
  for (int r = blockIdx.y; r < rows; r += gridDim.y) {
    const uint4 v = ld_nc_u4(x + static_cast<long>(r) * ld + c);
    float f[8] = {bf16_lo(v.x), bf16_hi(v.x), bf16_lo(v.y), bf16_hi(v.y),
                  bf16_lo(v.z), bf16_hi(v.z), bf16_lo(v.w), bf16_hi(v.w)};
    act_apply_chunk<8>(act, f);
    if (p > 0.f) {
      bool k0[4], k1[4];
      dropout_keep4(seed, stream, static_cast<uint64_t>(r) * cols + c, thresh, k0);
      dropout_keep4(seed, stream, static_cast<uint64_t>(r) * cols + c + 4, thresh, k1);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        f[e] = k0[e] ? f[e] * scale : 0.f;
        f[e + 4] = k1[e] ? f[e + 4] * scale : 0.f;
      }
    }
    uint4 o;
    o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]);
    o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
    *reinterpret_cast<uint4*>(y + static_cast<long>(r) * ld + c) = o;
  }
}

__global__ void __launch_bounds__(128) act_dropout_bwd_vec_kernel(const __nv_bfloat16* __restrict__ dy,
                                                                  const __nv_bfloat16* __restrict__ y,
                                                                  __nv_bfloat16* __restrict__ dx, long ld, int rows,
                                                                  int cols, int act, float p, uint32_t thresh, float scale,
                                                                  uint64_t seed, uint64_t stream,
                                                                  const long long* step_ptr) {
  if (step_ptr) stream += static_cast<uint64_t>(*step_ptr) * 1024ull;
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (c >= cols) return;
  const float unscale = p > 0.f ? 1.f - p : 1.f;
#pragma unroll 4
  for (int r = blockIdx.y; r < rows; r += gridDim.y) {
    const long i = static_cast<long>(r) * ld + c;
    const uint4 vy = ld_nc_u4(y + i), vg = ld_nc_u4(dy + i);
    float yy[8] = {bf16_lo(vy.x), bf16_hi(vy.x), bf16_lo(vy.y), bf16_hi(vy.y),
                   bf16_lo(vy.z), bf16_hi(vy.z), bf16_lo(vy.w), bf16_hi(vy.w)};
    float g[8] = {bf16_lo(vg.x), bf16_hi(vg.x), bf16_lo(vg.y), bf16_hi(vg.y),
                  bf16_lo(vg.z), bf16_hi(vg.z), bf16_lo(vg.w), bf16_hi(vg.w)};
    float d[8];
    act_grad_chunk<8>(act, yy, unscale, d);  // act'(forward output before the dropout scaling)
    bool k0[4] = {true, true, true, true}, k1[4] = {true, true, true, true};
    if (p > 0.f) {
      dropout_keep4(seed, stream, static_cast<uint64_t>(r) * cols + c, thresh, k0);
      dropout_keep4(seed, stream, static_cast<uint64_t>(r) * cols + c + 4, thresh, k1);
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      g[e] = k0[e] ? g[e] * scale * d[e] : 0.f;
      g[e + 4] = k1[e] ? g[e + 4] * scale * d[e + 4] : 0.f;
    }
    uint4 o;
    o.x = pack_bf16x2(g[0], g[1]); o.y = pack_bf16x2(g[2], g[3]);
    o.z = pack_bf16x2(g[4], g[5]); o.w = pack_bf16x2(g[6], g[7]);
    *reinterpret_cast<uint4*>(dx + i) = o;
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
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool vec = cols % 8 == 0 && ld_dst % 8 == 0 && ld_src % 4 == 0 && rows < (1L << 31) &&
                   !(reinterpret_cast<uintptr_t>(src) & 15) && !(reinterpret_cast<uintptr_t>(dst) & 15);
  if (vec) {
    const int gx = static_cast<int>((ld_dst / 8 + 127) / 128);
    stage_f32_to_bf16_vec_kernel<<<rows_grid(gx, rows), 128, 0, s>>>(src, ld_src, static_cast<__nv_bfloat16*>(dst), ld_dst,
                                                                   static_cast<int>(rows), cols);
    EB_LAUNCH_OK("stage_f32_to_bf16_vec_kernel");
    return 0;
  }
  stage_f32_to_bf16_kernel<<<grid_for(rows * ld_dst, 256, 4), 256, 0, s>>>(
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
  if (cols % 8 == 0 && ld % 8 == 0 && !(reinterpret_cast<uintptr_t>(x) & 15)) {
    const int cb = (cols + 255) / 256;
    int rb = (num_sms() * 4 + cb - 1) / cb;
    if (rb > (rows + 31) / 32) rb = (rows + 31) / 32;
    if (rb < 1) rb = 1;
    colsum_bf16_vec_kernel<<<dim3(cb, rb), block, 0, s>>>(static_cast<const __nv_bfloat16*>(x), ld, rows, cols, out);
    EB_LAUNCH_OK("colsum_bf16_vec_kernel");
    return 0;
  }
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
  if (cols % 8 == 0 && ld % 8 == 0 && !(reinterpret_cast<uintptr_t>(x) & 15) && !(reinterpret_cast<uintptr_t>(y) & 15)) {
    const int gx = (cols / 8 + 127) / 128;
    act_dropout_fwd_vec_kernel<<<rows_grid(gx, rows), 128, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), ld, rows, cols, act, drop_p,
        dropout_threshold(drop_p), scale, seed, drop_stream, static_cast<const long long*>(step_ptr));
    EB_LAUNCH_OK("act_dropout_fwd_vec_kernel");
    return 0;
  }
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
  if (cols % 8 == 0 && ld % 8 == 0 && !(reinterpret_cast<uintptr_t>(dy) & 15) && !(reinterpret_cast<uintptr_t>(y) & 15) &&
      !(reinterpret_cast<uintptr_t>(dx) & 15)) {
    const int gx = (cols / 8 + 127) / 128;
    act_dropout_bwd_vec_kernel<<<rows_grid(gx, rows), 128, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(y), static_cast<__nv_bfloat16*>(dx), ld,
        rows, cols, act, drop_p, dropout_threshold(drop_p), scale, seed, drop_stream,
        static_cast<const long long*>(step_ptr));
    EB_LAUNCH_OK("act_dropout_bwd_vec_kernel");
    return 0;
  }
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
