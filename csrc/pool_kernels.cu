// K9, pooling over channels-last bf16 activations [N][H][W][C]: max pooling (any window / stride / padding, the padding
// value either -inf - Keras padding='same' - or 0 - a ZeroPadding2D folded into the pool) and global average pooling,
// forward and backward. A thread owns 8 consecutive channels of one pixel (16-byte loads / stores).
//
// Max pooling keeps the winning tap of every output element in a byte map; the backward kernel is a GATHER over the
// input pixels (each input pixel looks at the few windows that contain it and adds the gradients of those it won), so
// it needs no atomics and its result does not depend on the launch geometry.
// Reference: Keras MaxPooling2D / GlobalAveragePooling2D inside model.fit (elephas/worker.py:42); SURVEY.md K9.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../include/elephas_b200.h"
#include "host.h"
#include "ptx.cuh"

namespace eb {
namespace pool {

struct Geo {
  int N, H, W, C, R, S, stride, pad_t, pad_l, P, Q;
};

constexpr unsigned kPadTap = 255u;  // the window's maximum is the padding value: no input element receives the gradient

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 o;
  o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]);
  o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
  return o;
}

// y[n][p][q][c] = max over the window (first maximum in row-major tap order wins, like the library kernels); idx = tap
__global__ void __launch_bounds__(256)
max_fwd_kernel(const __nv_bfloat16* __restrict__ x, const Geo g, int pad_is_zero, __nv_bfloat16* __restrict__ y,
               uint8_t* __restrict__ idx) {
  const int groups = g.C / 8;
  const long long total = static_cast<long long>(g.N) * g.P * g.Q * groups;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * 256) {
    const int grp = static_cast<int>(i % groups);
    const long long pix = i / groups;
    const int q = static_cast<int>(pix % g.Q), p = static_cast<int>((pix / g.Q) % g.P);
    const long long n = pix / (static_cast<long long>(g.Q) * g.P);
    const __nv_bfloat16* img = x + n * g.H * g.W * g.C + grp * 8;
    const int h0 = p * g.stride - g.pad_t, w0 = q * g.stride - g.pad_l;
    float best[8];
    unsigned tap[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; tap[j] = kPadTap; }
    for (int r = 0; r < g.R; ++r) {
      const int h = h0 + r;
      for (int s = 0; s < g.S; ++s) {
        const int w = w0 + s;
        if (h < 0 || h >= g.H || w < 0 || w >= g.W) {
          // a padding element takes part like any other tap when it is a zero (ties go to the earlier tap)
          if (pad_is_zero) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
              if (0.f > best[j]) { best[j] = 0.f; tap[j] = kPadTap; }
          }
          continue;
        }
        float v[8];
        unpack8(ld_nc_u4(img + (static_cast<long long>(h) * g.W + w) * g.C), v);
        const unsigned t = static_cast<unsigned>(r * g.S + s);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (v[j] > best[j]) { best[j] = v[j]; tap[j] = t; }
      }
    }
    st_v4(y + pix * g.C + grp * 8, pack8(best));
    uint2 packed;
    packed.x = tap[0] | (tap[1] << 8) | (tap[2] << 16) | (tap[3] << 24);
    packed.y = tap[4] | (tap[5] << 8) | (tap[6] << 16) | (tap[7] << 24);
    *reinterpret_cast<uint2*>(idx + pix * g.C + grp * 8) = packed;
  }
}

// dx[n][h][w][c] = sum of dy over the windows that contain (h, w) and whose maximum is this element
__global__ void __launch_bounds__(256)
max_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const uint8_t* __restrict__ idx, const Geo g,
               __nv_bfloat16* __restrict__ dx) {
  const int groups = g.C / 8;
  const long long total = static_cast<long long>(g.N) * g.H * g.W * groups;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * 256) {
    const int grp = static_cast<int>(i % groups);
    const long long pix = i / groups;
    const int w = static_cast<int>(pix % g.W), h = static_cast<int>((pix / g.W) % g.H);
    const long long n = pix / (static_cast<long long>(g.W) * g.H);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    // windows p with p * stride - pad_t <= h < p * stride - pad_t + R
    const int hp = h + g.pad_t, wp = w + g.pad_l;
    int p_lo = hp - g.R + 1;
    p_lo = p_lo <= 0 ? 0 : (p_lo + g.stride - 1) / g.stride;
    int p_hi = hp / g.stride;
    if (p_hi > g.P - 1) p_hi = g.P - 1;
    int q_lo = wp - g.S + 1;
    q_lo = q_lo <= 0 ? 0 : (q_lo + g.stride - 1) / g.stride;
    int q_hi = wp / g.stride;
    if (q_hi > g.Q - 1) q_hi = g.Q - 1;
    for (int p = p_lo; p <= p_hi; ++p) {
      const int r = hp - p * g.stride;
      for (int q = q_lo; q <= q_hi; ++q) {
        const unsigned t = static_cast<unsigned>(r * g.S + (wp - q * g.stride));
        const long long o = ((n * g.P + p) * g.Q + q) * g.C + grp * 8;
        const uint2 taps = *reinterpret_cast<const uint2*>(idx + o);
        float d[8];
        unpack8(ld_nc_u4(dy + o), d);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const unsigned tj = ((j < 4 ? taps.x : taps.y) >> (8 * (j & 3))) & 0xFFu;
          if (tj == t) acc[j] += d[j];
        }
      }
    }
    st_v4(dx + pix * g.C + grp * 8, pack8(acc));
  }
}

// y[n][c] = mean over the HW pixels of image n (fp32 accumulation)
__global__ void __launch_bounds__(256)
gap_fwd_kernel(const __nv_bfloat16* __restrict__ x, int N, int HW, int C, __nv_bfloat16* __restrict__ y) {
  const int groups = C / 8;
  const long long total = static_cast<long long>(N) * groups;
  const float inv = 1.f / static_cast<float>(HW);
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * 256) {
    const int grp = static_cast<int>(i % groups);
    const long long n = i / groups;
    const __nv_bfloat16* img = x + n * HW * C + grp * 8;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int m = 0;
    for (; m + 4 <= HW; m += 4) {  // four rows in flight
      uint4 raw[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) raw[k] = ld_nc_u4(img + static_cast<long long>(m + k) * C);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float v[8];
        unpack8(raw[k], v);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += v[j];
      }
    }
    for (; m < HW; ++m) {
      float v[8];
      unpack8(ld_nc_u4(img + static_cast<long long>(m) * C), v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] *= inv;
    st_v4(y + n * C + grp * 8, pack8(acc));
  }
}

// dx[n][m][c] = dy[n][c] / HW
__global__ void __launch_bounds__(256)
gap_bwd_kernel(const __nv_bfloat16* __restrict__ dy, int N, int HW, int C, __nv_bfloat16* __restrict__ dx) {
  const int groups = C / 8;
  const long long total = static_cast<long long>(N) * HW * groups;
  const float inv = 1.f / static_cast<float>(HW);
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * 256) {
    const int grp = static_cast<int>(i % groups);
    const long long pix = i / groups, n = pix / HW;
    float v[8];
    unpack8(ld_nc_u4(dy + n * C + grp * 8), v);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] *= inv;
    st_v4(dx + pix * C + grp * 8, pack8(v));
  }
}

static int grid_of(long long items) {
  long long blocks = (items + 255) / 256;
  const long long cap = static_cast<long long>(num_sms()) * 16;
  if (blocks > cap) blocks = cap;
  return blocks < 1 ? 1 : static_cast<int>(blocks);
}

static int check_geo(const char* who, const Geo& g, const void* a, const void* b) {
  if (g.N <= 0 || g.H <= 0 || g.W <= 0 || g.C <= 0 || g.P <= 0 || g.Q <= 0) return fail("%s: empty tensor", who);
  if (g.C % 8) return fail("%s: the channel count must be a multiple of 8 (got %d)", who, g.C);
  if (g.R < 1 || g.S < 1 || g.R * g.S >= static_cast<int>(kPadTap)) return fail("%s: window %dx%d not supported", who, g.R, g.S);
  if (g.stride < 1 || g.pad_t < 0 || g.pad_l < 0 || g.pad_t >= g.R || g.pad_l >= g.S)
    return fail("%s: bad stride / padding (stride %d, pads %d %d)", who, g.stride, g.pad_t, g.pad_l);
  if ((g.P - 1) * g.stride - g.pad_t >= g.H || (g.Q - 1) * g.stride - g.pad_l >= g.W)
    return fail("%s: output %dx%d does not fit the input %dx%d", who, g.P, g.Q, g.H, g.W);
  if ((reinterpret_cast<uintptr_t>(a) & 15) || (reinterpret_cast<uintptr_t>(b) & 15))
    return fail("%s: tensors must be 16-byte aligned", who);
  return 0;
}

}  // namespace pool
}  // namespace eb

using namespace eb;

extern "C" int eb_pool_max_fwd(const void* x, int32_t N, int32_t H, int32_t W, int32_t C, int32_t R, int32_t S, int32_t stride,
                               int32_t pad_t, int32_t pad_l, int32_t P, int32_t Q, int32_t pad_is_zero, void* y, void* idx,
                               void* stream) {
  const pool::Geo g{N, H, W, C, R, S, stride, pad_t, pad_l, P, Q};
  if (pool::check_geo("eb_pool_max_fwd", g, x, y)) return 1;
  if (idx == nullptr || (reinterpret_cast<uintptr_t>(idx) & 7)) return fail("eb_pool_max_fwd: the tap map must be 8-byte aligned");
  const long long items = static_cast<long long>(N) * P * Q * (C / 8);
  pool::max_fwd_kernel<<<pool::grid_of(items), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), g, pad_is_zero, static_cast<__nv_bfloat16*>(y), static_cast<uint8_t*>(idx));
  EB_LAUNCH_OK("pool_max_fwd_kernel");
  return 0;
}

extern "C" int eb_pool_max_bwd(const void* dy, const void* idx, int32_t N, int32_t H, int32_t W, int32_t C, int32_t R, int32_t S,
                               int32_t stride, int32_t pad_t, int32_t pad_l, int32_t P, int32_t Q, void* dx, void* stream) {
  const pool::Geo g{N, H, W, C, R, S, stride, pad_t, pad_l, P, Q};
  if (pool::check_geo("eb_pool_max_bwd", g, dy, dx)) return 1;
  if (idx == nullptr || (reinterpret_cast<uintptr_t>(idx) & 7)) return fail("eb_pool_max_bwd: the tap map must be 8-byte aligned");
  const long long items = static_cast<long long>(N) * H * W * (C / 8);
  pool::max_bwd_kernel<<<pool::grid_of(items), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy), static_cast<const uint8_t*>(idx), g, static_cast<__nv_bfloat16*>(dx));
  EB_LAUNCH_OK("pool_max_bwd_kernel");
  return 0;
}

extern "C" int eb_pool_gap_fwd(const void* x, int32_t N, int32_t HW, int32_t C, void* y, void* stream) {
  if (N <= 0 || HW <= 0 || C <= 0 || (C % 8)) return fail("eb_pool_gap_fwd: bad shape (C must be a multiple of 8)");
  if ((reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(y) & 15)) return fail("eb_pool_gap_fwd: tensors must be 16-byte aligned");
  pool::gap_fwd_kernel<<<pool::grid_of(static_cast<long long>(N) * (C / 8)), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), N, HW, C, static_cast<__nv_bfloat16*>(y));
  EB_LAUNCH_OK("pool_gap_fwd_kernel");
  return 0;
}

extern "C" int eb_pool_gap_bwd(const void* dy, int32_t N, int32_t HW, int32_t C, void* dx, void* stream) {
  if (N <= 0 || HW <= 0 || C <= 0 || (C % 8)) return fail("eb_pool_gap_bwd: bad shape (C must be a multiple of 8)");
  if ((reinterpret_cast<uintptr_t>(dy) & 15) || (reinterpret_cast<uintptr_t>(dx) & 15)) return fail("eb_pool_gap_bwd: tensors must be 16-byte aligned");
  pool::gap_bwd_kernel<<<pool::grid_of(static_cast<long long>(N) * HW * (C / 8)), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy), N, HW, C, static_cast<__nv_bfloat16*>(dx));
  EB_LAUNCH_OK("pool_gap_bwd_kernel");
  return 0;
}
