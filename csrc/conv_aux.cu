// K9, the bandwidth-bound companions of the convolution GEMMs: BatchNormalization (statistics, apply + residual + ReLU,
// backward) over channels-last bf16 activations, and the patch matrix of the 3-channel input layer.
//
// Layout: activations are [M][C] with C contiguous (M = N*H*W pixels). A thread owns 8 consecutive channels (one
// 16-byte load per pixel); the threads of a block tile (channel groups) x (pixel slots), walk the pixels with a
// block stride and combine their partial sums in shared memory before ONE fp32 atomic per channel and block.
// Reference: Keras BatchNormalization / Activation / Add inside model.fit (elephas/worker.py:42); SURVEY.md K9.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../include/elephas_b200.h"
#include "host.h"
#include "ptx.cuh"

namespace eb {
namespace bn {

constexpr int kThreads = 256;

struct Shape {
  long long M;
  int C, groups, lanes, slots;  // groups = C / 8; lanes = min(groups, 256) threads across channels; slots = 256 / lanes
};

static Shape shape_of(long long M, int C) {
  Shape s;
  s.M = M; s.C = C; s.groups = C / 8;
  s.lanes = s.groups < kThreads ? s.groups : kThreads;
  s.slots = kThreads / s.lanes;
  return s;
}

static int grid_for(const Shape& s, long long max_blocks = 0) {
  long long blocks = (s.M + s.slots * 8 - 1) / (s.slots * 8);  // >= 8 pixels per thread before another block pays off
  long long cap = static_cast<long long>(num_sms()) * 8;  // 8 x 256 threads: every SM fully occupied
  if (max_blocks > 0 && max_blocks < cap) cap = max_blocks;  // (reductions: one row of partial sums per block)
  if (blocks > cap) blocks = cap;
  return blocks < 1 ? 1 : static_cast<int>(blocks);
}

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
__device__ __forceinline__ void load8(const float* p, float (&f)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

// Rows of this thread, kUnroll at a time: `load(m, k)` issues the loads of row m into register set k, `use(k)` consumes
// them. All loads of a batch are issued before the first use: the kernels are latency-bound otherwise (a thread owns
// ~40 rows of 16 bytes, one dependent load at a time hides nothing).
constexpr int kUnroll = 4;
template <typename Load, typename Use>
__device__ __forceinline__ void walk_rows(const Shape& s, int slot, Load load, Use use) {
  const long long stride = static_cast<long long>(gridDim.x) * s.slots;
  long long m = static_cast<long long>(blockIdx.x) * s.slots + slot;
  for (; m + (kUnroll - 1) * stride < s.M; m += kUnroll * stride) {
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) load(m + k * stride, k);
#pragma unroll
    for (int k = 0; k < kUnroll; ++k) use(m + k * stride, k);
  }
  for (; m < s.M; m += stride) {
    load(m, 0);
    use(m, 0);
  }
}

// Combine the per-thread partial sums (two quantities x 8 channels) of all pixel slots of a channel group and store the
// block's totals as row `blockIdx.x` of `partials` ([blocks][2][C]): no atomics (a thousand blocks adding into the same
// few hundred addresses serialise in the L2), no zeroing, and a fixed summation order - a second, tiny kernel adds the
// rows up.
__device__ __forceinline__ void block_reduce_pair(const Shape& s, int lane, int slot, int g, float (&a)[8], float (&b)[8],
                                                  float* partials) {
  __shared__ float sh[kThreads][17];
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sh[threadIdx.x][j] = a[j];
    sh[threadIdx.x][8 + j] = b[j];
  }
  __syncthreads();
  if (slot == 0 && g < s.groups) {
    float ta[8], tb[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      ta[j] = 0.f;
      tb[j] = 0.f;
      for (int k = 0; k < s.slots; ++k) {
        ta[j] += sh[k * s.lanes + lane][j];
        tb[j] += sh[k * s.lanes + lane][8 + j];
      }
    }
    float* row = partials + static_cast<long long>(blockIdx.x) * 2 * s.C + g * 8;
    *reinterpret_cast<float4*>(row) = make_float4(ta[0], ta[1], ta[2], ta[3]);
    *reinterpret_cast<float4*>(row + 4) = make_float4(ta[4], ta[5], ta[6], ta[7]);
    *reinterpret_cast<float4*>(row + s.C) = make_float4(tb[0], tb[1], tb[2], tb[3]);
    *reinterpret_cast<float4*>(row + s.C + 4) = make_float4(tb[4], tb[5], tb[6], tb[7]);
  }
}

// Totals of 32 columns over the rows of `partials` by a block of 32 x 32 threads: thread (x, y) adds rows y, y + 32, ...
// of column c = blockIdx.x * 32 + x (coalesced 128-byte reads, four independent loads in flight), then the 32 row
// groups are combined through shared memory. The result is valid in the threads with threadIdx.y == 0.
__device__ __forceinline__ void sum_partials(const float* __restrict__ partials, int nblocks, int C, float& a, float& b) {
  __shared__ float sh[2][32][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
  if (c < C) {
    int r = threadIdx.y;
    for (; r + 32 < nblocks; r += 64) {
      a0 += partials[(2ll * r) * C + c];
      b0 += partials[(2ll * r + 1) * C + c];
      a1 += partials[(2ll * (r + 32)) * C + c];
      b1 += partials[(2ll * (r + 32) + 1) * C + c];
    }
    if (r < nblocks) {
      a0 += partials[(2ll * r) * C + c];
      b0 += partials[(2ll * r + 1) * C + c];
    }
  }
  sh[0][threadIdx.y][threadIdx.x] = a0 + a1;
  sh[1][threadIdx.y][threadIdx.x] = b0 + b1;
  __syncthreads();
  a = b = 0.f;
  if (threadIdx.y == 0) {
#pragma unroll 8
    for (int k = 0; k < 32; ++k) {
      a += sh[0][k][threadIdx.x];
      b += sh[1][k][threadIdx.x];
    }
  }
}

__global__ void __launch_bounds__(kThreads) stats_kernel(const __nv_bfloat16* __restrict__ x, const Shape s, float* partials) {
  const int lane = threadIdx.x % s.lanes, slot = threadIdx.x / s.lanes;
  for (int g0 = 0; g0 < s.groups; g0 += s.lanes) {  // (one pass unless C > 2048; uniform trip count: barriers inside)
    const int g = g0 + lane;
    float a[8] = {0, 0, 0, 0, 0, 0, 0, 0}, b[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (slot < s.slots && g < s.groups) {
      uint4 raw[kUnroll];
      walk_rows(s, slot, [&](long long m, int k) { raw[k] = ld_nc_u4(x + m * s.C + g * 8); },
                [&](long long, int k) {
                  float f[8];
                  unpack8(raw[k], f);
#pragma unroll
                  for (int j = 0; j < 8; ++j) { a[j] += f[j]; b[j] += f[j] * f[j]; }
                });
    }
    block_reduce_pair(s, lane, slot, g, a, b, partials);
  }
}

__global__ void finalize_kernel(const float* __restrict__ partials, int nblocks, long long M, int C, float eps, float momentum,
                                float* mean, float* rstd, float* running_mean, float* running_var) {
  const int c = blockIdx.x * 32 + threadIdx.x;
  float sum, sumsq;
  sum_partials(partials, nblocks, C, sum, sumsq);
  if (c >= C || threadIdx.y != 0) return;
  const float inv = 1.f / static_cast<float>(M);
  const float mu = sum * inv;
  float var = sumsq * inv - mu * mu;
  var = var > 0.f ? var : 0.f;
  mean[c] = mu;
  rstd[c] = rsqrtf(var + eps);
  if (running_mean != nullptr) running_mean[c] = running_mean[c] * momentum + mu * (1.f - momentum);
  if (running_var != nullptr) {
    const float unbiased = M > 1 ? var * static_cast<float>(M) / static_cast<float>(M - 1) : var;
    running_var[c] = running_var[c] * momentum + unbiased * (1.f - momentum);
  }
}

// sum_dz[c] (+)= sum over blocks, sum_dzx[c] (+)= ...: one writer per channel (the destinations may be gradient buffers)
__global__ void bwd_finalize_kernel(const float* __restrict__ partials, int nblocks, int C, float* sum_dz, float* sum_dzx,
                                    int accumulate) {
  const int c = blockIdx.x * 32 + threadIdx.x;
  float a, b;
  sum_partials(partials, nblocks, C, a, b);
  if (c >= C || threadIdx.y != 0) return;
  sum_dz[c] = accumulate ? sum_dz[c] + a : a;
  sum_dzx[c] = accumulate ? sum_dzx[c] + b : b;
}

__global__ void __launch_bounds__(kThreads)
apply_kernel(const __nv_bfloat16* __restrict__ x, const Shape s, const float* __restrict__ mean, const float* __restrict__ rstd,
             const float* __restrict__ gamma, const float* __restrict__ beta, const __nv_bfloat16* __restrict__ res, int relu,
             __nv_bfloat16* __restrict__ y) {
  const int lane = threadIdx.x % s.lanes, slot = threadIdx.x / s.lanes;
  if (slot >= s.slots) return;
  for (int g = lane; g < s.groups; g += s.lanes) {
    float mu[8], rs[8], ga[8], be[8], scale[8], shift[8];
    load8(mean + g * 8, mu);
    load8(rstd + g * 8, rs);
    if (gamma) load8(gamma + g * 8, ga);
    if (beta) load8(beta + g * 8, be);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      scale[j] = rs[j] * (gamma ? ga[j] : 1.f);
      shift[j] = (beta ? be[j] : 0.f) - mu[j] * scale[j];
    }
    uint4 rx[kUnroll], rr[kUnroll];
    walk_rows(s, slot,
              [&](long long m, int k) {
                const long long off = m * s.C + g * 8;
                rx[k] = ld_nc_u4(x + off);
                if (res != nullptr) rr[k] = ld_nc_u4(res + off);
              },
              [&](long long m, int k) {
                float f[8];
                unpack8(rx[k], f);
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] = f[j] * scale[j] + shift[j];
                if (res != nullptr) {
                  float r[8];
                  unpack8(rr[k], r);
#pragma unroll
                  for (int j = 0; j < 8; ++j) f[j] += r[j];
                }
                if (relu) {
#pragma unroll
                  for (int j = 0; j < 8; ++j) f[j] = f[j] > 0.f ? f[j] : 0.f;
                }
                st_v4(y + m * s.C + g * 8, pack8(f));
              });
  }
}

// dz = dy * (y > 0) [ReLU]; sums[0][c] += sum dz, sums[1][c] += sum dz * xhat
__global__ void __launch_bounds__(kThreads)
bwd_reduce_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ x,
                  const Shape s, const float* __restrict__ mean, const float* __restrict__ rstd, int relu, float* partials) {
  const int lane = threadIdx.x % s.lanes, slot = threadIdx.x / s.lanes;
  for (int g0 = 0; g0 < s.groups; g0 += s.lanes) {
    const int g = g0 + lane;
    float a[8] = {0, 0, 0, 0, 0, 0, 0, 0}, b[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (slot < s.slots && g < s.groups) {
      float mu[8], rs[8];
      load8(mean + g * 8, mu);
      load8(rstd + g * 8, rs);
      uint4 rd[kUnroll], rx[kUnroll], ry[kUnroll];
      walk_rows(s, slot,
                [&](long long m, int k) {
                  const long long off = m * s.C + g * 8;
                  rd[k] = ld_nc_u4(dy + off);
                  rx[k] = ld_nc_u4(x + off);
                  if (relu) ry[k] = ld_nc_u4(y + off);
                },
                [&](long long, int k) {
                  float d[8], xv[8];
                  unpack8(rd[k], d);
                  unpack8(rx[k], xv);
                  if (relu) {
                    float yv[8];
                    unpack8(ry[k], yv);
#pragma unroll
                    for (int j = 0; j < 8; ++j) d[j] = yv[j] > 0.f ? d[j] : 0.f;
                  }
#pragma unroll
                  for (int j = 0; j < 8; ++j) {
                    a[j] += d[j];
                    b[j] += d[j] * (xv[j] - mu[j]) * rs[j];
                  }
                });
    }
    block_reduce_pair(s, lane, slot, g, a, b, partials);
  }
}

__global__ void __launch_bounds__(kThreads)
bwd_apply_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ x,
                 const Shape s, const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ gamma,
                 const float* __restrict__ sum_dz, const float* __restrict__ sum_dzx, int relu, __nv_bfloat16* __restrict__ dx,
                 __nv_bfloat16* __restrict__ dres) {
  const int lane = threadIdx.x % s.lanes, slot = threadIdx.x / s.lanes;
  if (slot >= s.slots) return;
  const float inv = 1.f / static_cast<float>(s.M);
  for (int g = lane; g < s.groups; g += s.lanes) {
    float mu[8], rs[8], ga[8], sa[8], sb[8];
    load8(mean + g * 8, mu);
    load8(rstd + g * 8, rs);
    if (gamma) load8(gamma + g * 8, ga);
    load8(sum_dz + g * 8, sa);
    load8(sum_dzx + g * 8, sb);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      sa[j] *= inv;
      sb[j] *= inv;
      ga[j] = rs[j] * (gamma ? ga[j] : 1.f);
    }
    uint4 rd[kUnroll], rx[kUnroll], ry[kUnroll];
    walk_rows(s, slot,
              [&](long long m, int k) {
                const long long off = m * s.C + g * 8;
                rd[k] = ld_nc_u4(dy + off);
                rx[k] = ld_nc_u4(x + off);
                if (relu) ry[k] = ld_nc_u4(y + off);
              },
              [&](long long m, int k) {
                const long long off = m * s.C + g * 8;
                float d[8], xv[8];
                unpack8(rd[k], d);
                unpack8(rx[k], xv);
                if (relu) {
                  float yv[8];
                  unpack8(ry[k], yv);
#pragma unroll
                  for (int j = 0; j < 8; ++j) d[j] = yv[j] > 0.f ? d[j] : 0.f;
                }
                if (dres != nullptr) st_v4(dres + off, pack8(d));
                float o[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = ga[j] * (d[j] - sa[j] - (xv[j] - mu[j]) * rs[j] * sb[j]);
                st_v4(dx + off, pack8(o));
              });
  }
}

// patch matrix of the input layer: a thread builds 8 consecutive columns of one pixel's row (one 16-byte store)
template <typename T>
__global__ void __launch_bounds__(256)
im2col_kernel(const T* __restrict__ x, int N, int H, int W, int C, int R, int S, int stride, int pad, int P, int Q,
              __nv_bfloat16* __restrict__ out, int ld) {
  const int groups = ld / 8, kvalid = R * S * C;
  const long long total = static_cast<long long>(N) * P * Q * groups;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * 256) {
    const int grp = static_cast<int>(i % groups);
    const long long pix = i / groups;
    const int q = static_cast<int>(pix % Q), p = static_cast<int>((pix / Q) % P), n = static_cast<int>(pix / (static_cast<long long>(Q) * P));
    const T* img = x + static_cast<long long>(n) * H * W * C;
    float v[8];
    int col = grp * 8, tap = col / C, c = col - tap * C;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      v[j] = 0.f;
      if (col + j < kvalid) {
        const int r = tap / S, sx = tap - r * S;
        const int h = p * stride - pad + r, w = q * stride - pad + sx;
        if (h >= 0 && h < H && w >= 0 && w < W) v[j] = static_cast<float>(img[(static_cast<long long>(h) * W + w) * C + c]);
      }
      if (++c == C) { c = 0; ++tap; }
    }
    st_v4(out + pix * ld + grp * 8, pack8(v));
  }
}

static int check(const char* who, const void* x, long long M, int C) {
  if (M <= 0 || C <= 0) return fail("%s: empty tensor", who);
  if (C % 8) return fail("%s: the channel count must be a multiple of 8 (got %d)", who, C);
  if (reinterpret_cast<uintptr_t>(x) & 15) return fail("%s: tensors must be 16-byte aligned", who);
  return 0;
}

}  // namespace bn
}  // namespace eb

using namespace eb;

extern "C" int eb_bn_stats(const void* x, int64_t M, int32_t C, float* partials, int32_t max_blocks, float eps,
                           float momentum, float* mean, float* rstd, float* running_mean, float* running_var, void* stream) {
  if (bn::check("eb_bn_stats", x, M, C)) return 1;
  if (partials == nullptr || max_blocks < 1) return fail("eb_bn_stats: scratch for the per-block partial sums is required");
  const bn::Shape s = bn::shape_of(M, C);
  const int grid = bn::grid_for(s, max_blocks);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  bn::stats_kernel<<<grid, bn::kThreads, 0, st>>>(static_cast<const __nv_bfloat16*>(x), s, partials);
  EB_LAUNCH_OK("bn_stats_kernel");
  bn::finalize_kernel<<<(C + 31) / 32, dim3(32, 32), 0, st>>>(partials, grid, M, C, eps, momentum, mean, rstd, running_mean, running_var);
  EB_LAUNCH_OK("bn_finalize_kernel");
  return 0;
}

extern "C" int eb_bn_apply(const void* x, int64_t M, int32_t C, const float* mean, const float* rstd, const float* gamma,
                           const float* beta, const void* residual, int32_t relu, void* y, void* stream) {
  if (bn::check("eb_bn_apply", x, M, C)) return 1;
  const bn::Shape s = bn::shape_of(M, C);
  bn::apply_kernel<<<bn::grid_for(s), bn::kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), s, mean, rstd, gamma, beta, static_cast<const __nv_bfloat16*>(residual), relu,
      static_cast<__nv_bfloat16*>(y));
  EB_LAUNCH_OK("bn_apply_kernel");
  return 0;
}

extern "C" int eb_bn_bwd_reduce(const void* dy, const void* y, const void* x, int64_t M, int32_t C, const float* mean,
                                const float* rstd, int32_t relu, float* partials, int32_t max_blocks, float* sum_dz,
                                float* sum_dzx, int32_t accumulate, void* stream) {
  if (bn::check("eb_bn_bwd_reduce", x, M, C)) return 1;
  if (relu && y == nullptr) return fail("eb_bn_bwd_reduce: the ReLU mask needs the forward output");
  if (partials == nullptr || max_blocks < 1) return fail("eb_bn_bwd_reduce: scratch for the per-block partial sums is required");
  const bn::Shape s = bn::shape_of(M, C);
  const int grid = bn::grid_for(s, max_blocks);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  bn::bwd_reduce_kernel<<<grid, bn::kThreads, 0, st>>>(
      static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(y), static_cast<const __nv_bfloat16*>(x), s, mean,
      rstd, relu, partials);
  EB_LAUNCH_OK("bn_bwd_reduce_kernel");
  bn::bwd_finalize_kernel<<<(C + 31) / 32, dim3(32, 32), 0, st>>>(partials, grid, C, sum_dz, sum_dzx, accumulate);
  EB_LAUNCH_OK("bn_bwd_finalize_kernel");
  return 0;
}

extern "C" int eb_bn_bwd_apply(const void* dy, const void* y, const void* x, int64_t M, int32_t C, const float* mean,
                               const float* rstd, const float* gamma, const float* sum_dz, const float* sum_dzx, int32_t relu,
                               void* dx, void* dres, void* stream) {
  if (bn::check("eb_bn_bwd_apply", x, M, C)) return 1;
  if (relu && y == nullptr) return fail("eb_bn_bwd_apply: the ReLU mask needs the forward output");
  const bn::Shape s = bn::shape_of(M, C);
  bn::bwd_apply_kernel<<<bn::grid_for(s), bn::kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(y), static_cast<const __nv_bfloat16*>(x), s, mean,
      rstd, gamma, sum_dz, sum_dzx, relu, static_cast<__nv_bfloat16*>(dx), static_cast<__nv_bfloat16*>(dres));
  EB_LAUNCH_OK("bn_bwd_apply_kernel");
  return 0;
}

extern "C" int eb_conv_im2col(const void* x, int32_t x_is_f32, int32_t N, int32_t H, int32_t W, int32_t C, int32_t R,
                              int32_t S, int32_t stride, int32_t pad, int32_t P, int32_t Q, void* out_bf16, int32_t ld,
                              void* stream) {
  if (N <= 0 || P <= 0 || Q <= 0 || ld < R * S * C || (ld % 8)) return fail("eb_conv_im2col: bad geometry (ld must be a multiple of 8 and >= R*S*C)");
  const long long total = static_cast<long long>(N) * P * Q * (ld / 8);
  long long blocks = (total + 255) / 256;
  const long long cap = static_cast<long long>(num_sms()) * 16;
  if (blocks > cap) blocks = cap;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (x_is_f32)
    bn::im2col_kernel<float><<<static_cast<int>(blocks), 256, 0, s>>>(static_cast<const float*>(x), N, H, W, C, R, S, stride, pad, P, Q,
                                                                      static_cast<__nv_bfloat16*>(out_bf16), ld);
  else
    bn::im2col_kernel<__nv_bfloat16><<<static_cast<int>(blocks), 256, 0, s>>>(static_cast<const __nv_bfloat16*>(x), N, H, W, C, R, S,
                                                                              stride, pad, P, Q, static_cast<__nv_bfloat16*>(out_bf16), ld);
  EB_LAUNCH_OK("im2col_kernel");
  return 0;
}
