// K9: convolutions as implicit GEMMs on tcgen05 (forward, input gradient, weight gradient).
//
// The reference trains any Keras model through model.fit on the workers (elephas/worker.py:42,108); BASELINE.json
// names a ResNet-50-shaped CNN. Here a convolution over channels-last bf16 activations is a GEMM whose A operand is
// never materialised: for every filter tap (r, s) and 64-channel slice, ONE TMA load fetches a box of
// pn x ph x pw output-pixel positions x 64 channels from a 4-D tensor map over the activation - shifted by the tap,
// sampled with the convolution stride (element strides of the map), zero filled outside the image (= padding) -
// straight into the 128B-swizzled K-major layout that tcgen05.mma reads. Weights are the OHWI kernel itself, read
// K-major for the forward pass and N-major for the input gradient (no transposed copy); the weight gradient reduces
// over pixel patches with both operands MN-major and spreads the reduction over CTAs (fp32 red.add epilogue).
//
// Kernel structure = the Dense GEMM's (gemm_sm100.cu): persistent CTAs, warp 0 TMA producer, warp 1 single-lane MMA
// issuer that owns the TMEM allocation, 8 epilogue warps, smem full/empty ring + double-buffered TMEM accumulators.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <mutex>

#include "../include/elephas_b200.h"
#include "host.h"
#include "ptx.cuh"

namespace eb {
namespace conv {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;
constexpr int UMMA_K = 16;
constexpr int kEpiWarps = 8;
constexpr int kThreads = 64 + 32 * kEpiWarps;
constexpr int kAccStages = 2;

__device__ __forceinline__ void tma_load_4d(const CUtensorMap* map, uint64_t* bar, void* smem, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(const CUtensorMap* map, uint64_t* bar, void* smem, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void red_add_f32(float* p, float v) {
  asm volatile("red.relaxed.gpu.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void red_add_v4_f32(float* p, float a, float b, float c, float d) {
  asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// ------------------------------------------------------------------ forward / input gradient
struct IgemmDev {
  int N, P, Q;           // patch-origin grid
  int C, K;              // reduction channels, output channels
  int R, S, stride, pad_h, pad_w, flip_taps;
  int pn, ph, pw;        // pixels per tile
  int tiles_n, tiles_h, tiles_w, tiles_k;  // tile counts over images, rows, columns, output channels
  int cblocks;           // 64-channel slices of the reduction
  __nv_bfloat16* out;
  int OH, OW, os;
  const float* bias;
};

template <int BN>
struct IgemmCfg {
  static constexpr int A_STAGE = BLOCK_M * BLOCK_K * 2;
  static constexpr int B_STAGE = BN * BLOCK_K * 2;
  static constexpr int STAGE = A_STAGE + B_STAGE;
  static constexpr int STAGES_RAW = (200 * 1024) / STAGE;
  static constexpr int STAGES = STAGES_RAW > 6 ? 6 : STAGES_RAW;
  static constexpr int TMEM_COLS = kAccStages * BN <= 128 ? 128 : (kAccStages * BN <= 256 ? 256 : 512);
  static constexpr int BAR_BYTES = (2 * STAGES + 2 * kAccStages) * 8 + 16;
  static constexpr int SMEM_BYTES = STAGES * STAGE + BAR_BYTES + 1024;
};

// W_MN = false: weights [K][taps][C] read K-major (forward). W_MN = true: weights [C_red][taps][K_out] read N-major
// (input gradient: the forward kernel array, its leading dimension is now the reduction).
template <int BN, bool W_MN>
__global__ void __launch_bounds__(kThreads, 1)
conv_igemm_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w, const IgemmDev p) {
  using Cfg = IgemmCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + Cfg::STAGES * Cfg::A_STAGE;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::STAGES * Cfg::STAGE);
  uint64_t* empty_bar = full_bar + Cfg::STAGES;
  uint64_t* tfull_bar = empty_bar + Cfg::STAGES;
  uint64_t* tempty_bar = tfull_bar + kAccStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + kAccStages);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_x);
    tma_prefetch_desc(&map_w);
    for (int i = 0; i < Cfg::STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < kAccStages; ++i) {
      mbar_init(&tfull_bar[i], 1);
      mbar_init(&tempty_bar[i], kEpiWarps);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc<1>(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  const int m_tiles = p.tiles_n * p.tiles_h * p.tiles_w;
  const int num_tiles = m_tiles * p.tiles_k;
  const int num_kb = p.R * p.S * p.cblocks;
  const int rows_box = p.pn * p.ph * p.pw;
  const uint32_t stage_bytes = static_cast<uint32_t>(rows_box) * (BLOCK_K * 2) + Cfg::B_STAGE;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        const int kt = t % p.tiles_k, mt = t / p.tiles_k;
        const int tw = mt % p.tiles_w, th = (mt / p.tiles_w) % p.tiles_h, tn = mt / (p.tiles_w * p.tiles_h);
        const int q0 = tw * p.pw, p0 = th * p.ph, n0 = tn * p.pn, k0 = kt * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          const int tap = kb / p.cblocks, c0 = (kb - tap * p.cblocks) * BLOCK_K;
          const int r = tap / p.S, s = tap - r * p.S;
          const int wtap = p.flip_taps ? p.R * p.S - 1 - tap : tap;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem_a + stage * Cfg::A_STAGE;
          uint8_t* sb = smem_b + stage * Cfg::B_STAGE;
          mbar_arrive_expect_tx(&full_bar[stage], stage_bytes);
          // A: the (r, s)-shifted patch of pn x ph x pw pixel positions x 64 channels; rows land densely, row = pixel
          tma_load_4d(&map_x, &full_bar[stage], sa, c0, q0 * p.stride - p.pad_w + s, p0 * p.stride - p.pad_h + r, n0);
          if constexpr (!W_MN) {
            tma_load_3d(&map_w, &full_bar[stage], sb, c0, wtap, k0);  // [BN output channels][64 reduction channels]
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)  // [64 reduction channels][64 output channels] per 64-wide N block
              tma_load_3d(&map_w, &full_bar[stage], sb + j * (64 * BLOCK_K * 2), k0 + 64 * j, wtap, c0);
          }
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // -------------------------------------------------------------- MMA issuer
    const uint32_t idesc = make_idesc_bf16(BLOCK_M, BN, 0, W_MN ? 1 : 0);
    const uint32_t b_lbo = W_MN ? 64 * BLOCK_K * 2 : 16;
    const uint32_t a_kstep = (UMMA_K * 2) >> 4, b_kstep = (W_MN ? UMMA_K * 128 : UMMA_K * 2) >> 4;
    const uint64_t desc_a0 = make_smem_desc_sw128(smem_u32(smem_a), 16, 1024);
    const uint64_t desc_b0 = make_smem_desc_sw128(smem_u32(smem_b), b_lbo, 1024);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * BN;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t da0 = desc_a0 + static_cast<uint32_t>(stage * (Cfg::A_STAGE >> 4));
          const uint64_t db0 = desc_b0 + static_cast<uint32_t>(stage * (Cfg::B_STAGE >> 4));
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
            umma_bf16<1>(tmem_d, da0 + k * a_kstep, db0 + k * b_kstep, idesc, (kb | k) != 0 ? 1u : 0u);
          umma_commit<1>(&empty_bar[stage]);
          if (kb == num_kb - 1) umma_commit<1>(&tfull_bar[acc]);
        }
        __syncwarp();
        if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
      }
      if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
    }
  } else {
    // ---------------------------------------------------------------- epilogue
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = quarter * 32 + lane;
    // pixel of this row inside the patch (rows >= rows_box hold stale shared memory: never stored)
    const int rpw = row % p.pw, rph = (row / p.pw) % p.ph, rpn = row / (p.pw * p.ph);
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      const int kt = t % p.tiles_k, mt = t / p.tiles_k;
      const int tw = mt % p.tiles_w, th = (mt / p.tiles_w) % p.tiles_h, tn = mt / (p.tiles_w * p.tiles_h);
      const int q = tw * p.pw + rpw, pp = th * p.ph + rph, n = tn * p.pn + rpn;
      const bool row_ok = row < rows_box && q < p.Q && pp < p.P && n < p.N;
      __nv_bfloat16* orow =
          p.out + ((static_cast<size_t>(n) * p.OH + static_cast<size_t>(pp) * p.os) * p.OW + static_cast<size_t>(q) * p.os) * p.K;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BN;
      constexpr int NCHUNK = BN / 32;
      constexpr int PER = NCHUNK / 2;  // BN >= 64: every warp has at least one 32-column chunk
#pragma unroll 1
      for (int c = half * PER; c < (half + 1) * PER; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(taddr + c * 32, v);
        tmem_ld_wait();
        if (c == (half + 1) * PER - 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty_bar[acc]);
        }
        const int col0 = kt * BN + c * 32;
        if (!row_ok || col0 >= p.K) continue;
        float x[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) x[j] = __uint_as_float(v[j]);
        const bool full = col0 + 32 <= p.K;
        if (p.bias != nullptr) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (full || col0 + j < p.K) x[j] += __ldg(p.bias + col0 + j);
        }
        if (full) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            uint4 o;
            o.x = pack_bf16x2(x[j + 0], x[j + 1]); o.y = pack_bf16x2(x[j + 2], x[j + 3]);
            o.z = pack_bf16x2(x[j + 4], x[j + 5]); o.w = pack_bf16x2(x[j + 6], x[j + 7]);
            st_v4(orow + col0 + j, o);
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (col0 + j < p.K) orow[col0 + j] = __float2bfloat16_rn(x[j]);
        }
      }
      if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, Cfg::TMEM_COLS);
  }
}

// ------------------------------------------------------------------ weight gradient
// dW[k][tap][c] = sum over pixels of dy[pixel][k] * x[pixel shifted by tap][c]: M = output channels (tile 128),
// N = input channels of one tap (tile BN), reduction over pixel patches of <= 64 positions. Both operands are MN-major
// (channels contiguous). A work item is (k tile, c tile, tap, split): `splits` CTAs share the pixel reduction of one
// output tile and add their partial sums into dW with red.add (dW zeroed by the caller).
struct WgradDev {
  int N, P, Q, C, K, R, S, stride, pad_h, pad_w;
  int pn, ph, pw;
  int tiles_n, tiles_h, tiles_w;  // pixel patches over images, rows, columns
  int tiles_k, tiles_c, splits;
  float* dw;
};

template <int BN>
struct WgradCfg {
  static constexpr int A_STAGE = BLOCK_K * BLOCK_M * 2;  // [64 pixels][128 output channels] as two 64-wide blocks
  static constexpr int B_STAGE = BLOCK_K * BN * 2;       // [64 pixels][BN input channels]
  static constexpr int STAGE = A_STAGE + B_STAGE;
  static constexpr int STAGES_RAW = (200 * 1024) / STAGE;
  static constexpr int STAGES = STAGES_RAW > 6 ? 6 : STAGES_RAW;
  static constexpr int TMEM_COLS = kAccStages * BN <= 128 ? 128 : (kAccStages * BN <= 256 ? 256 : 512);
  static constexpr int BAR_BYTES = (2 * STAGES + 2 * kAccStages) * 8 + 16;
  static constexpr int SMEM_BYTES = STAGES * STAGE + BAR_BYTES + 1024;
};

template <int BN>
__global__ void __launch_bounds__(kThreads, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap map_dy, const __grid_constant__ CUtensorMap map_x, const WgradDev p) {
  using Cfg = WgradCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + Cfg::STAGES * Cfg::A_STAGE;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::STAGES * Cfg::STAGE);
  uint64_t* empty_bar = full_bar + Cfg::STAGES;
  uint64_t* tfull_bar = empty_bar + Cfg::STAGES;
  uint64_t* tempty_bar = tfull_bar + kAccStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + kAccStages);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  // A patch with fewer than 64 pixels leaves the last rows of a stage untouched; those rows take part in the reduction,
  // so they must be zero (0 x 0) and stay zero: clear all stages once, TMA never writes beyond its box.
  for (int i = threadIdx.x; i < Cfg::STAGES * Cfg::STAGE / 16; i += kThreads)
    reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async();
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_dy);
    tma_prefetch_desc(&map_x);
    for (int i = 0; i < Cfg::STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < kAccStages; ++i) {
      mbar_init(&tfull_bar[i], 1);
      mbar_init(&tempty_bar[i], kEpiWarps);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc<1>(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  const int taps = p.R * p.S;
  const int patches = p.tiles_n * p.tiles_h * p.tiles_w;
  const int per_split = (patches + p.splits - 1) / p.splits;
  const int num_work = p.tiles_k * p.tiles_c * taps * p.splits;
  const int rows_box = p.pn * p.ph * p.pw;
  const uint32_t stage_bytes = static_cast<uint32_t>(rows_box) * 128u * (BLOCK_M / 64 + BN / 64);

  // work item -> (k tile, c tile, tap, split); splits of one tile are neighbours so that they run at the same time
  auto decode = [&](int w, int& kt, int& ct, int& tap, int& lo, int& hi) {
    const int sp = w % p.splits;
    int rest = w / p.splits;
    ct = rest % p.tiles_c; rest /= p.tiles_c;
    tap = rest % taps;
    kt = rest / taps;
    lo = sp * per_split;
    hi = min(patches, lo + per_split);
  };

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int w = blockIdx.x; w < num_work; w += gridDim.x) {
        int kt, ct, tap, lo, hi;
        decode(w, kt, ct, tap, lo, hi);
        const int r = tap / p.S, s = tap - r * p.S;
        for (int pt = lo; pt < hi; ++pt) {
          const int tw = pt % p.tiles_w, th = (pt / p.tiles_w) % p.tiles_h, tn = pt / (p.tiles_w * p.tiles_h);
          const int q0 = tw * p.pw, p0 = th * p.ph, n0 = tn * p.pn;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem_a + stage * Cfg::A_STAGE;
          uint8_t* sb = smem_b + stage * Cfg::B_STAGE;
          mbar_arrive_expect_tx(&full_bar[stage], stage_bytes);
#pragma unroll
          for (int j = 0; j < BLOCK_M / 64; ++j)
            tma_load_4d(&map_dy, &full_bar[stage], sa + j * (64 * BLOCK_K * 2), kt * BLOCK_M + 64 * j, q0, p0, n0);
#pragma unroll
          for (int j = 0; j < BN / 64; ++j)
            tma_load_4d(&map_x, &full_bar[stage], sb + j * (64 * BLOCK_K * 2), ct * BN + 64 * j,
                        q0 * p.stride - p.pad_w + s, p0 * p.stride - p.pad_h + r, n0);
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc = make_idesc_bf16(BLOCK_M, BN, 1, 1);
    const uint32_t lbo = 64 * BLOCK_K * 2, kstep = (UMMA_K * 128) >> 4;
    const uint64_t desc_a0 = make_smem_desc_sw128(smem_u32(smem_a), lbo, 1024);
    const uint64_t desc_b0 = make_smem_desc_sw128(smem_u32(smem_b), lbo, 1024);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int w = blockIdx.x; w < num_work; w += gridDim.x) {
      int kt, ct, tap, lo, hi;
      decode(w, kt, ct, tap, lo, hi);
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * BN;
      for (int pt = lo; pt < hi; ++pt) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t da0 = desc_a0 + static_cast<uint32_t>(stage * (Cfg::A_STAGE >> 4));
          const uint64_t db0 = desc_b0 + static_cast<uint32_t>(stage * (Cfg::B_STAGE >> 4));
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
            umma_bf16<1>(tmem_d, da0 + k * kstep, db0 + k * kstep, idesc, (pt > lo || k > 0) ? 1u : 0u);
          umma_commit<1>(&empty_bar[stage]);
          if (pt == hi - 1) umma_commit<1>(&tfull_bar[acc]);
        }
        __syncwarp();
        if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
      }
      if (lo >= hi) {  // an empty split still owes the epilogue its signal (accumulator content is ignored)
        if (elect_one()) umma_commit<1>(&tfull_bar[acc]);
        __syncwarp();
      }
      if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
    }
  } else {
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = quarter * 32 + lane;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int w = blockIdx.x; w < num_work; w += gridDim.x) {
      int kt, ct, tap, lo, hi;
      decode(w, kt, ct, tap, lo, hi);
      const int k = kt * BLOCK_M + row;
      float* orow = p.dw + (static_cast<size_t>(k) * taps + tap) * p.C;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BN;
      constexpr int PER = BN / 64;
#pragma unroll 1
      for (int c = half * PER; c < (half + 1) * PER; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(taddr + c * 32, v);
        tmem_ld_wait();
        if (c == (half + 1) * PER - 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty_bar[acc]);
        }
        const int col0 = ct * BN + c * 32;
        if (lo >= hi || k >= p.K || col0 >= p.C) continue;
        if (col0 + 32 <= p.C) {
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            red_add_v4_f32(orow + col0 + j, __uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                           __uint_as_float(v[j + 3]));
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (col0 + j < p.C) red_add_f32(orow + col0 + j, __uint_as_float(v[j]));
        }
      }
      if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, Cfg::TMEM_COLS);
  }
}

}  // namespace conv
}  // namespace eb

// ------------------------------------------------------------------ host side
namespace eb {
namespace conv {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  });
  return fn;
}

// bf16 tensor map of rank 3 or 4 over a dense array: dims[0] is the contiguous dimension; 128-byte swizzle, zero fill.
static int make_map_nd(CUtensorMap* map, const void* base, int rank, const uint64_t* dims, const uint32_t* box,
                       const uint32_t* estride) {
  EncodeTiledFn enc = encode_fn();
  if (enc == nullptr) return fail("cuTensorMapEncodeTiled not available from the driver");
  cuuint64_t gdim[4], gstride[3];
  cuuint32_t b[4], e[4];
  uint64_t bytes = 2;
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    b[i] = box[i];
    e[i] = estride[i];
    bytes *= dims[i];
    if (i + 1 < rank) gstride[i] = bytes;
  }
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), gdim, gstride, b, e,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail("cuTensorMapEncodeTiled (rank %d) failed (%d): dims %llu %llu %llu %llu box %u %u %u %u stride %u %u %u %u",
                rank, (int)r, (unsigned long long)dims[0], (unsigned long long)dims[1], (unsigned long long)dims[2],
                (unsigned long long)(rank > 3 ? dims[3] : 0), box[0], box[1], box[2], rank > 3 ? box[3] : 0, estride[0],
                estride[1], estride[2], rank > 3 ? estride[3] : 0);
  return 0;
}

template <typename Kern>
static int prepare(Kern kern, int smem_bytes, bool* done) {
  static std::mutex m;
  std::lock_guard<std::mutex> guard(m);
  int dev = 0;
  EB_CUDA_OK(cudaGetDevice(&dev));
  if (!done[dev & 63]) {
    EB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    done[dev & 63] = true;
  }
  return 0;
}

template <int BN, bool W_MN>
static int launch_igemm(const EbConvArgs& a, IgemmDev p, cudaStream_t stream) {
  using Cfg = IgemmCfg<BN>;
  CUtensorMap map_x, map_w;
  const uint32_t st = static_cast<uint32_t>(a.stride);
  {
    const uint64_t dims[4] = {(uint64_t)a.C, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.N};
    const uint32_t box[4] = {BLOCK_K, (uint32_t)a.pw * st, (uint32_t)a.ph * st, (uint32_t)a.pn};
    const uint32_t es[4] = {1, st, st, 1};
    if (make_map_nd(&map_x, a.in, 4, dims, box, es)) return 1;
  }
  {
    const uint32_t es[3] = {1, 1, 1};
    if (!W_MN) {
      const uint64_t dims[3] = {(uint64_t)a.C, (uint64_t)(a.R * a.S), (uint64_t)a.K};
      const uint32_t box[3] = {BLOCK_K, 1, BN};
      if (make_map_nd(&map_w, a.w, 3, dims, box, es)) return 1;
    } else {
      const uint64_t dims[3] = {(uint64_t)a.K, (uint64_t)(a.R * a.S), (uint64_t)a.C};
      const uint32_t box[3] = {64, 1, BLOCK_K};
      if (make_map_nd(&map_w, a.w, 3, dims, box, es)) return 1;
    }
  }
  p.tiles_k = (a.K + BN - 1) / BN;
  auto kern = conv_igemm_kernel<BN, W_MN>;
  static bool done[64] = {false};
  if (prepare(kern, Cfg::SMEM_BYTES, done)) return 1;
  const long tiles = static_cast<long>(p.tiles_n) * p.tiles_h * p.tiles_w * p.tiles_k;
  const int grid = static_cast<int>(tiles < num_sms() ? tiles : num_sms());
  kern<<<grid, kThreads, Cfg::SMEM_BYTES, stream>>>(map_x, map_w, p);
  EB_LAUNCH_OK("conv_igemm_kernel");
  return 0;
}

template <int BN>
static int launch_wgrad(const EbConvWgradArgs& a, WgradDev p, cudaStream_t stream) {
  using Cfg = WgradCfg<BN>;
  CUtensorMap map_dy, map_x;
  const uint32_t st = static_cast<uint32_t>(a.stride);
  {
    const uint64_t dims[4] = {(uint64_t)a.K, (uint64_t)a.Q, (uint64_t)a.P, (uint64_t)a.N};
    const uint32_t box[4] = {64, (uint32_t)a.pw, (uint32_t)a.ph, (uint32_t)a.pn};
    const uint32_t es[4] = {1, 1, 1, 1};
    if (make_map_nd(&map_dy, a.dy, 4, dims, box, es)) return 1;
  }
  {
    const uint64_t dims[4] = {(uint64_t)a.C, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.N};
    const uint32_t box[4] = {64, (uint32_t)a.pw * st, (uint32_t)a.ph * st, (uint32_t)a.pn};
    const uint32_t es[4] = {1, st, st, 1};
    if (make_map_nd(&map_x, a.x, 4, dims, box, es)) return 1;
  }
  p.tiles_k = (a.K + BLOCK_M - 1) / BLOCK_M;
  p.tiles_c = (a.C + BN - 1) / BN;
  const int patches = p.tiles_n * p.tiles_h * p.tiles_w;
  const int base = p.tiles_k * p.tiles_c * a.R * a.S;
  int splits = a.splits;
  if (splits <= 0) {
    splits = (3 * num_sms() + base - 1) / base;
    const int min_per = 4;  // a split should at least amortise its accumulator round trip
    if (splits > (patches + min_per - 1) / min_per) splits = (patches + min_per - 1) / min_per;
  }
  if (splits > patches) splits = patches;
  if (splits < 1) splits = 1;
  p.splits = splits;
  auto kern = conv_wgrad_kernel<BN>;
  static bool done[64] = {false};
  if (prepare(kern, Cfg::SMEM_BYTES, done)) return 1;
  const long work = static_cast<long>(base) * splits;
  const int grid = static_cast<int>(work < num_sms() ? work : num_sms());
  kern<<<grid, kThreads, Cfg::SMEM_BYTES, stream>>>(map_dy, map_x, p);
  EB_LAUNCH_OK("conv_wgrad_kernel");
  return 0;
}

static int check_patch(const char* who, int pn, int ph, int pw, int limit, int stride) {
  if (pn < 1 || ph < 1 || pw < 1 || pn * ph * pw > limit)
    return fail("%s: patch %dx%dx%d must hold between 1 and %d pixels", who, pn, ph, pw, limit);
  if (pw * stride > 256 || ph * stride > 256 || pn > 256)
    return fail("%s: patch %dx%dx%d with stride %d exceeds the 256-element TMA box limit", who, pn, ph, pw, stride);
  return 0;
}

}  // namespace conv
}  // namespace eb

using namespace eb;

extern "C" int eb_conv_igemm(const EbConvArgs* args, void* stream_) {
  if (args == nullptr) return fail("eb_conv_igemm: null args");
  const EbConvArgs& a = *args;
  if (a.N <= 0 || a.H <= 0 || a.W <= 0 || a.C <= 0 || a.K <= 0 || a.R <= 0 || a.S <= 0 || a.P <= 0 || a.Q <= 0)
    return fail("eb_conv_igemm: empty problem");
  if ((a.C % 8) || (a.K % 8)) return fail("eb_conv_igemm: channel counts must be multiples of 8 (got %d, %d)", a.C, a.K);
  if (a.stride < 1 || a.stride > 8 || a.os < 1) return fail("eb_conv_igemm: stride %d / output stride %d unsupported", a.stride, a.os);
  if ((reinterpret_cast<uintptr_t>(a.in) & 15) || (reinterpret_cast<uintptr_t>(a.w) & 15) || (reinterpret_cast<uintptr_t>(a.out) & 15))
    return fail("eb_conv_igemm: tensors must be 16-byte aligned");
  if (conv::check_patch("eb_conv_igemm", a.pn, a.ph, a.pw, conv::BLOCK_M, a.stride)) return 1;
  if ((a.P - 1) * a.os >= a.OH || (a.Q - 1) * a.os >= a.OW) return fail("eb_conv_igemm: output grid does not fit the output tensor");
  conv::IgemmDev p;
  p.N = a.N; p.P = a.P; p.Q = a.Q; p.C = a.C; p.K = a.K;
  p.R = a.R; p.S = a.S; p.stride = a.stride; p.pad_h = a.pad_h; p.pad_w = a.pad_w; p.flip_taps = a.flip_taps ? 1 : 0;
  p.pn = a.pn; p.ph = a.ph; p.pw = a.pw;
  p.tiles_n = (a.N + a.pn - 1) / a.pn; p.tiles_h = (a.P + a.ph - 1) / a.ph; p.tiles_w = (a.Q + a.pw - 1) / a.pw;
  p.tiles_k = 0;
  p.cblocks = (a.C + conv::BLOCK_K - 1) / conv::BLOCK_K;
  p.out = static_cast<__nv_bfloat16*>(a.out); p.OH = a.OH; p.OW = a.OW; p.os = a.os;
  p.bias = a.bias;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  int bn = a.block_n ? a.block_n : (a.K >= 128 ? 128 : 64);
  if (bn != 64 && bn != 128) return fail("eb_conv_igemm: block_n must be 64 or 128 (got %d)", bn);
  if (a.w_mn) return bn == 128 ? conv::launch_igemm<128, true>(a, p, stream) : conv::launch_igemm<64, true>(a, p, stream);
  return bn == 128 ? conv::launch_igemm<128, false>(a, p, stream) : conv::launch_igemm<64, false>(a, p, stream);
}

extern "C" int eb_conv_wgrad(const EbConvWgradArgs* args, void* stream_) {
  if (args == nullptr) return fail("eb_conv_wgrad: null args");
  const EbConvWgradArgs& a = *args;
  if (a.N <= 0 || a.H <= 0 || a.W <= 0 || a.C <= 0 || a.K <= 0 || a.R <= 0 || a.S <= 0 || a.P <= 0 || a.Q <= 0)
    return fail("eb_conv_wgrad: empty problem");
  if ((a.C % 8) || (a.K % 8)) return fail("eb_conv_wgrad: channel counts must be multiples of 8 (got %d, %d)", a.C, a.K);
  if (a.stride < 1 || a.stride > 8) return fail("eb_conv_wgrad: stride %d unsupported", a.stride);
  if ((reinterpret_cast<uintptr_t>(a.x) & 15) || (reinterpret_cast<uintptr_t>(a.dy) & 15) || (reinterpret_cast<uintptr_t>(a.dw) & 15))
    return fail("eb_conv_wgrad: tensors must be 16-byte aligned");
  if (conv::check_patch("eb_conv_wgrad", a.pn, a.ph, a.pw, conv::BLOCK_K, a.stride)) return 1;
  conv::WgradDev p;
  p.N = a.N; p.P = a.P; p.Q = a.Q; p.C = a.C; p.K = a.K; p.R = a.R; p.S = a.S;
  p.stride = a.stride; p.pad_h = a.pad_h; p.pad_w = a.pad_w;
  p.pn = a.pn; p.ph = a.ph; p.pw = a.pw;
  p.tiles_n = (a.N + a.pn - 1) / a.pn; p.tiles_h = (a.P + a.ph - 1) / a.ph; p.tiles_w = (a.Q + a.pw - 1) / a.pw;
  p.tiles_k = p.tiles_c = p.splits = 0;
  p.dw = a.dw;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  return a.C >= 128 ? conv::launch_wgrad<128>(a, p, stream) : conv::launch_wgrad<64>(a, p, stream);
}
