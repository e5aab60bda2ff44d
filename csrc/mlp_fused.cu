// K8: whole-MLP forward in ONE persistent kernel (predict / evaluate of small Dense chains such as the reference's
// MNIST MLP 784-128-128-10). A CTA owns a tile of 128 rows and carries it through every layer:
//
//   layer 0      A tiles (128 x 64) and W0 tiles stream through a TMA ring, tcgen05.mma accumulates in TMEM
//   epilogue     the four epilogue warps read the accumulator (tcgen05.ld; one thread = one row), add the bias, apply
//                the activation and write the bf16 result straight into shared memory in the 128-byte-swizzled
//                K-major layout the next MMA's A descriptor expects - the activations never touch HBM
//   layers 1..   weights are resident in shared memory (loaded once per CTA), one burst of MMAs per layer
//   last layer   fp32 pre-activations go to global memory; the output activation (softmax / elementwise), loss and
//                metrics are the existing K4 kernel
//
// Limits: 2-4 layers, hidden widths multiples of 16, input width a multiple of 8; up to 128 units per layer in this kernel,
// up to 512 (and an input width up to 512) in the streamed-weights variant further down. Everything else runs the K1 GEMM
// chain. Reference call sites: model.predict / evaluate in the workers
// (elephas/spark_model.py:244-255,284-308; elephas/ml_model.py:199-233).
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <mutex>

#include "../include/elephas_b200.h"
#include "host.h"
#include "math.cuh"
#include "ptx.cuh"
#include "tma_host.h"

namespace eb {

constexpr int kMlpThreads = 192;      // warp 0: TMA, warp 1: MMA + TMEM, warps 2..5: epilogue (one per lane quarter)
constexpr int kMlpTileM = 128;
constexpr int kMlpKBlock = 64;        // bf16 elements per 128-byte swizzled row
constexpr int kMlpUmmaK = 16;
constexpr int kMlpChunkBytesA = kMlpTileM * 128;  // one 64-wide K chunk of a 128-row operand
constexpr int kMlpTmemCols = 128;

struct MlpDev {
  int rows, layers;
  int dims[EB_MLP_MAX_LAYERS + 1];  // dims[0] input width, dims[l + 1] units of layer l
  int npad[EB_MLP_MAX_LAYERS];      // units rounded up to a multiple of 16 (UMMA N)
  const float* bias[EB_MLP_MAX_LAYERS];
  int act[EB_MLP_MAX_LAYERS];
  float* z;
  long ldz;
  int stages;                        // depth of the layer-0 ring
  int ring_off, wres_off[EB_MLP_MAX_LAYERS], act_off, bar_off;  // byte offsets into the (1024-aligned) smem window
  int wres_bytes;                    // total bytes of the resident weights (layers 1..)
  int tiles;
};

}  // namespace eb

namespace eb {

// Bounded wait: a protocol error in this kernel must abort the launch (trap -> cudaErrorLaunchFailure), not hang the GPU.
__device__ __forceinline__ void mlp_wait(uint64_t* bar, uint32_t phase) {
  for (uint32_t spins = 0; !mbar_try_wait(bar, phase); ++spins)
    if (spins > (1u << 22)) asm volatile("trap;");
}

__global__ void __launch_bounds__(kMlpThreads, 1)
mlp_forward_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w0,
                   const __grid_constant__ CUtensorMap map_w1, const __grid_constant__ CUtensorMap map_w2,
                   const __grid_constant__ CUtensorMap map_w3, const MlpDev p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int stage_a = kMlpChunkBytesA;            // A tile of the ring: 128 rows x 128 B
  const int stage_b = p.npad[0] * 128;            // W0 tile: units x 128 B
  const int stage_bytes = stage_a + stage_b;
  uint8_t* ring = smem + p.ring_off;
  uint8_t* act_sm = smem + p.act_off;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + p.bar_off);
  uint64_t* empty_bar = full_bar + 4;
  uint64_t* acc_full = empty_bar + 4;    // MMA -> epilogue: the accumulator of the current layer is complete
  uint64_t* act_ready = acc_full + 1;    // epilogue -> MMA: TMEM drained (and, for hidden layers, A operand written)
  uint64_t* wres_full = act_ready + 1;   // resident weights have landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wres_full + 1);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const CUtensorMap* map_w[EB_MLP_MAX_LAYERS] = {&map_w0, &map_w1, &map_w2, &map_w3};

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_x);
    for (int l = 0; l < p.layers; ++l) tma_prefetch_desc(map_w[l]);
    for (int i = 0; i < 4; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(acc_full, 1);
    mbar_init(act_ready, 4);  // one arrival per epilogue warp
    mbar_init(wres_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc<1>(tmem_slot, kMlpTmemCols);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  const int num_kb = (p.dims[0] + kMlpKBlock - 1) / kMlpKBlock;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      // resident weights of layers 1.. : one 64-wide K chunk (units x 128 B) per TMA box
      mbar_arrive_expect_tx(wres_full, p.wres_bytes);
      for (int l = 1; l < p.layers; ++l) {
        const int chunks = (p.dims[l] + kMlpKBlock - 1) / kMlpKBlock;
        for (int c = 0; c < chunks; ++c)
          tma_load_2d<1>(map_w[l], wres_full, smem + p.wres_off[l] + c * p.npad[l] * 128, c * kMlpKBlock, 0);
      }
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < p.tiles; t += gridDim.x) {
        for (int kb = 0; kb < num_kb; ++kb) {
          mlp_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = ring + stage * stage_bytes;
          mbar_arrive_expect_tx(&full_bar[stage], stage_bytes);
          tma_load_2d<1>(&map_x, &full_bar[stage], sa, kb * kMlpKBlock, t * kMlpTileM);
          tma_load_2d<1>(&map_w0, &full_bar[stage], sa + stage_a, kb * kMlpKBlock, 0);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // -------------------------------------------------------------------- MMA issuer
    int stage = 0;
    uint32_t phase = 0, act_phase = 0;
    bool first_tile = true;
    for (int t = blockIdx.x; t < p.tiles; t += gridDim.x) {
      if (!first_tile) {
        // the previous tile's last epilogue must have drained the accumulator
        mlp_wait(act_ready, act_phase);
        act_phase ^= 1;
        tc_fence_after();
      }
      first_tile = false;
      const uint32_t idesc0 = make_idesc_bf16(kMlpTileM, p.npad[0], 0, 0);
      for (int kb = 0; kb < num_kb; ++kb) {
        mlp_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(ring + stage * stage_bytes);
          const uint64_t da = make_smem_desc_sw128(sa, 16, 1024);
          const uint64_t db = make_smem_desc_sw128(sa + stage_a, 16, 1024);
#pragma unroll
          for (int k = 0; k < kMlpKBlock / kMlpUmmaK; ++k)
            umma_bf16<1>(tmem_base, da + 2 * k, db + 2 * k, idesc0, (kb | k) != 0 ? 1u : 0u);
          umma_commit<1>(&empty_bar[stage]);
          if (kb == num_kb - 1) umma_commit<1>(acc_full);
        }
        __syncwarp();
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
      for (int l = 1; l < p.layers; ++l) {
        if (l == 1) mlp_wait(wres_full, 0);  // (returns at once after the first tile)
        mlp_wait(act_ready, act_phase);      // layer l-1's output is in shared memory, TMEM is free again
        act_phase ^= 1;
        tc_fence_after();
        if (elect_one()) {
          const uint32_t idesc = make_idesc_bf16(kMlpTileM, p.npad[l], 0, 0);
          const int ksteps = p.dims[l] / kMlpUmmaK;
          const uint32_t a0 = smem_u32(act_sm), b0 = smem_u32(smem + p.wres_off[l]);
          for (int k = 0; k < ksteps; ++k) {
            const int chunk = k >> 2, sub = k & 3;
            const uint64_t da = make_smem_desc_sw128(a0 + chunk * kMlpChunkBytesA, 16, 1024) + 2 * sub;
            const uint64_t db = make_smem_desc_sw128(b0 + chunk * p.npad[l] * 128, 16, 1024) + 2 * sub;
            umma_bf16<1>(tmem_base, da, db, idesc, k != 0 ? 1u : 0u);
          }
          umma_commit<1>(acc_full);
        }
        __syncwarp();
      }
    }
  } else {
    // ---------------------------------------------------------------------- epilogue
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
    uint32_t acc_phase = 0;
    for (int t = blockIdx.x; t < p.tiles; t += gridDim.x) {
      const long grow = static_cast<long>(t) * kMlpTileM + row;
      for (int l = 0; l < p.layers; ++l) {
        mlp_wait(acc_full, acc_phase);
        acc_phase ^= 1;
        tc_fence_after();
        const bool last = l == p.layers - 1;
        const int n_units = p.dims[l + 1];
        const float* bias = p.bias[l];
        for (int n0 = 0; n0 < p.npad[l]; n0 += 16) {
          uint32_t v[32];
          tmem_ld_32x16(taddr + n0, v);
          tmem_ld_wait();
          float x[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int n = n0 + j;
            x[j] = __uint_as_float(v[j]) + ((bias != nullptr && n < n_units) ? bias[n] : 0.f);
          }
          if (last) {
            if (grow < p.rows) {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (n0 + j < n_units) p.z[grow * p.ldz + n0 + j] = x[j];
            }
          } else {
            act_apply_chunk<16>(p.act[l], x);
            // two 16-byte groups of the row, at the swizzled position of the K-major SWIZZLE_128B layout
            const int chunk = n0 >> 6, g0 = (n0 & 63) >> 3;
            uint8_t* base = act_sm + chunk * kMlpChunkBytesA + row * 128;
            uint4 o0, o1;
            o0.x = pack_bf16x2(x[0], x[1]); o0.y = pack_bf16x2(x[2], x[3]);
            o0.z = pack_bf16x2(x[4], x[5]); o0.w = pack_bf16x2(x[6], x[7]);
            o1.x = pack_bf16x2(x[8], x[9]); o1.y = pack_bf16x2(x[10], x[11]);
            o1.z = pack_bf16x2(x[12], x[13]); o1.w = pack_bf16x2(x[14], x[15]);
            *reinterpret_cast<uint4*>(base + ((g0 ^ (row & 7)) << 4)) = o0;
            *reinterpret_cast<uint4*>(base + (((g0 + 1) ^ (row & 7)) << 4)) = o1;
          }
        }
        // accumulator read (and the next layer's A operand written): hand over to the MMA warp
        tc_fence_before();
        if (!last) fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(act_ready);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, kMlpTmemCols);
  }
}


// ---------------------------------------------------------------------------------------------------------------------
// Wide variant (hidden widths up to 512, e.g. the Otto model 93-512-512-512-9): the weights no longer fit in shared
// memory, so they STREAM - every layer's weights pass through the TMA ring in 128 x 64 tiles (they come out of L2: the
// whole model is a megabyte) - while the activations stay resident: the input tile is loaded by TMA into the same
// swizzled K-major buffer the epilogue writes the hidden activations into (8 chunks of 128 rows x 64 columns = 128 KB),
// and the accumulator of a layer occupies up to all 512 TMEM columns (N chunks of 128 columns, one burst of MMAs each).
// Same roles and barriers as above, plus in_full / in_free for the input tile.
constexpr int kWideTmemCols = 512;
constexpr int kWideActChunks = 8;
constexpr int kWideTileN = 128;
constexpr int kWideStageBytes = kWideTileN * 128;  // one weight tile: 128 units x 64 inputs (bf16)

struct MlpWideDev {
  int rows, layers;
  int dims[EB_MLP_MAX_LAYERS + 1];
  int npad[EB_MLP_MAX_LAYERS];
  const float* bias[EB_MLP_MAX_LAYERS];
  int act[EB_MLP_MAX_LAYERS];
  float* z;
  long ldz;
  int stages, act_off, ring_off, bar_off, tiles;
};

__global__ void __launch_bounds__(kMlpThreads, 1)
mlp_forward_wide_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w0,
                        const __grid_constant__ CUtensorMap map_w1, const __grid_constant__ CUtensorMap map_w2,
                        const __grid_constant__ CUtensorMap map_w3, const MlpWideDev p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* act_sm = smem + p.act_off;
  uint8_t* ring = smem + p.ring_off;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + p.bar_off);
  uint64_t* empty_bar = full_bar + 8;
  uint64_t* acc_full = empty_bar + 8;    // MMA -> epilogue: the accumulator of the current layer is complete
  uint64_t* act_ready = acc_full + 1;    // epilogue -> MMA: TMEM drained (and, for hidden layers, A operand written)
  uint64_t* in_full = act_ready + 1;     // TMA -> MMA: the input tile has landed in the activation buffer
  uint64_t* in_free = in_full + 1;       // MMA -> TMA: the last layer's MMAs have read the activation buffer
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(in_free + 1);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const CUtensorMap* map_w[EB_MLP_MAX_LAYERS] = {&map_w0, &map_w1, &map_w2, &map_w3};

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_x);
    for (int l = 0; l < p.layers; ++l) tma_prefetch_desc(map_w[l]);
    for (int i = 0; i < 8; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(acc_full, 1);
    mbar_init(act_ready, 4);  // one arrival per epilogue warp
    mbar_init(in_full, 1);
    mbar_init(in_free, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc<1>(tmem_slot, kWideTmemCols);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0, free_phase = 0;
      bool first_tile = true;
      const int in_chunks = (p.dims[0] + kMlpKBlock - 1) / kMlpKBlock;
      for (int t = blockIdx.x; t < p.tiles; t += gridDim.x) {
        if (!first_tile) {
          mlp_wait(in_free, free_phase);  // the previous tile's last layer no longer reads the activation buffer
          free_phase ^= 1;
        }
        first_tile = false;
        mbar_arrive_expect_tx(in_full, in_chunks * kMlpChunkBytesA);
        for (int c = 0; c < in_chunks; ++c)
          tma_load_2d<1>(&map_x, in_full, act_sm + c * kMlpChunkBytesA, c * kMlpKBlock, t * kMlpTileM);
        for (int l = 0; l < p.layers; ++l) {
          const int kcs = (p.dims[l] + kMlpKBlock - 1) / kMlpKBlock;
          const int ncs = (p.npad[l] + kWideTileN - 1) / kWideTileN;
          for (int nc = 0; nc < ncs; ++nc)
            for (int kb = 0; kb < kcs; ++kb) {
              mlp_wait(&empty_bar[stage], phase ^ 1);
              mbar_arrive_expect_tx(&full_bar[stage], kWideStageBytes);
              tma_load_2d<1>(map_w[l], &full_bar[stage], ring + stage * kWideStageBytes, kb * kMlpKBlock, nc * kWideTileN);
              if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
        }
      }
    }
  } else if (warp == 1) {
    // -------------------------------------------------------------------- MMA issuer
    int stage = 0;
    uint32_t phase = 0, act_phase = 0, in_phase = 0;
    bool first_tile = true;
    for (int t = blockIdx.x; t < p.tiles; t += gridDim.x) {
      mlp_wait(in_full, in_phase);
      in_phase ^= 1;
      if (!first_tile) {
        mlp_wait(act_ready, act_phase);  // the previous tile's last epilogue has drained the accumulator
        act_phase ^= 1;
      }
      first_tile = false;
      tc_fence_after();
      for (int l = 0; l < p.layers; ++l) {
        if (l > 0) {
          mlp_wait(act_ready, act_phase);  // layer l-1's output is in the activation buffer, TMEM is free again
          act_phase ^= 1;
          tc_fence_after();
        }
        const int kcs = (p.dims[l] + kMlpKBlock - 1) / kMlpKBlock;
        const int ncs = (p.npad[l] + kWideTileN - 1) / kWideTileN;
        const bool last = l == p.layers - 1;
        for (int nc = 0; nc < ncs; ++nc) {
          const int ncols = min(kWideTileN, p.npad[l] - nc * kWideTileN);
          const uint32_t idesc = make_idesc_bf16(kMlpTileM, ncols, 0, 0);
          for (int kb = 0; kb < kcs; ++kb) {
            mlp_wait(&full_bar[stage], phase);
            tc_fence_after();
            if (elect_one()) {
              const uint64_t da = make_smem_desc_sw128(smem_u32(act_sm + kb * kMlpChunkBytesA), 16, 1024);
              const uint64_t db = make_smem_desc_sw128(smem_u32(ring + stage * kWideStageBytes), 16, 1024);
#pragma unroll
              for (int k = 0; k < kMlpKBlock / kMlpUmmaK; ++k)
                umma_bf16<1>(tmem_base + nc * kWideTileN, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
              umma_commit<1>(&empty_bar[stage]);
              if (nc == ncs - 1 && kb == kcs - 1) {
                umma_commit<1>(acc_full);
                if (last) umma_commit<1>(in_free);
              }
            }
            __syncwarp();
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else {
    // ---------------------------------------------------------------------- epilogue
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
    uint32_t acc_phase = 0;
    for (int t = blockIdx.x; t < p.tiles; t += gridDim.x) {
      const long grow = static_cast<long>(t) * kMlpTileM + row;
      for (int l = 0; l < p.layers; ++l) {
        mlp_wait(acc_full, acc_phase);
        acc_phase ^= 1;
        tc_fence_after();
        const bool last = l == p.layers - 1;
        const int n_units = p.dims[l + 1];
        const float* bias = p.bias[l];
        for (int n0 = 0; n0 < p.npad[l]; n0 += 16) {
          uint32_t v[32];
          tmem_ld_32x16(taddr + n0, v);
          tmem_ld_wait();
          float x[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int n = n0 + j;
            x[j] = __uint_as_float(v[j]) + ((bias != nullptr && n < n_units) ? bias[n] : 0.f);
          }
          if (last) {
            if (grow < p.rows) {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (n0 + j < n_units) p.z[grow * p.ldz + n0 + j] = x[j];
            }
          } else {
            act_apply_chunk<16>(p.act[l], x);
            const int chunk = n0 >> 6, g0 = (n0 & 63) >> 3;
            uint8_t* base = act_sm + chunk * kMlpChunkBytesA + row * 128;
            uint4 o0, o1;
            o0.x = pack_bf16x2(x[0], x[1]); o0.y = pack_bf16x2(x[2], x[3]);
            o0.z = pack_bf16x2(x[4], x[5]); o0.w = pack_bf16x2(x[6], x[7]);
            o1.x = pack_bf16x2(x[8], x[9]); o1.y = pack_bf16x2(x[10], x[11]);
            o1.z = pack_bf16x2(x[12], x[13]); o1.w = pack_bf16x2(x[14], x[15]);
            *reinterpret_cast<uint4*>(base + ((g0 ^ (row & 7)) << 4)) = o0;
            *reinterpret_cast<uint4*>(base + (((g0 + 1) ^ (row & 7)) << 4)) = o1;
          }
        }
        if (!last) {
          // the rest of the last 64-wide chunk is read by the next layer's MMAs (against zero weights): no stale NaNs there
          const int upto = (p.npad[l] + 63) / 64 * 64;
          for (int n0 = p.npad[l]; n0 < upto; n0 += 8) {
            uint8_t* base = act_sm + (n0 >> 6) * kMlpChunkBytesA + row * 128;
            *reinterpret_cast<uint4*>(base + ((((n0 & 63) >> 3) ^ (row & 7)) << 4)) = make_uint4(0u, 0u, 0u, 0u);
          }
        }
        tc_fence_before();
        if (!last) fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(act_ready);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, kWideTmemCols);
  }
}

}  // namespace eb

using namespace eb;

// Layers wider than 128 units: the streamed-weights kernel (input width and every layer up to 512).
static int mlp_forward_wide(const EbMlpArgs* a, void* stream) {
  if (a->dims[0] > kWideActChunks * kMlpKBlock)
    return fail("eb_mlp_forward: input width %d exceeds %d (wide layers keep the input tile on chip)", a->dims[0],
                kWideActChunks * kMlpKBlock);
  MlpWideDev p;
  p.rows = a->rows;
  p.layers = a->layers;
  p.dims[0] = a->dims[0];
  for (int l = 0; l < a->layers; ++l) {
    const int units = a->dims[l + 1];
    const bool last = l == a->layers - 1;
    if (units <= 0 || units > kWideTmemCols) return fail("eb_mlp_forward: layer %d has %d units (1..%d supported)", l, units, kWideTmemCols);
    if (!last && (units % 16)) return fail("eb_mlp_forward: hidden width %d is not a multiple of 16", units);
    if (a->w[l] == nullptr || (a->ldw[l] % 8) || (reinterpret_cast<uintptr_t>(a->w[l]) & 15) || a->ldw[l] < a->dims[l])
      return fail("eb_mlp_forward: weights of layer %d must be 16-byte aligned with ldw %% 8 == 0 and ldw >= width", l);
    p.dims[l + 1] = units;
    p.npad[l] = (units + 15) / 16 * 16;
    p.bias[l] = a->bias[l];
    p.act[l] = a->act[l];
  }
  for (int l = a->layers; l < EB_MLP_MAX_LAYERS; ++l) {
    p.dims[l + 1] = 0; p.npad[l] = 16; p.bias[l] = nullptr; p.act[l] = 0;
  }
  p.z = a->z;
  p.ldz = a->ldz;
  p.tiles = (a->rows + kMlpTileM - 1) / kMlpTileM;
  // shared-memory plan: [activation buffer: 8 chunks of 128 rows x 128 B][weight ring][barriers]
  p.act_off = 0;
  p.ring_off = kWideActChunks * kMlpChunkBytesA;
  p.stages = (227 * 1024 - 1024 - 256 - p.ring_off) / kWideStageBytes;
  if (p.stages > 6) p.stages = 6;
  p.bar_off = p.ring_off + p.stages * kWideStageBytes;
  const int smem_bytes = p.bar_off + 256 + 1024;

  CUtensorMap mx, mw[EB_MLP_MAX_LAYERS];
  if (make_map(&mx, a->x, a->rows, a->dims[0], a->ldx, kMlpTileM, kMlpKBlock)) return 1;
  for (int l = 0; l < EB_MLP_MAX_LAYERS; ++l) {
    if (l < a->layers) {
      if (make_map(&mw[l], a->w[l], p.dims[l + 1], p.dims[l], a->ldw[l], kWideTileN, kMlpKBlock)) return 1;
    } else {
      mw[l] = mw[0];
    }
  }
  {
    static std::mutex init_mutex;
    static bool attr_set[64] = {false};
    int dev = 0;
    EB_CUDA_OK(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> guard(init_mutex);
    if (!attr_set[dev & 63]) {
      EB_CUDA_OK(cudaFuncSetAttribute(mlp_forward_wide_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
      attr_set[dev & 63] = true;
    }
  }
  int grid = num_sms();
  if (grid > p.tiles) grid = p.tiles;
  mlp_forward_wide_kernel<<<grid, kMlpThreads, smem_bytes, static_cast<cudaStream_t>(stream)>>>(mx, mw[0], mw[1], mw[2],
                                                                                                mw[3], p);
  EB_LAUNCH_OK("mlp_forward_wide_kernel");
  return 0;
}

extern "C" int eb_mlp_forward(const EbMlpArgs* a, void* stream) {
  if (a == nullptr) return fail("eb_mlp_forward: null args");
  if (a->rows <= 0) return 0;
  if (a->layers < 2 || a->layers > EB_MLP_MAX_LAYERS)
    return fail("eb_mlp_forward: 2..%d layers supported (got %d)", EB_MLP_MAX_LAYERS, a->layers);
  if (a->x == nullptr || a->z == nullptr) return fail("eb_mlp_forward: input / output missing");
  if (a->dims[0] <= 0 || (a->dims[0] % 8) || (a->ldx % 8) || (reinterpret_cast<uintptr_t>(a->x) & 15))
    return fail("eb_mlp_forward: input width and ldx must be multiples of 8, x 16-byte aligned");
  bool wide = false;
  for (int l = 0; l < a->layers; ++l) wide |= a->dims[l + 1] > 128;
  if (wide) return mlp_forward_wide(a, stream);
  MlpDev p;
  p.rows = a->rows;
  p.layers = a->layers;
  p.dims[0] = a->dims[0];
  for (int l = 0; l < a->layers; ++l) {
    const int units = a->dims[l + 1];
    const bool last = l == a->layers - 1;
    if (units <= 0 || units > 128) return fail("eb_mlp_forward: layer %d has %d units (1..128 supported)", l, units);
    if (!last && (units % 16)) return fail("eb_mlp_forward: hidden width %d is not a multiple of 16", units);
    if (a->w[l] == nullptr || (a->ldw[l] % 8) || (reinterpret_cast<uintptr_t>(a->w[l]) & 15) || a->ldw[l] < a->dims[l])
      return fail("eb_mlp_forward: weights of layer %d must be 16-byte aligned with ldw %% 8 == 0 and ldw >= width", l);
    p.dims[l + 1] = units;
    p.npad[l] = (units + 15) / 16 * 16;
    p.bias[l] = a->bias[l];
    p.act[l] = a->act[l];
  }
  for (int l = a->layers; l < EB_MLP_MAX_LAYERS; ++l) {
    p.dims[l + 1] = 0; p.npad[l] = 16; p.bias[l] = nullptr; p.act[l] = 0;
  }
  p.z = a->z;
  p.ldz = a->ldz;
  p.tiles = (a->rows + kMlpTileM - 1) / kMlpTileM;

  // shared-memory plan: [resident weights of layers 1..][activation buffer][ring][barriers]
  int off = 0;
  p.wres_bytes = 0;
  for (int l = 0; l < EB_MLP_MAX_LAYERS; ++l) p.wres_off[l] = 0;
  for (int l = 1; l < a->layers; ++l) {
    const int chunks = (p.dims[l] + kMlpKBlock - 1) / kMlpKBlock;
    p.wres_off[l] = off;
    off += chunks * p.npad[l] * 128;
    off = (off + 1023) / 1024 * 1024;
    p.wres_bytes += chunks * p.npad[l] * 128;
  }
  p.act_off = off;
  off += 2 * kMlpChunkBytesA;  // hidden widths up to 128: two 64-wide chunks
  p.ring_off = off;
  const int stage_bytes = kMlpChunkBytesA + p.npad[0] * 128;
  const int budget = 227 * 1024 - 1024 /* alignment slack */ - 256 /* barriers */ - off;
  int stages = budget / stage_bytes;
  if (stages > 4) stages = 4;
  if (stages < 2) return fail("eb_mlp_forward: layers too wide for the shared-memory plan");
  p.stages = stages;
  off += stages * stage_bytes;
  p.bar_off = off;
  const int smem_bytes = off + 256 + 1024;

  CUtensorMap mx, mw[EB_MLP_MAX_LAYERS];
  if (make_map(&mx, a->x, a->rows, a->dims[0], a->ldx, kMlpTileM, kMlpKBlock)) return 1;
  for (int l = 0; l < EB_MLP_MAX_LAYERS; ++l) {
    if (l < a->layers) {
      if (make_map(&mw[l], a->w[l], p.dims[l + 1], p.dims[l], a->ldw[l], p.npad[l], kMlpKBlock)) return 1;
    } else {
      mw[l] = mw[0];
    }
  }
  {
    static std::mutex init_mutex;
    static bool attr_set[64] = {false};
    int dev = 0;
    EB_CUDA_OK(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> guard(init_mutex);
    if (!attr_set[dev & 63]) {
      EB_CUDA_OK(cudaFuncSetAttribute(mlp_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
      attr_set[dev & 63] = true;
    }
  }
  int grid = num_sms();
  if (grid > p.tiles) grid = p.tiles;
  mlp_forward_kernel<<<grid, kMlpThreads, smem_bytes, static_cast<cudaStream_t>(stream)>>>(mx, mw[0], mw[1], mw[2],
                                                                                           mw[3], p);
  EB_LAUNCH_OK("mlp_forward_kernel");
  return 0;
}
