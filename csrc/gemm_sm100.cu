// K1: Dense GEMM for sm_100a. C[M,N] = epilogue(A x B^T), bf16 in, fp32 accumulate.
//
// Design (B200-first, no library GEMM):
//   * persistent, warp-specialised CTA of 192 threads:
//       warp 0      TMA producer   (cp.async.bulk.tensor, 128B swizzle, OOB zero fill handles all tails)
//       warp 1      MMA issuer     (one lane issues tcgen05.mma; owns the TMEM allocation)
//       warps 2..5  epilogue       (tcgen05.ld -> bias/activation/dropout or activation-gradient -> global)
//   * three mbarrier pipelines: smem full/empty ring (TMA <-> MMA), TMEM full/empty
//     double buffer (MMA <-> epilogue) so the epilogue of tile i overlaps the MMAs of tile i+1
//   * optional CTA pair (cta_group::2, UMMA M=256): each CTA loads its own 128 rows of A and
//     half of the B tile; the leader CTA issues the MMAs for both and multicasts the commits
//   * both operands may be K-major or MN-major (shared-memory descriptors differ, data is
//     never transposed in memory), which covers forward, dgrad and wgrad of a Dense layer.
//
// Absorbs the Keras Dense forward/backward the reference reaches through
// elephas/worker.py:42,108,123 and elephas/spark_model.py:248,255,291.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <stdlib.h>

#include <mutex>

#include "../include/elephas_b200.h"
#include "host.h"
#include "math.cuh"
#include "ptx.cuh"

namespace eb {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // 64 bf16 = one 128-byte swizzle atom
constexpr int UMMA_K = 16;
constexpr int kEpiWarps = 8;                       // two warps per TMEM lane quarter, each takes half the columns
constexpr int kGemmThreads = 64 + 32 * kEpiWarps;  // warp 0: TMA, warp 1: MMA, warps 2..9: epilogue
constexpr int kAccStages = 2;
constexpr int kGroupM = 8;  // tile rasterisation: 8 m-tiles share each B tile through L2

struct GemmDev {
  int M, N, K;
  int a_mn, b_mn;
  float alpha;
  int epi, act;
  const float* bias;
  int bias_vec;  // bias pointer is 16-byte aligned: float4 loads
  const __nv_bfloat16* aux;
  int ldaux;
  float drop_p;
  uint32_t drop_thresh;
  float drop_scale;
  uint64_t seed, drop_stream;
  const long long* step_ptr;
  __nv_bfloat16* out_bf16;
  int ldo;
  float* out_f32;
  int ldo32;
  float beta;
  int bf16_from_f32;
  int tiles_m, tiles_n;
  int b_copy_on;  // side-write every landed B tile through map_c (K2: the GEMM is the weight pull)
  int full_work, num_work;  // work items: [0, full_work) whole tiles, then the tail tiles as two half-N items each
};

template <int BN, int CG>
struct GemmCfg {
  static constexpr int B_ROWS = BN / CG;
  static constexpr int A_STAGE = BLOCK_M * BLOCK_K * 2;
  static constexpr int B_STAGE = B_ROWS * BLOCK_K * 2;
  static constexpr int STAGE = A_STAGE + B_STAGE;
  static constexpr int STAGES_RAW = (225 * 1024) / STAGE;  // 227 KB per CTA minus barriers and alignment slack
  static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
  static constexpr int TMEM_COLS = (kAccStages * BN <= 32)    ? 32
                                   : (kAccStages * BN <= 64)  ? 64
                                   : (kAccStages * BN <= 128) ? 128
                                   : (kAccStages * BN <= 256) ? 256
                                                              : 512;
  static constexpr int BAR_BYTES = (2 * STAGES + 2 * kAccStages) * 8 + 16;
  static constexpr int SMEM_BYTES = STAGES * STAGE + BAR_BYTES + 1024;  // +1024: manual alignment slack
  static_assert(BN % 16 == 0 && BN >= 16 && BN <= 256, "UMMA N constraint");
  static_assert(CG == 1 || B_ROWS % 64 == 0, "CTA pair needs >= 64 B rows per CTA");
};

__device__ __forceinline__ void tile_coords(int t, int tiles_m, int tiles_n, int& tm, int& tn) {
  const int per_group = kGroupM * tiles_n;
  const int group = t / per_group;
  const int first_m = group * kGroupM;
  const int gsize = min(tiles_m - first_m, kGroupM);
  const int r = t - group * per_group;
  tm = first_m + r % gsize;
  tn = r / gsize;
}

// Work item -> tile coordinates. The last, partially filled wave of a persistent grid leaves most SMs idle;
// when at most half of the CTA groups would get a tile there, each tail tile is issued as two half-width
// (BN/2) work items instead, so the tail wave is shorter and busier.
struct WorkItem {
  int tm, tn, n_off, width;
};

template <int BN>
__device__ __forceinline__ WorkItem decode_work(int w, const GemmDev& p) {
  WorkItem it;
  int tile = w;
  it.n_off = 0;
  it.width = BN;
  if (w >= p.full_work) {
    const int s = w - p.full_work;
    tile = p.full_work + (s >> 1);
    it.n_off = (s & 1) * (BN / 2);
    it.width = BN / 2;
  }
  tile_coords(tile, p.tiles_m, p.tiles_n, it.tm, it.tn);
  return it;
}

template <int BN, int CG, int MC>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 const __grid_constant__ CUtensorMap map_c, const GemmDev p) {
  using Cfg = GemmCfg<BN, CG>;
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B needs 1024-byte aligned tiles; the dynamic smem base is only 16-byte aligned by contract.
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + Cfg::STAGES * Cfg::A_STAGE;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::STAGES * Cfg::STAGE);
  uint64_t* empty_bar = full_bar + Cfg::STAGES;
  uint64_t* tfull_bar = empty_bar + Cfg::STAGES;
  uint64_t* tempty_bar = tfull_bar + kAccStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + kAccStages);

  // warp index through a shuffle: provably warp-uniform, so role dispatch compiles to uniform branches
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  // cluster = MC CTA groups stacked along M that share (multicast) the B tile; cta_rank is the rank inside the group
  const uint32_t cluster_rank = (CG * MC > 1) ? cluster_ctarank() : 0u;
  const uint32_t cta_rank = cluster_rank & (CG - 1);
  const uint32_t group = cluster_rank / CG;
  const bool leader = cta_rank == 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    for (int i = 0; i < Cfg::STAGES; ++i) {
      mbar_init(&full_bar[i], CG);  // leader: arrive.expect_tx; peer CTA: remote arrive
      mbar_init(&empty_bar[i], MC);  // one tcgen05.commit per CTA group that reads this CTA's B rows
    }
    for (int i = 0; i < kAccStages; ++i) {
      mbar_init(&tfull_bar[i], 1);        // one tcgen05.commit
      mbar_init(&tempty_bar[i], CG * kEpiWarps);  // one arrival per epilogue warp of every CTA in the group
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc<CG>(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish<CG>();
  }
  tc_fence_before();
  if constexpr (CG * MC > 1) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  const int num_tiles = p.num_work;  // work items (whole tiles and half-width tail items)
  const int num_kb = (p.K + BLOCK_K - 1) / BLOCK_K;
  const int first_tile = blockIdx.x / (CG * MC);
  const int tile_step = gridDim.x / (CG * MC);

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      // K2: B tiles of the first m-tile row are written back to the local copy one pipeline step after their
      // load was issued (by then they have usually landed, so the producer rarely stalls)
      int pend_stage = -1, pend_k0 = 0, pend_n0 = 0;
      uint32_t pend_phase = 0;
      for (int t = first_tile; t < num_tiles; t += tile_step) {
        const WorkItem it = decode_work<BN>(t, p);
        const bool copy_tile = (CG * MC == 1) && p.b_copy_on && it.tm == 0;
        const int m0 = ((it.tm * MC + group) * CG + cta_rank) * BLOCK_M;
        // a half-width item still loads full-height B boxes (constant transaction size); only width/CG rows are used
        const int n0 = it.tn * BN + it.n_off + cta_rank * (it.width / CG);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if constexpr (CG * MC == 1) {
            if (p.b_copy_on) tma_store_wait_read();  // an earlier side-write may still be reading this slot
          }
          const int k0 = kb * BLOCK_K;
          uint8_t* sa = smem_a + stage * Cfg::A_STAGE;
          uint8_t* sb = smem_b + stage * Cfg::B_STAGE;
          if (CG == 1 || leader)
            mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE * CG);
          else
            mbar_arrive_leader(&full_bar[stage]);
          if (!p.a_mn) {
            tma_load_2d<CG>(&map_a, &full_bar[stage], sa, k0, m0);
          } else {
#pragma unroll
            for (int j = 0; j < BLOCK_M / 64; ++j)
              tma_load_2d<CG>(&map_a, &full_bar[stage], sa + j * (64 * BLOCK_K * 2), m0 + 64 * j, k0);
          }
          if constexpr (MC == 1) {
            if (!p.b_mn) {
              tma_load_2d<CG>(&map_b, &full_bar[stage], sb, k0, n0);
            } else {
              if constexpr (Cfg::B_ROWS % 64 == 0) {
#pragma unroll
                for (int j = 0; j < Cfg::B_ROWS / 64; ++j)
                  tma_load_2d<CG>(&map_b, &full_bar[stage], sb + j * (64 * BLOCK_K * 2), n0 + 64 * j, k0);
              }
            }
          } else {
            // Both CTA groups of the cluster need the same B rows: each loads half of them (64 rows) and
            // multicasts the box to its counterpart (same in-group rank) in the other group.
            const uint16_t mask = static_cast<uint16_t>((1u << cta_rank) | (1u << (cta_rank + CG)));
            uint8_t* dst = sb + group * (64 * BLOCK_K * 2);
            if (!p.b_mn) tma_load_2d_mcast_pair(&map_b, &full_bar[stage], dst, k0, n0 + 64 * group, mask);
            else tma_load_2d_mcast_pair(&map_b, &full_bar[stage], dst, n0 + 64 * group, k0, mask);
          }
          if constexpr (CG * MC == 1) {
            if (pend_stage >= 0) {
              mbar_wait(&full_bar[pend_stage], pend_phase);
              tma_store_2d(&map_c, smem_b + pend_stage * Cfg::B_STAGE, pend_k0, pend_n0);
              tma_store_commit();
              pend_stage = -1;
            }
            if (copy_tile) { pend_stage = stage; pend_phase = phase; pend_k0 = k0; pend_n0 = n0; }
          }
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
        }
      }
      if constexpr (CG * MC == 1) {
        if (pend_stage >= 0) {
          mbar_wait(&full_bar[pend_stage], pend_phase);
          tma_store_2d(&map_c, smem_b + pend_stage * Cfg::B_STAGE, pend_k0, pend_n0);
          tma_store_commit();
        }
        if (p.b_copy_on) tma_store_wait();
      }
    }
  } else if (warp == 1) {
    // -------------------------------------------------------------- MMA issuer
    if (leader) {
      const uint32_t idesc_full = make_idesc_bf16(BLOCK_M * CG, BN, p.a_mn, p.b_mn);
      const uint32_t idesc_half = make_idesc_bf16(BLOCK_M * CG, BN / 2, p.a_mn, p.b_mn);
      // K-major:  8-row groups are 1024 B apart (SBO); LBO is unused with swizzling (canonical value 16 B).
      // MN-major: 8 k-rows of one 64-wide block are 1024 B apart (SBO); 64-wide blocks are 8192 B apart (LBO).
      const uint32_t a_lbo = p.a_mn ? 64 * BLOCK_K * 2 : 16, b_lbo = p.b_mn ? 64 * BLOCK_K * 2 : 16;
      const uint32_t a_kstep = (p.a_mn ? UMMA_K * 128 : UMMA_K * 2) >> 4, b_kstep = (p.b_mn ? UMMA_K * 128 : UMMA_K * 2) >> 4;
      // Descriptors differ between stages / k-steps only in the 14-bit start-address field (smem < 256 KB, so
      // adding to the low word never carries out of it): build them once, then add 16-byte-unit offsets.
      const uint64_t desc_a0 = make_smem_desc_sw128(smem_u32(smem_a), a_lbo, 1024);
      const uint64_t desc_b0 = make_smem_desc_sw128(smem_u32(smem_b), b_lbo, 1024);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = first_tile; t < num_tiles; t += tile_step) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        const uint32_t idesc = t >= p.full_work ? idesc_half : idesc_full;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t da0 = desc_a0 + static_cast<uint32_t>(stage * (Cfg::A_STAGE >> 4));
            const uint64_t db0 = desc_b0 + static_cast<uint32_t>(stage * (Cfg::B_STAGE >> 4));
#pragma unroll
            for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
              umma_bf16<CG>(tmem_d, da0 + k * a_kstep, db0 + k * b_kstep, idesc, (kb | k) != 0 ? 1u : 0u);
            // smem slot reusable once these MMAs retire: tell every CTA that writes into it
            umma_commit<CG>(&empty_bar[stage], static_cast<uint16_t>((1u << (CG * MC)) - 1));
            // accumulator complete: tell the epilogue warps of this CTA group
            if (kb == num_kb - 1) umma_commit<CG>(&tfull_bar[acc], static_cast<uint16_t>(0x3u << (CG * group)));
          }
          __syncwarp();
          if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
        }
        if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ---------------------------------------------------------------- epilogue
    const int quarter = warp & 3;          // TMEM lane quarter this warp may access (hardware: warp id % 4)
    const int half = (warp - 2) >> 2;      // which half of the tile's column chunks this warp handles
    const int row = quarter * 32 + lane;
    int acc = 0;
    uint32_t acc_phase = 0;
    const bool has_drop = p.drop_p > 0.f;
    // the Philox stream is offset by the device-side step counter so that a CUDA-graph replay draws fresh masks
    const uint64_t drop_stream = p.drop_stream + (p.step_ptr ? static_cast<uint64_t>(*p.step_ptr) * 1024ull : 0ull);
    for (int t = first_tile; t < num_tiles; t += tile_step) {
      const WorkItem it = decode_work<BN>(t, p);
      const int m = ((it.tm * MC + group) * CG + cta_rank) * BLOCK_M + row;
      const int n_tile = it.tn * BN + it.n_off;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BN;
      constexpr int CHUNK = BN >= 32 ? 32 : 16;
      const int nchunk = it.width / CHUNK > 0 ? it.width / CHUNK : 1;
      // chunks [c_lo, c_hi) belong to this warp; with a single chunk the second warp of the quarter only
      // takes part in the barrier protocol
      const int per = (nchunk + 1) / 2;
      const int c_lo = half * per, c_hi = min(nchunk, c_lo + per);
      if (c_lo >= c_hi) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if constexpr (CG == 1) mbar_arrive(&tempty_bar[acc]); else mbar_arrive_leader(&tempty_bar[acc]);
        }
      }
#pragma unroll 1
      for (int c = c_lo; c < c_hi; ++c) {
        uint32_t v[32];
        if constexpr (CHUNK == 32) tmem_ld_32x32(taddr + c * 32, v); else tmem_ld_32x16(taddr + c * 16, v);
        tmem_ld_wait();
        if (c == c_hi - 1) {
          // all of this warp's accumulator reads are complete: hand the TMEM stage back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if constexpr (CG == 1) mbar_arrive(&tempty_bar[acc]); else mbar_arrive_leader(&tempty_bar[acc]);
          }
        }
        const int n0 = n_tile + c * CHUNK;
        if (m >= p.M || n0 >= p.N) continue;
        float x[CHUNK];
#pragma unroll
        for (int j = 0; j < CHUNK; ++j) x[j] = p.alpha * __uint_as_float(v[j]);
        const bool full = (n0 + CHUNK <= p.N);
        if (p.epi == EB_EPI_FWD) {
          if (p.bias != nullptr) {
            if (full && p.bias_vec) {
              const float4* b4 = reinterpret_cast<const float4*>(p.bias + n0);
#pragma unroll
              for (int j = 0; j < CHUNK; j += 4) {
                const float4 b = __ldg(b4 + (j >> 2));
                x[j] += b.x; x[j + 1] += b.y; x[j + 2] += b.z; x[j + 3] += b.w;
              }
            } else {
#pragma unroll
              for (int j = 0; j < CHUNK; ++j)
                if (full || n0 + j < p.N) x[j] += __ldg(p.bias + n0 + j);
            }
          }
          if (p.act != ACT_LINEAR) act_apply_chunk<CHUNK>(p.act, x);
          if (has_drop) {
#pragma unroll
            for (int j = 0; j < CHUNK; j += 4) {
              bool keep[4];
              dropout_keep4(p.seed, drop_stream, static_cast<uint64_t>(m) * p.N + n0 + j, p.drop_thresh, keep);
#pragma unroll
              for (int e = 0; e < 4; ++e) x[j + e] = keep[e] ? x[j + e] * p.drop_scale : 0.f;
            }
          }
        } else {  // EB_EPI_ACTGRAD
          const __nv_bfloat16* arow = p.aux + static_cast<size_t>(m) * p.ldaux + n0;
          float y[CHUNK];
          if (full) {
#pragma unroll
            for (int j = 0; j < CHUNK; j += 8) {
              uint4 q = ld_nc_v4(arow + j);
              y[j + 0] = bf16_lo(q.x); y[j + 1] = bf16_hi(q.x);
              y[j + 2] = bf16_lo(q.y); y[j + 3] = bf16_hi(q.y);
              y[j + 4] = bf16_lo(q.z); y[j + 5] = bf16_hi(q.z);
              y[j + 6] = bf16_lo(q.w); y[j + 7] = bf16_hi(q.w);
            }
          } else {
#pragma unroll
            for (int j = 0; j < CHUNK; ++j) y[j] = (n0 + j < p.N) ? __bfloat162float(arow[j]) : 0.f;
          }
          if (has_drop || p.act != ACT_LINEAR) {
            float g[CHUNK];
            act_grad_chunk<CHUNK>(p.act, y, has_drop ? 1.f - p.drop_p : 1.f, g);
            if (has_drop) {
#pragma unroll
              for (int j = 0; j < CHUNK; j += 4) {
                bool keep[4];
                dropout_keep4(p.seed, drop_stream, static_cast<uint64_t>(m) * p.N + n0 + j, p.drop_thresh, keep);
#pragma unroll
                for (int e = 0; e < 4; ++e) x[j + e] = keep[e] ? x[j + e] * p.drop_scale * g[j + e] : 0.f;
              }
            } else {
#pragma unroll
              for (int j = 0; j < CHUNK; ++j) x[j] *= g[j];
            }
          }
        }
        if (p.out_f32 != nullptr) {
          float* orow = p.out_f32 + static_cast<size_t>(m) * p.ldo32 + n0;
          if (full) {
#pragma unroll
            for (int j = 0; j < CHUNK; j += 4) {
              float4 o = make_float4(x[j], x[j + 1], x[j + 2], x[j + 3]);
              if (p.beta != 0.f) {
                const float4 old = *reinterpret_cast<const float4*>(orow + j);
                o.x += p.beta * old.x; o.y += p.beta * old.y; o.z += p.beta * old.z; o.w += p.beta * old.w;
              }
              *reinterpret_cast<float4*>(orow + j) = o;
              if (p.bf16_from_f32) { x[j] = o.x; x[j + 1] = o.y; x[j + 2] = o.z; x[j + 3] = o.w; }
            }
          } else {
#pragma unroll
            for (int j = 0; j < CHUNK; ++j)
              if (n0 + j < p.N) {
                const float o = (p.beta != 0.f ? p.beta * orow[j] : 0.f) + x[j];
                orow[j] = o;
                if (p.bf16_from_f32) x[j] = o;
              }
          }
        }
        if (p.out_bf16 != nullptr) {
          __nv_bfloat16* orow = p.out_bf16 + static_cast<size_t>(m) * p.ldo + n0;
          if (full) {
#pragma unroll
            for (int j = 0; j < CHUNK; j += 8) {
              uint4 q;
              q.x = pack_bf16x2(x[j + 0], x[j + 1]);
              q.y = pack_bf16x2(x[j + 2], x[j + 3]);
              q.z = pack_bf16x2(x[j + 4], x[j + 5]);
              q.w = pack_bf16x2(x[j + 6], x[j + 7]);
              st_v4(orow + j, q);
            }
          } else {
#pragma unroll
            for (int j = 0; j < CHUNK; ++j)
              if (n0 + j < p.N) orow[j] = __float2bfloat16_rn(x[j]);
          }
        }
      }
      if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
    }
  }

  // ------------------------------------------------------------------ teardown
  tc_fence_before();
  if constexpr (CG * MC > 1) cluster_sync_all(); else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<CG>(tmem_base, Cfg::TMEM_COLS);
  }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
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

// 2-D bf16 tensor map over a row-major [rows][ld] array whose logical extent is [rows][cols];
// box = [box_rows][box_cols], 128-byte swizzle, out-of-bounds elements read as zero.
int make_map(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
             uint32_t box_cols) {
  EncodeTiledFn enc = get_encode_fn();
  if (enc == nullptr) return fail("cuTensorMapEncodeTiled not available from the driver");
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {ld * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail("cuTensorMapEncodeTiled failed (%d): base=%p rows=%llu cols=%llu ld=%llu box=%ux%u", (int)r, base,
                (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld, box_rows, box_cols);
  return 0;
}

template <int BN, int CG, int MC>
static int launch_gemm(const EbGemmArgs& a, GemmDev p, cudaStream_t stream) {
  using Cfg = GemmCfg<BN, CG>;
  static_assert(MC == 1 || (CG == 2 && Cfg::B_ROWS == 128), "B multicast is built for 256-wide CTA-pair tiles");
  CUtensorMap map_a, map_b, map_c;
  if (!a.a_mn) { if (make_map(&map_a, a.A, a.M, a.K, a.lda, BLOCK_M, BLOCK_K)) return 1; }
  else         { if (make_map(&map_a, a.A, a.K, a.M, a.lda, BLOCK_K, 64)) return 1; }
  if (!a.b_mn) { if (make_map(&map_b, a.B, a.N, a.K, a.ldb, Cfg::B_ROWS / MC, BLOCK_K)) return 1; }
  else         { if (make_map(&map_b, a.B, a.K, a.N, a.ldb, BLOCK_K, 64)) return 1; }
  // K2: destination of the side-written B tiles (same geometry as B); otherwise an unused copy of map_b
  if (a.b_copy) { if (make_map(&map_c, a.b_copy, a.N, a.K, a.ldb, Cfg::B_ROWS / MC, BLOCK_K)) return 1; }
  else map_c = map_b;
  p.b_copy_on = a.b_copy != nullptr ? 1 : 0;
  constexpr int CLUSTER = CG * MC;
  p.tiles_m = (a.M + BLOCK_M * CLUSTER - 1) / (BLOCK_M * CLUSTER);
  p.tiles_n = (a.N + BN - 1) / BN;
  const int tiles = p.tiles_m * p.tiles_n;
  auto kern = gemm_bf16_kernel<BN, CG, MC>;
  static bool attr_set[64] = {false};  // per instantiation, per device
  static int max_clusters[64] = {0};
  int dev = 0;
  EB_CUDA_OK(cudaGetDevice(&dev));
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CLUSTER;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  {
    // first use of this instantiation on this device (several worker threads may race here)
    static std::mutex init_mutex;
    std::lock_guard<std::mutex> guard(init_mutex);
    if (!attr_set[dev & 63]) {
      EB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
      int n = num_sms() / CLUSTER;
      if (CLUSTER > 2) {
        // clusters of 4 do not tile every GPC evenly: ask how many can actually be co-resident
        cfg.gridDim = dim3((num_sms() / CLUSTER) * CLUSTER);
        int q = 0;
        if (cudaOccupancyMaxActiveClusters(&q, kern, &cfg) == cudaSuccess && q > 0 && q < n) n = q;
        cudaGetLastError();
      }
      max_clusters[dev & 63] = n > 0 ? n : 1;
      attr_set[dev & 63] = true;
    }
  }
  int clusters = max_clusters[dev & 63];
  // tail handling: split the tiles of a sparsely filled last wave into two half-width work items
  const int rem = tiles % clusters;
  const bool split = BN == 256 && tiles > clusters && rem > 0 && 2 * rem <= clusters && !a.no_tail_split && !a.b_copy;
  p.full_work = split ? tiles - rem : tiles;
  p.num_work = split ? tiles + rem : tiles;
  if (clusters > p.num_work) clusters = p.num_work;
  cfg.gridDim = dim3(clusters * CLUSTER);
  EB_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, map_a, map_b, map_c, p));
  count_launch();
  return 0;
}

static void pick_config(const EbGemmArgs& a, int& bn, int& cg, int& mc) {
  bn = a.block_n;
  cg = a.cta_group;
  mc = a.multicast;
  if (bn == 0) {
    const int min_bn = a.b_mn ? 64 : 16;
    const int sms = num_sms();
    // Largest tile that still gives every SM work; small problems favour more, smaller tiles.
    bn = 256;
    while (bn > min_bn) {
      const long tiles = static_cast<long>((a.M + BLOCK_M - 1) / BLOCK_M) * ((a.N + bn - 1) / bn);
      if (tiles >= sms && bn <= a.N) break;
      bn >>= 1;
    }
    if (bn > a.N) {
      int fit = min_bn;
      while (fit < a.N && fit < 256) fit <<= 1;
      bn = fit < bn ? fit : bn;
    }
  }
  if (cg == 0) {
    const long tiles2 = static_cast<long>((a.M + 2 * BLOCK_M - 1) / (2 * BLOCK_M)) * ((a.N + bn - 1) / bn);
    cg = (bn >= 128 && tiles2 >= num_sms() / 2 && a.K >= 512) ? 2 : 1;
  }
  if (mc == 0) {
    // B multicast across two CTA pairs pays once the kernel is limited by L2 -> SM traffic: big, deep GEMMs
    const long tiles4 = static_cast<long>((a.M + 4 * BLOCK_M - 1) / (4 * BLOCK_M)) * ((a.N + bn - 1) / bn);
    static const int env = getenv("EB_GEMM_MC") ? atoi(getenv("EB_GEMM_MC")) : 0;
    mc = env ? env : ((cg == 2 && bn == 256 && tiles4 >= num_sms() / 8 && a.K >= 512 && (a.M % 512 == 0 || a.M >= 2048)) ? 2 : 1);
  }
  if (cg != 2 || bn != 256) mc = 1;
}

}  // namespace eb

extern "C" int eb_gemm_bf16(const EbGemmArgs* args, void* stream_) {
  using namespace eb;
  if (args == nullptr) return fail("eb_gemm_bf16: null args");
  const EbGemmArgs& a = *args;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (a.M <= 0 || a.N <= 0 || a.K <= 0) return fail("eb_gemm_bf16: empty problem %dx%dx%d", a.M, a.N, a.K);
  if ((a.lda % 8) || (a.ldb % 8)) return fail("eb_gemm_bf16: lda/ldb must be multiples of 8 (got %d, %d)", a.lda, a.ldb);
  if ((reinterpret_cast<uintptr_t>(a.A) & 15) || (reinterpret_cast<uintptr_t>(a.B) & 15))
    return fail("eb_gemm_bf16: A and B must be 16-byte aligned");
  if (a.out_bf16 && ((a.ldo % 8) || (reinterpret_cast<uintptr_t>(a.out_bf16) & 15)))
    return fail("eb_gemm_bf16: bf16 output must be 16-byte aligned with ldo %% 8 == 0");
  if (a.out_f32 && ((a.ldo32 % 4) || (reinterpret_cast<uintptr_t>(a.out_f32) & 15)))
    return fail("eb_gemm_bf16: fp32 output must be 16-byte aligned with ldo32 %% 4 == 0");
  if (a.epi == EB_EPI_ACTGRAD && (a.aux == nullptr || (a.ldaux % 8) || (reinterpret_cast<uintptr_t>(a.aux) & 15)))
    return fail("eb_gemm_bf16: EB_EPI_ACTGRAD needs a 16-byte aligned aux with ldaux %% 8 == 0");
  if (a.drop_p < 0.f || a.drop_p >= 1.f) return fail("eb_gemm_bf16: dropout rate %f outside [0,1)", a.drop_p);
  if (a.drop_p > 0.f && (a.N % 4)) return fail("eb_gemm_bf16: dropout needs N %% 4 == 0 (got %d)", a.N);

  GemmDev p;
  p.M = a.M; p.N = a.N; p.K = a.K;
  p.a_mn = a.a_mn ? 1 : 0; p.b_mn = a.b_mn ? 1 : 0;
  p.alpha = a.alpha; p.epi = a.epi; p.act = a.act;
  p.bias = a.bias;
  p.bias_vec = (reinterpret_cast<uintptr_t>(a.bias) & 15) == 0 ? 1 : 0;
  p.aux = static_cast<const __nv_bfloat16*>(a.aux); p.ldaux = a.ldaux;
  p.drop_p = a.drop_p; p.drop_thresh = dropout_threshold(a.drop_p);
  p.drop_scale = a.drop_p > 0.f ? 1.f / (1.f - a.drop_p) : 1.f;
  p.seed = a.seed; p.drop_stream = a.drop_stream;
  p.step_ptr = static_cast<const long long*>(a.step_ptr);
  p.out_bf16 = static_cast<__nv_bfloat16*>(a.out_bf16); p.ldo = a.ldo;
  p.out_f32 = a.out_f32; p.ldo32 = a.ldo32; p.beta = a.beta;
  p.bf16_from_f32 = a.bf16_from_f32;
  p.tiles_m = p.tiles_n = 0;

  int bn, cg, mc;
  pick_config(a, bn, cg, mc);
  // TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory
  // hung the GPU: CTA pairs and the B multicast are only ever used when the caller has declared both operands to be
  // memory of this GPU; everything else (peer / IPC mappings, undeclared callers, the pull-fused GEMM) gets one plain
  // TMA load per CTA.
  if (!a.operands_local || a.b_copy) {
    if (a.cta_group == 2 || a.multicast == 2)
      return fail("eb_gemm_bf16: cta_group 2 / multicast 2 need operands_local = 1 (and no b_copy): TMA multicast loads "
                  "read only this GPU's memory");
    cg = 1;
    mc = 1;
  }
  if (a.b_copy) {
    // the pull GEMM runs on single CTAs: every B tile lands whole in one CTA's shared memory and is stored from there
    if (a.b_mn) return fail("eb_gemm_bf16: b_copy (fused weight pull) needs a K-major B operand");
    if (reinterpret_cast<uintptr_t>(a.b_copy) & 15) return fail("eb_gemm_bf16: b_copy must be 16-byte aligned");
    cg = 1;
    mc = 1;
  }
  if (a.b_mn && bn < 64) return fail("eb_gemm_bf16: MN-major B needs block_n >= 64 (got %d)", bn);
  if (cg == 2 && bn < 128) return fail("eb_gemm_bf16: cta_group 2 needs block_n >= 128 (got %d)", bn);
  if (mc == 2) return launch_gemm<256, 2, 2>(a, p, stream);
  switch (cg * 1000 + bn) {
    case 1016: return launch_gemm<16, 1, 1>(a, p, stream);
    case 1032: return launch_gemm<32, 1, 1>(a, p, stream);
    case 1064: return launch_gemm<64, 1, 1>(a, p, stream);
    case 1128: return launch_gemm<128, 1, 1>(a, p, stream);
    case 1256: return launch_gemm<256, 1, 1>(a, p, stream);
    case 2128: return launch_gemm<128, 2, 1>(a, p, stream);
    case 2256: return launch_gemm<256, 2, 1>(a, p, stream);
    default: return fail("eb_gemm_bf16: unsupported config block_n=%d cta_group=%d", bn, cg);
  }
}
