// K6 / K7: collectives and the parameter server as kernels over NVLink peer memory.
//
// Every rank maps every other rank's communication buffers (CUDA IPC, see eb_ipc_*), so a kernel on
// GPU r can load from and store / reduce into the HBM of GPU p directly; NVSwitch gives each pair
// full bandwidth, so the kernels use flat all-to-all access patterns rather than rings.
//
//  * allreduce_update (K6): the synchronous-mode round of the reference - collect every worker's
//    delta, average, subtract from the master weights (elephas/spark_model.py:220-231) - as ONE
//    kernel: cross-GPU flag barrier, P2P reduction of the peers' deltas, 1/N scaling, fp32 master
//    update and bf16 weight refresh. One-shot variant (every rank reduces everything; latency bound
//    payloads) and two-shot variant (reduce-scatter of a 1/N slice + P2P all-gather of the averaged
//    slice + local apply; bandwidth bound payloads, 2(N-1)/N bytes per element over NVLink).
//  * allreduce_sum: plain one-shot sum (evaluate statistics, generic delta lists).
//  * ps_push / ps_pull / ps_lock / ps_unlock (K7): the parameter server of elephas/parameter/server.py
//    with the master weights resident in (peer) HBM: push is red.global.add.f32 of -delta straight
//    into the owner's memory (fused with delta = before - after), pull is a vectorised (peer) copy
//    fused with the bf16 cast; 'asynchronous' mode brackets them with a system-scope ticket lock
//    (the device-side counterpart of utils/rwlock.py), 'hogwild' does not (server.py:109,120).
//
// Barrier protocol: each rank owns a signal pad of 32-bit epoch flags (one per peer). A rank
// announces "my delta for epoch e is visible" by storing e into slot [rank] of every peer's pad
// (st.release.sys after a system fence) and waits until all of its own slots are >= e
// (ld.acquire.sys). Delta buffers are double buffered by the caller, which makes the start barrier
// of epoch e+1 double as the "everyone finished reading epoch e" barrier.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <string.h>

#include <mutex>

#include "../include/elephas_b200.h"
#include "host.h"
#include "ptx.cuh"

namespace eb {

constexpr int kMaxRanks = 8;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_relaxed_sys_f32(const float* p) {
  float v;
  asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.relaxed.sys.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}
__device__ __forceinline__ void red_add_v4_f32_sys(float* p, float a, float b, float c, float d) {
  asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}
__device__ __forceinline__ void red_add_f32_sys(float* p, float a) {
  asm volatile("red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(p), "f"(a) : "memory");
}

__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Watchdog: every cross-GPU wait gives up after this long, records the fact in a status word the host reads
// back, and lets the kernel finish (with meaningless results) instead of hanging the GPU - a peer that died or
// never launched must not take the whole node down (the reference has no failure detection at all).
constexpr unsigned long long kBarrierTimeoutNs = 20ull * 1000ull * 1000ull * 1000ull;
constexpr unsigned long long kLockTimeoutNs = 10ull * 1000ull * 1000ull * 1000ull;

struct CommDev {
  const void* delta[kMaxRanks];  // every rank's delta buffer for this epoch (peer mapped)
  void* avg[kMaxRanks];          // two-shot: every rank's averaged-delta buffer (peer mapped)
  uint32_t* pad[kMaxRanks];      // every rank's signal pad: [0..8) start flags, [8..16) phase flags
  uint32_t* local;               // local scratch: [0] epoch counter, [1] finished-block counter, [2] phase counter
  float* w32;
  __nv_bfloat16* w16;
  float* out32;                  // allreduce_sum output
  long n;
  int rank, world;
  float scale;
};

// All blocks wait until every peer has published epoch `e` in this rank's pad (slot base `slot0`).
__device__ __forceinline__ void wait_flags(const CommDev& p, int slot0, uint32_t e) {
  if (threadIdx.x < p.world) {
    const uint32_t* f = p.pad[p.rank] + slot0 + threadIdx.x;
    unsigned long long t0 = 0;
    uint32_t spins = 0;
    while (ld_acquire_sys(f) < e) {
      if ((++spins & 0x3FFu) == 0) {
        const unsigned long long now = global_timer_ns();
        if (t0 == 0) t0 = now;
        else if (now - t0 > kBarrierTimeoutNs) {
          atomicExch(p.local + 8, 1u + threadIdx.x);  // status word: 1 + rank of the peer that never arrived
          break;
        }
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ void publish_flags(const CommDev& p, int slot0, uint32_t e) {
  __threadfence_system();
  if (threadIdx.x < p.world) st_release_sys(p.pad[threadIdx.x] + slot0 + p.rank, e);
}

template <typename T>
__device__ __forceinline__ void accumulate8(const void* base, long i, float (&s)[8]);

template <>
__device__ __forceinline__ void accumulate8<__nv_bfloat16>(const void* base, long i, float (&s)[8]) {
  const uint4 v = ld_peer_v4(static_cast<const __nv_bfloat16*>(base) + i);
  s[0] += bf16_lo(v.x); s[1] += bf16_hi(v.x); s[2] += bf16_lo(v.y); s[3] += bf16_hi(v.y);
  s[4] += bf16_lo(v.z); s[5] += bf16_hi(v.z); s[6] += bf16_lo(v.w); s[7] += bf16_hi(v.w);
}

template <>
__device__ __forceinline__ void accumulate8<float>(const void* base, long i, float (&s)[8]) {
  const uint4 a = ld_peer_v4(static_cast<const float*>(base) + i);
  const uint4 b = ld_peer_v4(static_cast<const float*>(base) + i + 4);
  s[0] += __uint_as_float(a.x); s[1] += __uint_as_float(a.y); s[2] += __uint_as_float(a.z); s[3] += __uint_as_float(a.w);
  s[4] += __uint_as_float(b.x); s[5] += __uint_as_float(b.y); s[6] += __uint_as_float(b.z); s[7] += __uint_as_float(b.w);
}

__device__ __forceinline__ void apply8(const CommDev& p, long i, const float (&d)[8]) {
  float4 w0 = *reinterpret_cast<float4*>(p.w32 + i), w1 = *reinterpret_cast<float4*>(p.w32 + i + 4);
  w0.x -= d[0]; w0.y -= d[1]; w0.z -= d[2]; w0.w -= d[3];
  w1.x -= d[4]; w1.y -= d[5]; w1.z -= d[6]; w1.w -= d[7];
  *reinterpret_cast<float4*>(p.w32 + i) = w0;
  *reinterpret_cast<float4*>(p.w32 + i + 4) = w1;
  if (p.w16) {
    uint4 o;
    o.x = pack_bf16x2(w0.x, w0.y); o.y = pack_bf16x2(w0.z, w0.w);
    o.z = pack_bf16x2(w1.x, w1.y); o.w = pack_bf16x2(w1.z, w1.w);
    *reinterpret_cast<uint4*>(p.w16 + i) = o;
  }
}

// mode 0: one-shot all-reduce + update; mode 1: one-shot sum into out32 (no update).
template <typename T, int MODE>
__global__ void __launch_bounds__(512) allreduce_oneshot_kernel(const CommDev p) {
  const uint32_t e = p.local[0] + 1;
  if (blockIdx.x == 0) publish_flags(p, 0, e);
  wait_flags(p, 0, e);
  const long n8 = p.n / 8;
  for (long g = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; g < n8;
       g += static_cast<long>(gridDim.x) * blockDim.x) {
    float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    // fixed summation order 0..world-1 on every rank: all replicas compute bit-identical results
#pragma unroll 1
    for (int peer = 0; peer < p.world; ++peer) accumulate8<T>(p.delta[peer], g * 8, s);
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] *= p.scale;
    if (MODE == 0) {
      apply8(p, g * 8, s);
    } else {
      *reinterpret_cast<float4*>(p.out32 + g * 8) = make_float4(s[0], s[1], s[2], s[3]);
      *reinterpret_cast<float4*>(p.out32 + g * 8 + 4) = make_float4(s[4], s[5], s[6], s[7]);
    }
  }
  // the last block to finish advances the local epoch counter
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(p.local + 1, 1u) == gridDim.x - 1) {
      p.local[1] = 0;
      p.local[0] = e;
    }
  }
}

// Two-shot: reduce-scatter my slice from all peers, write the averaged slice into every peer's avg
// buffer, cross-GPU barrier, apply the full averaged delta locally. The grid must be co-resident.
template <typename T>
__global__ void __launch_bounds__(512) allreduce_twoshot_kernel(const CommDev p) {
  const uint32_t e = p.local[0] + 1;
  if (blockIdx.x == 0) publish_flags(p, 0, e);
  wait_flags(p, 0, e);
  const long n8 = p.n / 8;
  const long per = (n8 + p.world - 1) / p.world;
  const long lo = per * p.rank, hi = min(n8, lo + per);
  for (long g = lo + blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; g < hi;
       g += static_cast<long>(gridDim.x) * blockDim.x) {
    float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll 1
    for (int k = 0; k < p.world; ++k) {
      int peer = p.rank + k;
      if (peer >= p.world) peer -= p.world;
      accumulate8<T>(p.delta[peer], g * 8, s);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] *= p.scale;
    if (sizeof(T) == 2) {
      uint4 o;
      o.x = pack_bf16x2(s[0], s[1]); o.y = pack_bf16x2(s[2], s[3]);
      o.z = pack_bf16x2(s[4], s[5]); o.w = pack_bf16x2(s[6], s[7]);
#pragma unroll 1
      for (int k = 0; k < p.world; ++k) {
        int peer = p.rank + k;
        if (peer >= p.world) peer -= p.world;
        st_v4(static_cast<__nv_bfloat16*>(p.avg[peer]) + g * 8, o);
      }
    } else {
      uint4 o0 = make_uint4(__float_as_uint(s[0]), __float_as_uint(s[1]), __float_as_uint(s[2]), __float_as_uint(s[3]));
      uint4 o1 = make_uint4(__float_as_uint(s[4]), __float_as_uint(s[5]), __float_as_uint(s[6]), __float_as_uint(s[7]));
#pragma unroll 1
      for (int k = 0; k < p.world; ++k) {
        int peer = p.rank + k;
        if (peer >= p.world) peer -= p.world;
        st_v4(static_cast<float*>(p.avg[peer]) + g * 8, o0);
        st_v4(static_cast<float*>(p.avg[peer]) + g * 8 + 4, o1);
      }
    }
  }
  // phase barrier: when every block of this rank has stored its part, tell all peers
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    if (atomicAdd(p.local + 2, 1u) == gridDim.x - 1) {
      p.local[2] = 0;
      __threadfence_system();
      for (int k = 0; k < p.world; ++k) st_release_sys(p.pad[k] + 8 + p.rank, e);
    }
  }
  wait_flags(p, 8, e);
  // apply the averaged delta (now complete in my avg buffer) to the local weights
  const void* avg = p.avg[p.rank];
  for (long g = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; g < n8;
       g += static_cast<long>(gridDim.x) * blockDim.x) {
    float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    accumulate8<T>(avg, g * 8, s);
    apply8(p, g * 8, s);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(p.local + 1, 1u) == gridDim.x - 1) {
      p.local[1] = 0;
      p.local[0] = e;
    }
  }
}

// Sharded-master variant (ZeRO-1 style): rank r owns slice r of the fp32 master weights. It reduces the peers'
// deltas for its slice over NVLink, updates its fp32 slice, and stores the refreshed bf16 weights of the slice
// straight into EVERY rank's (peer-mapped) compute copy; a second flag barrier tells each rank that all slices
// of its bf16 copy have arrived. Per rank and element this moves (N-1)/N delta reads + (N-1)/N bf16 stores over
// NVLink (the two-shot optimum) but touches only 1/N of the fp32 master in local HBM - instead of every rank
// re-reading and re-writing all of it - which matters because the GEMMs running next to it are L2-bound.
template <typename T>
__device__ __forceinline__ void add8(const uint4 (&v)[sizeof(T) / 2], float (&s)[8]);

template <>
__device__ __forceinline__ void add8<__nv_bfloat16>(const uint4 (&v)[1], float (&s)[8]) {
  s[0] += bf16_lo(v[0].x); s[1] += bf16_hi(v[0].x); s[2] += bf16_lo(v[0].y); s[3] += bf16_hi(v[0].y);
  s[4] += bf16_lo(v[0].z); s[5] += bf16_hi(v[0].z); s[6] += bf16_lo(v[0].w); s[7] += bf16_hi(v[0].w);
}

template <>
__device__ __forceinline__ void add8<float>(const uint4 (&v)[2], float (&s)[8]) {
  s[0] += __uint_as_float(v[0].x); s[1] += __uint_as_float(v[0].y); s[2] += __uint_as_float(v[0].z);
  s[3] += __uint_as_float(v[0].w); s[4] += __uint_as_float(v[1].x); s[5] += __uint_as_float(v[1].y);
  s[6] += __uint_as_float(v[1].z); s[7] += __uint_as_float(v[1].w);
}

// MAXW: compile-time bound on the number of ranks; U: groups of 8 elements per thread and iteration. All U x world
// peer loads of an iteration are issued before the first one is consumed: with a ~2 us NVLink round trip the
// bytes in flight (not the block count) decide the bandwidth, and a small grid leaves the SMs to the GEMMs.
template <typename T, int MAXW, int U>
__global__ void __launch_bounds__(512) allreduce_shard_update_kernel(const CommDev p) {
  constexpr int V = sizeof(T) / 2;  // 16-byte vectors per group of 8 elements
  const uint32_t e = p.local[0] + 1;
  if (blockIdx.x == 0) publish_flags(p, 0, e);
  wait_flags(p, 0, e);
  const long n8 = p.n / 8;
  const long per = (n8 + p.world - 1) / p.world;
  const long lo = per * p.rank, hi = min(n8, lo + per);
  const long stride = static_cast<long>(gridDim.x) * blockDim.x;
  for (long g0 = lo + blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; g0 < hi; g0 += stride * U) {
    uint4 raw[U][MAXW][V];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long g = g0 + u * stride;
      if (g < hi) {
#pragma unroll
        for (int k = 0; k < MAXW; ++k) {
          if (k < p.world) {
            int peer = p.rank + k;
            if (peer >= p.world) peer -= p.world;
            const T* src = static_cast<const T*>(p.delta[peer]) + g * 8;
#pragma unroll
            for (int v = 0; v < V; ++v) raw[u][k][v] = ld_peer_v4(src + v * (8 / V));
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long g = g0 + u * stride;
      if (g < hi) {
        float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int k = 0; k < MAXW; ++k)
          if (k < p.world) add8<T>(raw[u][k], s);
        float4 w0 = *reinterpret_cast<float4*>(p.w32 + g * 8), w1 = *reinterpret_cast<float4*>(p.w32 + g * 8 + 4);
        w0.x -= s[0] * p.scale; w0.y -= s[1] * p.scale; w0.z -= s[2] * p.scale; w0.w -= s[3] * p.scale;
        w1.x -= s[4] * p.scale; w1.y -= s[5] * p.scale; w1.z -= s[6] * p.scale; w1.w -= s[7] * p.scale;
        *reinterpret_cast<float4*>(p.w32 + g * 8) = w0;
        *reinterpret_cast<float4*>(p.w32 + g * 8 + 4) = w1;
        uint4 o;
        o.x = pack_bf16x2(w0.x, w0.y); o.y = pack_bf16x2(w0.z, w0.w);
        o.z = pack_bf16x2(w1.x, w1.y); o.w = pack_bf16x2(w1.z, w1.w);
#pragma unroll
        for (int k = 0; k < MAXW; ++k) {
          if (k < p.world) {
            int peer = p.rank + k;
            if (peer >= p.world) peer -= p.world;
            st_v4(static_cast<__nv_bfloat16*>(p.avg[peer]) + g * 8, o);  // avg[]: the peers' bf16 weight buffers
          }
        }
      }
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    if (atomicAdd(p.local + 2, 1u) == gridDim.x - 1) {
      p.local[2] = 0;
      __threadfence_system();
      for (int k = 0; k < p.world; ++k) st_release_sys(p.pad[k] + 8 + p.rank, e);
    }
  }
  wait_flags(p, 8, e);
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(p.local + 1, 1u) == gridDim.x - 1) {
      p.local[1] = 0;
      p.local[0] = e;
    }
  }
}

// NVLS variant of the sharded-master kernel: the deltas of ALL ranks are added inside the NVSwitch
// (multimem.ld_reduce on the multicast address of the delta buffers: one reduced stream arrives instead of N-1), rank r
// updates slice r of its fp32 master, and ONE multimem.st per vector replicates the refreshed bf16 weights into every
// rank's compute copy. Per rank and direction about S bytes cross NVLink instead of 2 (N-1)/N S, and the SMs issue 1/N
// of the loads. p.delta[0] / p.avg[0] are the MULTICAST addresses of the delta and bf16 weight buffers (cuMulticast*
// objects, csrc/vmm.cu); the flag pads are ordinary peer mappings.
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st_bf16x8(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

template <int U>
__global__ void __launch_bounds__(512) allreduce_nvls_update_kernel(const CommDev p) {
  const uint32_t e = p.local[0] + 1;
  if (blockIdx.x == 0) publish_flags(p, 0, e);
  wait_flags(p, 0, e);
  const long n8 = p.n / 8;
  const long per = (n8 + p.world - 1) / p.world;
  const long lo = per * p.rank, hi = min(n8, lo + per);
  const long stride = static_cast<long>(gridDim.x) * blockDim.x;
  const __nv_bfloat16* mc_delta = static_cast<const __nv_bfloat16*>(p.delta[0]);
  __nv_bfloat16* mc_w16 = static_cast<__nv_bfloat16*>(p.avg[0]);
  for (long g0 = lo + blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; g0 < hi; g0 += stride * U) {
    uint4 raw[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long g = g0 + u * stride;
      if (g < hi) raw[u] = multimem_ld_reduce_bf16x8(mc_delta + g * 8);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long g = g0 + u * stride;
      if (g < hi) {
        const uint4 v = raw[u];
        float4 w0 = *reinterpret_cast<float4*>(p.w32 + g * 8), w1 = *reinterpret_cast<float4*>(p.w32 + g * 8 + 4);
        w0.x -= bf16_lo(v.x) * p.scale; w0.y -= bf16_hi(v.x) * p.scale; w0.z -= bf16_lo(v.y) * p.scale; w0.w -= bf16_hi(v.y) * p.scale;
        w1.x -= bf16_lo(v.z) * p.scale; w1.y -= bf16_hi(v.z) * p.scale; w1.z -= bf16_lo(v.w) * p.scale; w1.w -= bf16_hi(v.w) * p.scale;
        *reinterpret_cast<float4*>(p.w32 + g * 8) = w0;
        *reinterpret_cast<float4*>(p.w32 + g * 8 + 4) = w1;
        uint4 o;
        o.x = pack_bf16x2(w0.x, w0.y); o.y = pack_bf16x2(w0.z, w0.w);
        o.z = pack_bf16x2(w1.x, w1.y); o.w = pack_bf16x2(w1.z, w1.w);
        multimem_st_bf16x8(mc_w16 + g * 8, o);
      }
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    if (atomicAdd(p.local + 2, 1u) == gridDim.x - 1) {
      p.local[2] = 0;
      __threadfence_system();
      for (int k = 0; k < p.world; ++k) st_release_sys(p.pad[k] + 8 + p.rank, e);
    }
  }
  wait_flags(p, 8, e);
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(p.local + 1, 1u) == gridDim.x - 1) {
      p.local[1] = 0;
      p.local[0] = e;
    }
  }
}

// ------------------------------------------------------------------ parameter server
__global__ void ps_lock_kernel(uint32_t* ctrl) {
  // ctrl[1]: next ticket, ctrl[2]: now serving, ctrl[0]: sequence word (odd while a writer is inside),
  // ctrl[4]: watchdog status (non-zero: a waiter gave up after kLockTimeoutNs and went ahead unlocked)
  const uint32_t ticket = atomicAdd_system(ctrl + 1, 1u);
  unsigned long long t0 = 0;
  uint32_t spins = 0;
  while (ld_acquire_sys(ctrl + 2) != ticket) {
    if ((++spins & 0x3FFu) == 0) {
      const unsigned long long now = global_timer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > kLockTimeoutNs) {
        atomicExch_system(ctrl + 4, 1u);
        break;
      }
    }
  }
  atomicAdd_system(ctrl + 0, 1u);
  __threadfence_system();
}

__global__ void ps_unlock_kernel(uint32_t* ctrl) {
  __threadfence_system();
  atomicAdd_system(ctrl + 0, 1u);
  __threadfence_system();
  atomicAdd_system(ctrl + 2, 1u);
}

// master -= delta, where delta is given directly (fp32) or as before - after. red.add into (peer) HBM.
__global__ void __launch_bounds__(256) ps_push_kernel(float* __restrict__ master, const float* __restrict__ delta,
                                                      const float* __restrict__ before,
                                                      const float* __restrict__ after, long n) {
  const long n4 = n / 4;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    float4 d;
    if (delta) {
      d = reinterpret_cast<const float4*>(delta)[i];
    } else {
      const float4 b = reinterpret_cast<const float4*>(before)[i], a = reinterpret_cast<const float4*>(after)[i];
      d = make_float4(b.x - a.x, b.y - a.y, b.z - a.z, b.w - a.w);
    }
    if (d.x != 0.f || d.y != 0.f || d.z != 0.f || d.w != 0.f) red_add_v4_f32_sys(master + 4 * i, -d.x, -d.y, -d.z, -d.w);
  }
  for (long i = n4 * 4 + blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const float d = delta ? delta[i] : before[i] - after[i];
    if (d != 0.f) red_add_f32_sys(master + i, -d);
  }
}

// Push for stores that publish a bf16 mirror of the master (the operand the pull-fused GEMM, K2, reads through TMA):
// master -= delta with red.add as above, then the updated values are read back (same thread, same address: the load
// observes at least this thread's own reduction) and the mirror entries are refreshed. delta: fp32, bf16 or
// before - after. Elements whose delta is zero are skipped - their mirror entry is already current.
__global__ void __launch_bounds__(256) ps_push_mirror_kernel(float* __restrict__ master, __nv_bfloat16* __restrict__ mirror,
                                                             const float* __restrict__ delta32,
                                                             const __nv_bfloat16* __restrict__ delta16,
                                                             const float* __restrict__ before,
                                                             const float* __restrict__ after, long n) {
  const long n4 = n / 4;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    float4 d;
    if (delta32) {
      d = reinterpret_cast<const float4*>(delta32)[i];
    } else if (delta16) {
      const uint2 v = reinterpret_cast<const uint2*>(delta16)[i];
      d = make_float4(bf16_lo(v.x), bf16_hi(v.x), bf16_lo(v.y), bf16_hi(v.y));
    } else {
      const float4 b = reinterpret_cast<const float4*>(before)[i], a = reinterpret_cast<const float4*>(after)[i];
      d = make_float4(b.x - a.x, b.y - a.y, b.z - a.z, b.w - a.w);
    }
    if (d.x != 0.f || d.y != 0.f || d.z != 0.f || d.w != 0.f) {
      red_add_v4_f32_sys(master + 4 * i, -d.x, -d.y, -d.z, -d.w);
      const uint4 w = ld_peer_v4(master + 4 * i);
      uint2 o;
      o.x = pack_bf16x2(__uint_as_float(w.x), __uint_as_float(w.y));
      o.y = pack_bf16x2(__uint_as_float(w.z), __uint_as_float(w.w));
      reinterpret_cast<uint2*>(mirror)[i] = o;
    }
  }
  for (long i = n4 * 4 + blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const float d = delta32 ? delta32[i] : delta16 ? __bfloat162float(delta16[i]) : before[i] - after[i];
    if (d != 0.f) {
      red_add_f32_sys(master + i, -d);
      mirror[i] = __float2bfloat16_rn(ld_relaxed_sys_f32(master + i));
    }
  }
}

// dst32 <- master (peer loads), optionally dst16 <- bf16(master) in the same pass.
__global__ void __launch_bounds__(256) ps_pull_kernel(const float* __restrict__ master, float* __restrict__ dst32,
                                                      __nv_bfloat16* __restrict__ dst16, long n) {
  const long n4 = n / 4;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const uint4 v = ld_peer_v4(master + 4 * i);
    if (dst32) reinterpret_cast<uint4*>(dst32)[i] = v;
    if (dst16) {
      uint2 o;
      o.x = pack_bf16x2(__uint_as_float(v.x), __uint_as_float(v.y));
      o.y = pack_bf16x2(__uint_as_float(v.z), __uint_as_float(v.w));
      reinterpret_cast<uint2*>(dst16)[i] = o;
    }
  }
  for (long i = n4 * 4 + blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const float v = master[i];
    if (dst32) dst32[i] = v;
    if (dst16) dst16[i] = __float2bfloat16_rn(v);
  }
}

static int fill_comm(const EbCommArgs* a, CommDev& p) {
  if (a->world < 1 || a->world > kMaxRanks) return fail("comm: world size %d outside [1, %d]", a->world, kMaxRanks);
  if (a->rank < 0 || a->rank >= a->world) return fail("comm: bad rank %d", a->rank);
  if (a->n % 8) return fail("comm: element count must be a multiple of 8 (got %lld)", (long long)a->n);
  for (int i = 0; i < kMaxRanks; ++i) {
    p.delta[i] = i < a->world ? a->delta[i] : nullptr;
    p.avg[i] = i < a->world ? a->avg[i] : nullptr;
    p.pad[i] = i < a->world ? static_cast<uint32_t*>(a->pad[i]) : nullptr;
    if (i < a->world && (p.delta[i] == nullptr || p.pad[i] == nullptr)) return fail("comm: missing peer pointer %d", i);
  }
  p.local = static_cast<uint32_t*>(a->local);
  if (p.local == nullptr) return fail("comm: local scratch missing");
  p.w32 = a->w32; p.w16 = static_cast<__nv_bfloat16*>(a->w16); p.out32 = a->out32;
  p.n = a->n; p.rank = a->rank; p.world = a->world; p.scale = a->scale;
  return 0;
}

}  // namespace eb

using namespace eb;

extern "C" int eb_allreduce_update(const EbCommArgs* a, void* stream) {
  if (a == nullptr) return fail("eb_allreduce_update: null args");
  if (a->n <= 0) return 0;
  CommDev p;
  if (fill_comm(a, p)) return 1;
  if (p.w32 == nullptr) return fail("eb_allreduce_update: fp32 weights missing");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int blocks = a->blocks > 0 ? a->blocks : num_sms();
  const long work = a->n / 8;
  const bool twoshot = a->algo == 2 || (a->algo == 0 && a->world > 2 && a->n * (a->dtype_bf16 ? 2 : 4) > (4l << 20));
  if (twoshot) {
    for (int i = 0; i < a->world; ++i)
      if (p.avg[i] == nullptr) return fail("eb_allreduce_update: two-shot needs the avg buffers");
    if (blocks > num_sms()) blocks = num_sms();  // co-resident grid: the kernel contains a grid-wide barrier
    if (a->dtype_bf16) allreduce_twoshot_kernel<__nv_bfloat16><<<blocks, 512, 0, s>>>(p);
    else allreduce_twoshot_kernel<float><<<blocks, 512, 0, s>>>(p);
    EB_LAUNCH_OK("allreduce_twoshot_kernel");
  } else {
    if (blocks > (work + 511) / 512) blocks = static_cast<int>((work + 511) / 512);
    if (blocks < 1) blocks = 1;
    if (a->dtype_bf16) allreduce_oneshot_kernel<__nv_bfloat16, 0><<<blocks, 512, 0, s>>>(p);
    else allreduce_oneshot_kernel<float, 0><<<blocks, 512, 0, s>>>(p);
    EB_LAUNCH_OK("allreduce_oneshot_kernel");
  }
  return 0;
}

extern "C" int eb_allreduce_shard_update(const EbCommArgs* a, void* stream) {
  if (a == nullptr) return fail("eb_allreduce_shard_update: null args");
  if (a->n <= 0) return 0;
  CommDev p;
  if (fill_comm(a, p)) return 1;
  if (p.w32 == nullptr) return fail("eb_allreduce_shard_update: fp32 weights missing");
  for (int i = 0; i < a->world; ++i)
    if (p.avg[i] == nullptr) return fail("eb_allreduce_shard_update: every rank's bf16 weight buffer is required (avg[])");
  int blocks = a->blocks > 0 ? a->blocks : num_sms() / 2;
  if (blocks > num_sms()) blocks = num_sms();  // co-resident grid: the kernel contains a grid-wide barrier
  const bool small_world = a->world <= 4;
  const int per_thread = a->dtype_bf16 ? (small_world ? 4 : 2) : (small_world ? 2 : 1);
  const long work = (a->n / 8 / a->world + 512L * per_thread - 1) / (512L * per_thread);
  if (blocks > work) blocks = static_cast<int>(work);
  if (blocks < 1) blocks = 1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (a->dtype_bf16) {
    if (small_world) allreduce_shard_update_kernel<__nv_bfloat16, 4, 4><<<blocks, 512, 0, s>>>(p);
    else allreduce_shard_update_kernel<__nv_bfloat16, 8, 2><<<blocks, 512, 0, s>>>(p);
  } else {
    if (small_world) allreduce_shard_update_kernel<float, 4, 2><<<blocks, 512, 0, s>>>(p);
    else allreduce_shard_update_kernel<float, 8, 1><<<blocks, 512, 0, s>>>(p);
  }
  EB_LAUNCH_OK("allreduce_shard_update_kernel");
  return 0;
}

extern "C" int eb_allreduce_nvls_update(const EbCommArgs* a, void* stream) {
  if (a == nullptr) return fail("eb_allreduce_nvls_update: null args");
  if (a->n <= 0) return 0;
  if (a->world < 2 || a->world > kMaxRanks) return fail("eb_allreduce_nvls_update: world size %d outside [2, %d]", a->world, kMaxRanks);
  if (a->rank < 0 || a->rank >= a->world) return fail("eb_allreduce_nvls_update: bad rank %d", a->rank);
  if (a->n % 8) return fail("eb_allreduce_nvls_update: element count must be a multiple of 8");
  if (!a->dtype_bf16) return fail("eb_allreduce_nvls_update: bf16 deltas only");
  if (a->delta[0] == nullptr || a->avg[0] == nullptr || a->w32 == nullptr || a->local == nullptr)
    return fail("eb_allreduce_nvls_update: multicast delta / weight addresses, fp32 weights and local scratch are required");
  if ((reinterpret_cast<uintptr_t>(a->delta[0]) & 15) || (reinterpret_cast<uintptr_t>(a->avg[0]) & 15))
    return fail("eb_allreduce_nvls_update: multicast addresses must be 16-byte aligned");
  CommDev p;
  for (int i = 0; i < kMaxRanks; ++i) {
    p.delta[i] = nullptr; p.avg[i] = nullptr;
    p.pad[i] = i < a->world ? static_cast<uint32_t*>(a->pad[i]) : nullptr;
    if (i < a->world && p.pad[i] == nullptr) return fail("eb_allreduce_nvls_update: missing signal pad %d", i);
  }
  p.delta[0] = a->delta[0]; p.avg[0] = a->avg[0];
  p.local = static_cast<uint32_t*>(a->local);
  p.w32 = a->w32; p.w16 = nullptr; p.out32 = nullptr;
  p.n = a->n; p.rank = a->rank; p.world = a->world; p.scale = a->scale;
  int blocks = a->blocks > 0 ? a->blocks : 32;
  if (blocks > num_sms()) blocks = num_sms();  // co-resident grid: the kernel contains a grid-wide barrier
  const long work = (a->n / 8 / a->world + 512L * 4 - 1) / (512L * 4);
  if (blocks > work) blocks = static_cast<int>(work);
  if (blocks < 1) blocks = 1;
  allreduce_nvls_update_kernel<4><<<blocks, 512, 0, static_cast<cudaStream_t>(stream)>>>(p);
  EB_LAUNCH_OK("allreduce_nvls_update_kernel");
  return 0;
}

// Replicate a local buffer into every rank's memory through the multicast address of a symmetric buffer: one
// multimem.st per 16 bytes, the NVSwitch fans the data out (the sender's link carries the payload once, not N-1 times).
// Used for the weight broadcast at the start of a fit and for the gather of the sharded fp32 master at its end
// (reference: Spark broadcast of the driver's weights, elephas/spark_model.py:206-207). Completion is ordered by the
// caller (stream synchronize + barrier) - the kernel itself is a pure store stream.
namespace eb {
__global__ void __launch_bounds__(512) nvls_store_kernel(const uint4* __restrict__ src, uint4* mc_dst, long n16) {
  const long stride = static_cast<long>(gridDim.x) * blockDim.x;
  long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x;
  for (; i + 3 * stride < n16; i += 4 * stride) {  // four loads in flight per thread
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = ld_nc_u4(src + i + u * stride);
#pragma unroll
    for (int u = 0; u < 4; ++u) multimem_st_bf16x8(mc_dst + i + u * stride, v[u]);
  }
  for (; i < n16; i += stride) multimem_st_bf16x8(mc_dst + i, ld_nc_u4(src + i));
}
}  // namespace eb

extern "C" int eb_nvls_store(const void* src, void* mc_dst, int64_t nbytes, int32_t blocks, void* stream) {
  if (nbytes <= 0) return 0;
  if (src == nullptr || mc_dst == nullptr) return fail("eb_nvls_store: null pointer");
  if ((nbytes & 15) || (reinterpret_cast<uintptr_t>(src) & 15) || (reinterpret_cast<uintptr_t>(mc_dst) & 15))
    return fail("eb_nvls_store: source, multicast destination and size must be multiples of 16 bytes");
  const long n16 = static_cast<long>(nbytes / 16);
  int grid = blocks > 0 ? blocks : 64;
  if (grid > (n16 + 511) / 512) grid = static_cast<int>((n16 + 511) / 512);
  nvls_store_kernel<<<grid, 512, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const uint4*>(src),
                                                                        static_cast<uint4*>(mc_dst), n16);
  EB_LAUNCH_OK("nvls_store_kernel");
  return 0;
}

extern "C" int eb_allreduce_sum(const EbCommArgs* a, void* stream) {
  if (a == nullptr) return fail("eb_allreduce_sum: null args");
  if (a->n <= 0) return 0;
  CommDev p;
  if (fill_comm(a, p)) return 1;
  if (p.out32 == nullptr) return fail("eb_allreduce_sum: output missing");
  int blocks = a->blocks > 0 ? a->blocks : num_sms();
  const long work = a->n / 8;
  if (blocks > (work + 511) / 512) blocks = static_cast<int>((work + 511) / 512);
  if (blocks < 1) blocks = 1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (a->dtype_bf16) allreduce_oneshot_kernel<__nv_bfloat16, 1><<<blocks, 512, 0, s>>>(p);
  else allreduce_oneshot_kernel<float, 1><<<blocks, 512, 0, s>>>(p);
  EB_LAUNCH_OK("allreduce_oneshot_kernel(sum)");
  return 0;
}

extern "C" int eb_ps_lock(void* ctrl, void* stream) {
  ps_lock_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<uint32_t*>(ctrl));
  EB_LAUNCH_OK("ps_lock_kernel");
  return 0;
}

extern "C" int eb_ps_unlock(void* ctrl, void* stream) {
  ps_unlock_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<uint32_t*>(ctrl));
  EB_LAUNCH_OK("ps_unlock_kernel");
  return 0;
}

extern "C" int eb_ps_push(float* master, const float* delta, const float* before, const float* after, int64_t n,
                          void* stream) {
  if (n <= 0) return 0;
  if (delta == nullptr && (before == nullptr || after == nullptr)) return fail("eb_ps_push: need delta or before/after");
  long blocks = (n / 4 + 255) / 256;
  if (blocks > num_sms() * 4) blocks = num_sms() * 4;
  if (blocks < 1) blocks = 1;
  ps_push_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(master, delta, before, after, n);
  EB_LAUNCH_OK("ps_push_kernel");
  return 0;
}

extern "C" int eb_ps_push_mirror(float* master, void* mirror_bf16, const float* delta_f32, const void* delta_bf16,
                                 const float* before, const float* after, int64_t n, void* stream) {
  if (n <= 0) return 0;
  if (master == nullptr || mirror_bf16 == nullptr) return fail("eb_ps_push_mirror: master and mirror are required");
  if (delta_f32 == nullptr && delta_bf16 == nullptr && (before == nullptr || after == nullptr))
    return fail("eb_ps_push_mirror: need a delta (fp32 or bf16) or before/after");
  long blocks = (n / 4 + 255) / 256;
  if (blocks > num_sms() * 4) blocks = num_sms() * 4;
  if (blocks < 1) blocks = 1;
  ps_push_mirror_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      master, static_cast<__nv_bfloat16*>(mirror_bf16), delta_f32, static_cast<const __nv_bfloat16*>(delta_bf16), before,
      after, n);
  EB_LAUNCH_OK("ps_push_mirror_kernel");
  return 0;
}

extern "C" int eb_ps_pull(const float* master, float* dst_f32, void* dst_bf16, int64_t n, void* stream) {
  if (n <= 0) return 0;
  long blocks = (n / 4 + 255) / 256;
  if (blocks > num_sms() * 4) blocks = num_sms() * 4;
  if (blocks < 1) blocks = 1;
  ps_pull_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      master, dst_f32, static_cast<__nv_bfloat16*>(dst_bf16), n);
  EB_LAUNCH_OK("ps_pull_kernel");
  return 0;
}

// ------------------------------------------------------------------ allocation-free entry point
// Scratch of eb_allreduce_update_noalloc: one copy per GPU that loads this library, resident until the
// process exits. Layout: [0,256) signal pad, [256,512) kernel counters, [512, ...) averaged-delta buffer.
constexpr long kNoAllocMaxElems = 1l << 20;  // largest n accepted (fp32 elements fill the buffer exactly)
constexpr long kNoAllocAvgOffset = 512;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocAvgOffset + kNoAllocMaxElems * 4];

namespace eb {
static void* g_noalloc_ptr[64] = {nullptr};  // device address of the array above, per CUDA device ordinal
static std::mutex g_noalloc_mutex;

// In host code `g_noalloc_scratch` is only a handle for the runtime, not an address a kernel can use:
// ask the runtime for the device address once per device (with that device current) and cache it.
static int noalloc_scratch_address(unsigned char** out) {
  int dev = 0;
  EB_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return fail("eb_allreduce_update_noalloc: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> guard(g_noalloc_mutex);
  if (g_noalloc_ptr[dev] == nullptr) {
    void* p = nullptr;
    EB_CUDA_OK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    g_noalloc_ptr[dev] = p;
  }
  *out = static_cast<unsigned char*>(g_noalloc_ptr[dev]);
  return 0;
}
}  // namespace eb

extern "C" int eb_allreduce_update_noalloc(const void* delta, int64_t n, int32_t dtype_bf16, float scale, float* w32,
                                           void* w16, void* stream) {
  if (delta == nullptr || w32 == nullptr) return fail("eb_allreduce_update_noalloc: delta and w32 are required");
  if (n <= 0) return 0;
  if (n > kNoAllocMaxElems)
    return fail("eb_allreduce_update_noalloc: n = %lld exceeds the built-in scratch capacity of %ld elements; "
                "use eb_allreduce_update with caller-provided buffers for larger payloads",
                (long long)n, kNoAllocMaxElems);
  if (n % 8) return fail("eb_allreduce_update_noalloc: n must be a multiple of 8 (got %lld)", (long long)n);
  if ((reinterpret_cast<uintptr_t>(delta) & 15) || (reinterpret_cast<uintptr_t>(w32) & 15) ||
      (w16 && (reinterpret_cast<uintptr_t>(w16) & 15)))
    return fail("eb_allreduce_update_noalloc: delta, w32 and w16 must be 16-byte aligned");
  unsigned char* scratch = nullptr;
  if (noalloc_scratch_address(&scratch)) return 1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  uint32_t* local = reinterpret_cast<uint32_t*>(scratch + 256);
  // The array keeps whatever the previous call left. The epoch word (local[0]) and the pad flags are
  // monotonic on purpose; the two block counters must be zero when the kernel starts.
  EB_CUDA_OK(cudaMemsetAsync(local + 1, 0, 2 * sizeof(uint32_t), s));
  // ... and so must the watchdog status word (local[8]): with one rank no wait can time out, a non-zero
  // value could only be left over from whatever ran on this array before
  EB_CUDA_OK(cudaMemsetAsync(local + 8, 0, sizeof(uint32_t), s));
  CommDev p;
  for (int i = 0; i < kMaxRanks; ++i) { p.delta[i] = nullptr; p.avg[i] = nullptr; p.pad[i] = nullptr; }
  p.delta[0] = delta;
  p.avg[0] = scratch + kNoAllocAvgOffset;
  p.pad[0] = reinterpret_cast<uint32_t*>(scratch);
  p.local = local;
  p.w32 = w32; p.w16 = static_cast<__nv_bfloat16*>(w16); p.out32 = nullptr;
  p.n = n; p.rank = 0; p.world = 1; p.scale = scale;
  int blocks = num_sms();
  const long work = (n / 8 + 511) / 512;
  if (blocks > work) blocks = static_cast<int>(work);
  if (blocks < 1) blocks = 1;
  if (dtype_bf16) allreduce_twoshot_kernel<__nv_bfloat16><<<blocks, 512, 0, s>>>(p);
  else allreduce_twoshot_kernel<float><<<blocks, 512, 0, s>>>(p);
  EB_LAUNCH_OK("allreduce_twoshot_kernel(noalloc)");
  return 0;
}

// ------------------------------------------------------------------ memory / IPC plumbing
extern "C" int eb_device_malloc(void** ptr, int64_t bytes) {
  EB_CUDA_OK(cudaMalloc(ptr, static_cast<size_t>(bytes)));
  EB_CUDA_OK(cudaMemset(*ptr, 0, static_cast<size_t>(bytes)));
  return 0;
}

extern "C" int eb_device_free(void* ptr) {
  EB_CUDA_OK(cudaFree(ptr));
  return 0;
}

extern "C" int eb_ipc_export(void* ptr, void* handle64) {
  cudaIpcMemHandle_t h;
  EB_CUDA_OK(cudaIpcGetMemHandle(&h, ptr));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  return 0;
}

extern "C" int eb_ipc_import(const void* handle64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  EB_CUDA_OK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

extern "C" int eb_ipc_close(void* ptr) {
  EB_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

extern "C" int eb_enable_peer_access(int32_t peer_device) {
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  EB_CUDA_OK(e);
  return 0;
}
