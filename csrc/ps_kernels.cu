// K7 (striped): the parameter server of elephas/parameter/server.py spread over the HBM of all GPUs.
//
// The reference keeps the master weights in one process and serialises every worker through it: a read lock around
// get_parameters (server.py:107-115, 212-218), a write lock around `weights -= delta` (server.py:117-132, 201-210),
// both from elephas/utils/rwlock.py ("several threads can hold this type of lock" for reading, one for writing).
// Here the master (fp32) and a bf16 mirror of it are cut into one stripe per rank (EbPsStore, include/elephas_b200.h);
// a worker's push is ONE kernel whose blocks each take one chunk of one stripe: count the writer in -> red.global.add.f32
// of the negated delta straight into the owner's memory over NVLink -> count it out. Additions are atomic per element,
// so writers do not exclude each other; a pull copies a chunk between two reads of the chunk's counters and repeats the
// copy when a writer was inside or arrived (seqlock), so any number of workers read at once and none of them holds up a
// writer - unless a reader keeps losing: it then raises the chunk's gate for one copy, which writers respect. All of it
// lives inside a kernel: whoever is waited for is a block that is running right now, never a kernel queued behind the
// one that waits.
//
// The optimizer step can ride in the push kernel (ps_optimizer_push): the update of an element is computed from the
// gradient and the optimizer slots in registers and leaves the SM as a red.add into peer memory - the per-batch loop of
// elephas/worker.py:116-127 (get_parameters / train_on_batch / subtract_params / update_parameters) then costs one
// pull kernel, the training step and this kernel, with no delta buffer in between.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../include/elephas_b200.h"
#include "host.h"
#include "optim.cuh"
#include "ptx.cuh"

namespace eb {

namespace ps {

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_sys_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.relaxed.sys.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
// The fence of the ONE thread that moves a chunk's counters; the block's other threads are ordered with it by bar.sync
// (barrier + release pattern of a single thread is cumulative over what the block did before the barrier). A store whose
// stripes all live on this GPU has no observer outside of it: device scope is enough there, and several times cheaper
// (ncu of the one-GPU loop: the system-scope fences of all 512 threads were 60 % of both kernels' stall cycles).
__device__ __forceinline__ void fence_counters(bool sys) {
  if (sys) asm volatile("fence.acq_rel.sys;" ::: "memory");
  else asm volatile("fence.acq_rel.gpu;" ::: "memory");
}
__device__ __forceinline__ unsigned long long now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

constexpr unsigned long long kLockTimeoutNs = 10ull * 1000ull * 1000ull * 1000ull;
constexpr int kThreads = 512;

struct Dev {
  float* master[EB_PS_MAX_RANKS];
  __nv_bfloat16* mirror[EB_PS_MAX_RANKS];
  uint32_t* ctrl[EB_PS_MAX_RANKS];
  long long stripe_len[EB_PS_MAX_RANKS];
  long long chunk_len;
  int chunks, world, rank, nspans;
  EbPsSpan spans[EB_PS_MAX_SPANS];
};

// Which (owner, chunk) this block serves. Owners are visited starting at the caller's own rank, so that the first blocks
// of the N ranks' kernels do not all start on stripe 0.
struct Work {
  int owner;
  long long lo, hi;  // stripe-local element range of the chunk
  uint32_t* lock;    // the chunk's lock record in the owner's memory
};

__device__ __forceinline__ bool block_work(const Dev& p, Work& w) {
  const int k = blockIdx.x / p.chunks, c = blockIdx.x % p.chunks;
  int owner = p.rank + k;
  if (owner >= p.world) owner -= p.world;
  w.owner = owner;
  w.lo = static_cast<long long>(c) * p.chunk_len;
  w.hi = min(p.stripe_len[owner], w.lo + p.chunk_len);
  w.lock = p.ctrl[owner] + c * EB_PS_LOCK_WORDS;
  return w.lo < w.hi;
}

// f(flat index of a group of 8 elements, stripe-local index of the same group), for every group of the chunk.
template <typename F>
__device__ __forceinline__ void for_each_group(const Dev& p, const Work& w, F f) {
  for (int s = 0; s < p.nspans; ++s) {
    const EbPsSpan& sp = p.spans[s];
    if (sp.owner != w.owner) continue;
    const long long a = max(w.lo, static_cast<long long>(sp.owner_off));
    const long long b = min(w.hi, static_cast<long long>(sp.owner_off + sp.n));
    const long long shift = sp.flat_off - sp.owner_off;
    for (long long g = a + 8ll * threadIdx.x; g < b; g += 8ll * kThreads) f(g + shift, g);
  }
}

// Lock record of a chunk (8 words in the owner's memory): [0] gate, [1] started - ONE 64-bit word, gate in the low half -
// [2] finished, [3] watchdog status. Pushes are element-wise atomic additions, so writers need no exclusion among
// themselves (any interleaving equals some serial order of the same updates); what a chunk's record provides is the
// READER's guarantee of elephas/utils/rwlock.py: a pull never observes a half-applied push.
//   writer   started += 1 (one 64-bit atomic that also returns the gate), red.adds, fence, finished += 1. Wait-free
//            unless the gate is up: then it steps back out (finished += 1 without having written) and waits for the gate.
//   reader   optimistic: read finished, then {gate, started}; equal counters = nobody inside; copy; re-read started;
//            retry when it moved. After EB_PS_SEQ_RETRIES failed attempts it raises the gate (gate += 1 in the same word
//            the writers increment, so every writer is ordered before or after it), waits until finished == started,
//            copies - no writer can be inside - and lowers the gate. That bounds a reader's wait under any write load.
// Counters wrap modulo 2^32 together (started sits in the HIGH half: its carry falls off the word, never into the gate).
// A waiter that is still not served after kLockTimeoutNs records that in the status word and goes ahead (the host
// reports it; nothing hangs). Thread 0 of the block calls these.
__device__ __forceinline__ unsigned long long ld_acquire_sys_u64(const void* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t gate_of(unsigned long long w) { return static_cast<uint32_t>(w); }
__device__ __forceinline__ uint32_t started_of(unsigned long long w) { return static_cast<uint32_t>(w >> 32); }

struct Watchdog {
  unsigned long long t0 = 0;
  uint32_t spins = 0;
  __device__ __forceinline__ bool expired(uint32_t* lk, uint32_t code) {
    if ((++spins & 0x3FFu) != 0) return false;
    const unsigned long long now = now_ns();
    if (t0 == 0) { t0 = now; return false; }
    if (now - t0 <= kLockTimeoutNs) return false;
    atomicExch_system(lk + 3, code);
    return true;
  }
};

__device__ __forceinline__ void writer_begin(uint32_t* lk) {
  unsigned long long* word = reinterpret_cast<unsigned long long*>(lk);
  Watchdog dog;
  for (;;) {
    const unsigned long long old = atomicAdd_system(word, 1ull << 32);  // started += 1; the same atomic reads the gate
    if (gate_of(old) == 0) return;
    atomicAdd_system(lk + 2, 1u);  // a reader insists on a quiet chunk: step back out ...
    bool give_up = false;
    while (gate_of(ld_acquire_sys_u64(word)) != 0 && !give_up) give_up = dog.expired(lk, 1u);  // ... until it is done
    if (give_up) {
      atomicAdd_system(word, 1ull << 32);
      return;
    }
  }
}

__device__ __forceinline__ void writer_end(uint32_t* lk) { atomicAdd_system(lk + 2, 1u); }

// Reader that has run out of optimistic attempts: raise the gate, wait for the writers inside to leave.
__device__ __forceinline__ void reader_gate_up(uint32_t* lk) {
  unsigned long long* word = reinterpret_cast<unsigned long long*>(lk);
  atomicAdd_system(word, 1ull);
  Watchdog dog;
  for (;;) {
    const uint32_t finished = ld_acquire_sys(lk + 2);
    if (started_of(ld_acquire_sys_u64(word)) == finished) return;
    if (dog.expired(lk, 2u)) return;
  }
}
__device__ __forceinline__ void reader_gate_down(uint32_t* lk) {
  atomicAdd_system(reinterpret_cast<unsigned long long*>(lk), ~0ull);  // gate -= 1 (never below zero: no borrow)
}

// Writer side, executed by the whole block around its reductions.
__device__ __forceinline__ void writer_enter(uint32_t* lk, bool sys) {
  if (threadIdx.x == 0) {
    writer_begin(lk);
    fence_counters(sys);  // the counter has moved before any reduction of this block lands
  }
  __syncthreads();
}

__device__ __forceinline__ void writer_leave(uint32_t* lk, bool sys) {
  __syncthreads();  // every thread's reductions are issued ...
  if (threadIdx.x == 0) {
    fence_counters(sys);  // ... and performed (release side of the counter, cumulative over the block) before ...
    writer_end(lk);       // ... the chunk counts this writer as finished
  }
}

enum { SRC_F32 = 0, SRC_BF16 = 1, SRC_DIFF = 2 };

struct PushSrc {
  const float* d32;
  const __nv_bfloat16* d16;
  const float* before;
  const float* after;
};

// master[g..g+8) -= d, then the bf16 mirror of the same elements is rewritten from the values read back (same thread,
// same address: the load observes at least this thread's own reduction). All-zero groups are skipped.
template <bool MIRROR>
__device__ __forceinline__ void apply_group(const Dev& p, int owner, long long g, const float (&d)[8]) {
  bool any = false;
#pragma unroll
  for (int j = 0; j < 8; ++j) any |= d[j] != 0.f;
  if (!any) return;
  float* m = p.master[owner] + g;
  red_add_v4(m, -d[0], -d[1], -d[2], -d[3]);
  red_add_v4(m + 4, -d[4], -d[5], -d[6], -d[7]);
  if (!MIRROR) return;
  const uint4 a = ld_sys_v4(m), b = ld_sys_v4(m + 4);
  uint4 o;
  o.x = pack_bf16x2(__uint_as_float(a.x), __uint_as_float(a.y));
  o.y = pack_bf16x2(__uint_as_float(a.z), __uint_as_float(a.w));
  o.z = pack_bf16x2(__uint_as_float(b.x), __uint_as_float(b.y));
  o.w = pack_bf16x2(__uint_as_float(b.z), __uint_as_float(b.w));
  st_v4(p.mirror[owner] + g, o);
}

template <int SRC, bool LOCKED>
__global__ void __launch_bounds__(kThreads) ps_push_kernel(const Dev p, const PushSrc src) {
  Work w;
  if (!block_work(p, w)) return;
  if (LOCKED) writer_enter(w.lock, p.world > 1);
  for_each_group(p, w, [&](long long f, long long g) {
    float d[8];
    if (SRC == SRC_F32) {
      const float4 a = *reinterpret_cast<const float4*>(src.d32 + f), b = *reinterpret_cast<const float4*>(src.d32 + f + 4);
      d[0] = a.x; d[1] = a.y; d[2] = a.z; d[3] = a.w; d[4] = b.x; d[5] = b.y; d[6] = b.z; d[7] = b.w;
    } else if (SRC == SRC_BF16) {
      const uint4 v = *reinterpret_cast<const uint4*>(src.d16 + f);
      d[0] = bf16_lo(v.x); d[1] = bf16_hi(v.x); d[2] = bf16_lo(v.y); d[3] = bf16_hi(v.y);
      d[4] = bf16_lo(v.z); d[5] = bf16_hi(v.z); d[6] = bf16_lo(v.w); d[7] = bf16_hi(v.w);
    } else {
      const float4 b0 = *reinterpret_cast<const float4*>(src.before + f), b1 = *reinterpret_cast<const float4*>(src.before + f + 4);
      const float4 a0 = *reinterpret_cast<const float4*>(src.after + f), a1 = *reinterpret_cast<const float4*>(src.after + f + 4);
      d[0] = b0.x - a0.x; d[1] = b0.y - a0.y; d[2] = b0.z - a0.z; d[3] = b0.w - a0.w;
      d[4] = b1.x - a1.x; d[5] = b1.y - a1.y; d[6] = b1.z - a1.z; d[7] = b1.w - a1.w;
    }
    apply_group<!LOCKED>(p, w.owner, g, d);
  });
  if (LOCKED) writer_leave(w.lock, p.world > 1);
}

// Optimizer step fused with the push: slots are updated in place from the flat gradient, the step of every element goes
// from registers into the owner's master. The delta never exists in memory.
template <int KIND, bool LOCKED>
__global__ void __launch_bounds__(kThreads) ps_optimizer_push_kernel(const Dev p, const OptimDev o) {
  Work w;
  if (!block_work(p, w)) return;
  __shared__ float sh_lr[2];
  if (threadIdx.x == 32) step_rates<KIND>(o, sh_lr[0], sh_lr[1]);
  if (LOCKED) writer_enter(w.lock, p.world > 1);
  else __syncthreads();
  const float lr_t = sh_lr[0], lr_adam = sh_lr[1];
  for_each_group(p, w, [&](long long f, long long g) {
    float d[8];
    {
      float st[4];
      step_quad<KIND>(o, f / 4, ld4(o.g32, f / 4), lr_t, lr_adam, st);
      d[0] = st[0]; d[1] = st[1]; d[2] = st[2]; d[3] = st[3];
      step_quad<KIND>(o, f / 4 + 1, ld4(o.g32, f / 4 + 1), lr_t, lr_adam, st);
      d[4] = st[0]; d[5] = st[1]; d[6] = st[2]; d[7] = st[3];
    }
    apply_group<!LOCKED>(p, w.owner, g, d);
  });
  if (LOCKED) writer_leave(w.lock, p.world > 1);
}

struct PullDst {
  float* f32;
  long long f32_lo, f32_hi;
  __nv_bfloat16* b16;
  long long b16_lo, b16_hi;
};

// FROM_MASTER: the bf16 destination is converted from the fp32 master (stores whose pushes run under the chunk locks
// do not maintain the mirror: the read-back it needs would double the time a lock is held); otherwise it is copied from
// the bf16 mirror (half the NVLink bytes; hogwild stores keep it current for the pull-fused GEMM).
template <bool FROM_MASTER>
__device__ __forceinline__ void copy_chunk(const Dev& p, const Work& w, const PullDst& d) {
  for_each_group(p, w, [&](long long f, long long g) {
    // (ranges are multiples of 8 elements: a group is wholly inside or outside)
    const bool want32 = d.f32 != nullptr && f >= d.f32_lo && f < d.f32_hi;
    const bool want16 = d.b16 != nullptr && f >= d.b16_lo && f < d.b16_hi;
    if (want32 || (want16 && FROM_MASTER)) {
      const float* m = p.master[w.owner] + g;
      const uint4 a = ld_sys_v4(m), b = ld_sys_v4(m + 4);
      if (want32) {
        *reinterpret_cast<uint4*>(d.f32 + f) = a;
        *reinterpret_cast<uint4*>(d.f32 + f + 4) = b;
      }
      if (want16 && FROM_MASTER) {
        uint4 o;
        o.x = pack_bf16x2(__uint_as_float(a.x), __uint_as_float(a.y));
        o.y = pack_bf16x2(__uint_as_float(a.z), __uint_as_float(a.w));
        o.z = pack_bf16x2(__uint_as_float(b.x), __uint_as_float(b.y));
        o.w = pack_bf16x2(__uint_as_float(b.z), __uint_as_float(b.w));
        *reinterpret_cast<uint4*>(d.b16 + f) = o;
      }
    }
    if (want16 && !FROM_MASTER) *reinterpret_cast<uint4*>(d.b16 + f) = ld_sys_v4(p.mirror[w.owner] + g);
  });
}

// Reader (see the lock record above): optimistic attempts between two reads of the counters; readers never exclude each
// other and never make a writer wait. A chunk that keeps changing under the reader (EB_PS_SEQ_RETRIES failed attempts) is
// copied behind the gate instead, which bounds the wait.
template <bool SEQLOCK>
__global__ void __launch_bounds__(kThreads) ps_pull_kernel(const Dev p, const PullDst d) {
  Work w;
  if (!block_work(p, w)) return;
  if constexpr (!SEQLOCK) {
    copy_chunk<false>(p, w, d);
  } else {
  __shared__ uint32_t sh_seq, sh_again;
  const unsigned long long* word = reinterpret_cast<const unsigned long long*>(w.lock);
  for (int attempt = 0;; ++attempt) {
    if (attempt >= EB_PS_SEQ_RETRIES) {
      if (threadIdx.x == 0) reader_gate_up(w.lock);
      __syncthreads();
      copy_chunk<true>(p, w, d);
      __syncthreads();
      if (threadIdx.x == 0) {
        fence_counters(p.world > 1);
        reader_gate_down(w.lock);
      }
      return;
    }
    if (threadIdx.x == 0) {
      // finished FIRST, then started: equal values mean that nobody was inside when `started` was read
      Watchdog dog;
      uint32_t finished = ld_acquire_sys(w.lock + 2);
      uint32_t started = started_of(ld_acquire_sys_u64(word));
      while (started != finished) {  // a writer is inside: its block is running, this ends within microseconds
        if (dog.expired(w.lock, 2u)) break;
        finished = ld_acquire_sys(w.lock + 2);
        started = started_of(ld_acquire_sys_u64(word));
      }
      sh_seq = started;
    }
    __syncthreads();
    copy_chunk<true>(p, w, d);
    __syncthreads();  // every thread's loads have returned (their values were stored) ...
    if (threadIdx.x == 0) {
      fence_counters(p.world > 1);  // ... before the counter is read again
      sh_again = started_of(ld_acquire_sys_u64(word)) != sh_seq;  // the version re-read
    }
    __syncthreads();
    if (!sh_again) return;
  }
  }
}

static int fill(const EbPsStore* a, Dev& p, const char* who) {
  if (a == nullptr) return fail("%s: null store", who);
  if (a->world < 1 || a->world > EB_PS_MAX_RANKS) return fail("%s: world size %d outside [1, %d]", who, a->world, EB_PS_MAX_RANKS);
  if (a->rank < 0 || a->rank >= a->world) return fail("%s: bad rank %d", who, a->rank);
  if (a->nspans < 1 || a->nspans > EB_PS_MAX_SPANS) return fail("%s: %d spans (1..%d supported)", who, a->nspans, EB_PS_MAX_SPANS);
  if (a->chunks < 1 || a->chunks > EB_PS_MAX_CHUNKS) return fail("%s: %d chunks per stripe (1..%d supported)", who, a->chunks, EB_PS_MAX_CHUNKS);
  if (a->chunk_len <= 0 || a->chunk_len % 64) return fail("%s: chunk length must be a positive multiple of 64", who);
  for (int r = 0; r < EB_PS_MAX_RANKS; ++r) {
    const bool live = r < a->world;
    p.master[r] = live ? static_cast<float*>(a->master[r]) : nullptr;
    p.mirror[r] = live ? static_cast<__nv_bfloat16*>(a->mirror[r]) : nullptr;
    p.ctrl[r] = live ? static_cast<uint32_t*>(a->ctrl[r]) : nullptr;
    p.stripe_len[r] = live ? a->stripe_len[r] : 0;
    if (live && a->stripe_len[r] > 0 && (p.master[r] == nullptr || p.mirror[r] == nullptr || p.ctrl[r] == nullptr))
      return fail("%s: stripe %d has no master / mirror / lock records", who, r);
    if (live && (a->stripe_len[r] % 64 || a->stripe_len[r] > a->chunk_len * a->chunks))
      return fail("%s: stripe %d length %lld does not fit %d chunks of %lld", who, r, (long long)a->stripe_len[r], a->chunks,
                  (long long)a->chunk_len);
  }
  for (int s = 0; s < a->nspans; ++s) {
    const EbPsSpan& sp = a->spans[s];
    if (sp.owner < 0 || sp.owner >= a->world || sp.n <= 0 || sp.n % 64 || sp.flat_off % 64 || sp.owner_off % 64 ||
        sp.owner_off + sp.n > a->stripe_len[sp.owner])
      return fail("%s: span %d is malformed", who, s);
    p.spans[s] = sp;
  }
  p.chunk_len = a->chunk_len;
  p.chunks = a->chunks; p.world = a->world; p.rank = a->rank; p.nspans = a->nspans;
  return 0;
}

}  // namespace ps
}  // namespace eb

using namespace eb;

extern "C" int eb_ps_striped_pull(const EbPsStore* store, float* dst_f32, int64_t f32_lo, int64_t f32_hi, void* dst_bf16,
                                  int64_t b16_lo, int64_t b16_hi, int32_t seqlock, void* stream) {
  ps::Dev p;
  if (ps::fill(store, p, "eb_ps_striped_pull")) return 1;
  if ((f32_lo | f32_hi | b16_lo | b16_hi) % 8) return fail("eb_ps_striped_pull: ranges must be multiples of 8 elements");
  if ((reinterpret_cast<uintptr_t>(dst_f32) & 15) || (reinterpret_cast<uintptr_t>(dst_bf16) & 15))
    return fail("eb_ps_striped_pull: destinations must be 16-byte aligned");
  ps::PullDst d{dst_f32, f32_lo, f32_hi, static_cast<__nv_bfloat16*>(dst_bf16), b16_lo, b16_hi};
  if ((d.f32 == nullptr || d.f32_lo >= d.f32_hi) && (d.b16 == nullptr || d.b16_lo >= d.b16_hi)) return 0;
  const int grid = p.world * p.chunks;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (seqlock) ps::ps_pull_kernel<true><<<grid, ps::kThreads, 0, s>>>(p, d);
  else ps::ps_pull_kernel<false><<<grid, ps::kThreads, 0, s>>>(p, d);
  EB_LAUNCH_OK("ps_pull_kernel(striped)");
  return 0;
}

extern "C" int eb_ps_striped_push(const EbPsStore* store, const float* delta_f32, const void* delta_bf16,
                                  const float* before, const float* after, int32_t locked, void* stream) {
  ps::Dev p;
  if (ps::fill(store, p, "eb_ps_striped_push")) return 1;
  const int given = (delta_f32 != nullptr) + (delta_bf16 != nullptr) + (before != nullptr && after != nullptr);
  if (given != 1) return fail("eb_ps_striped_push: give exactly one of delta_f32, delta_bf16 or before/after");
  if ((reinterpret_cast<uintptr_t>(delta_f32) & 15) || (reinterpret_cast<uintptr_t>(delta_bf16) & 15) ||
      (reinterpret_cast<uintptr_t>(before) & 15) || (reinterpret_cast<uintptr_t>(after) & 15))
    return fail("eb_ps_striped_push: sources must be 16-byte aligned");
  ps::PushSrc src{delta_f32, static_cast<const __nv_bfloat16*>(delta_bf16), before, after};
  const int grid = p.world * p.chunks;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (delta_f32) {
    if (locked) ps::ps_push_kernel<ps::SRC_F32, true><<<grid, ps::kThreads, 0, s>>>(p, src);
    else ps::ps_push_kernel<ps::SRC_F32, false><<<grid, ps::kThreads, 0, s>>>(p, src);
  } else if (delta_bf16) {
    if (locked) ps::ps_push_kernel<ps::SRC_BF16, true><<<grid, ps::kThreads, 0, s>>>(p, src);
    else ps::ps_push_kernel<ps::SRC_BF16, false><<<grid, ps::kThreads, 0, s>>>(p, src);
  } else {
    if (locked) ps::ps_push_kernel<ps::SRC_DIFF, true><<<grid, ps::kThreads, 0, s>>>(p, src);
    else ps::ps_push_kernel<ps::SRC_DIFF, false><<<grid, ps::kThreads, 0, s>>>(p, src);
  }
  EB_LAUNCH_OK("ps_push_kernel(striped)");
  return 0;
}

namespace {
template <int KIND>
void launch_opt_push(const ps::Dev& p, const OptimDev& o, bool locked, int grid, cudaStream_t s) {
  if (locked) ps::ps_optimizer_push_kernel<KIND, true><<<grid, ps::kThreads, 0, s>>>(p, o);
  else ps::ps_optimizer_push_kernel<KIND, false><<<grid, ps::kThreads, 0, s>>>(p, o);
}
}  // namespace

extern "C" int eb_ps_striped_optimizer_push(const EbPsStore* store, const EbOptimArgs* opt, int32_t locked, void* stream) {
  ps::Dev p;
  if (ps::fill(store, p, "eb_ps_striped_optimizer_push")) return 1;
  if (opt == nullptr) return fail("eb_ps_striped_optimizer_push: null optimizer args");
  EbOptimArgs a = *opt;
  a.apply = 0; a.w32 = nullptr; a.w16 = nullptr; a.delta_bf16 = nullptr; a.delta_f32 = nullptr;
  OptimDev o;
  if (fill_optim(&a, o, "eb_ps_striped_optimizer_push")) return 1;
  const int grid = p.world * p.chunks;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (a.kind) {
    case EB_OPT_SGD: launch_opt_push<EB_OPT_SGD>(p, o, locked != 0, grid, s); break;
    case EB_OPT_RMSPROP: launch_opt_push<EB_OPT_RMSPROP>(p, o, locked != 0, grid, s); break;
    case EB_OPT_ADAM: launch_opt_push<EB_OPT_ADAM>(p, o, locked != 0, grid, s); break;
    case EB_OPT_ADAGRAD: launch_opt_push<EB_OPT_ADAGRAD>(p, o, locked != 0, grid, s); break;
    case EB_OPT_ADADELTA: launch_opt_push<EB_OPT_ADADELTA>(p, o, locked != 0, grid, s); break;
    default: launch_opt_push<EB_OPT_ADAMAX>(p, o, locked != 0, grid, s); break;
  }
  EB_LAUNCH_OK("ps_optimizer_push_kernel");
  if (a.advance) {
    launch_advance_step(a.step, s);
    EB_LAUNCH_OK("advance_step_kernel");
  }
  return 0;
}
