/*
 * elephas_b200 native library: C ABI of libelephas_b200.so (sm_100a only).
 *
 * Every entry point returns 0 on success and a non-zero code on failure;
 * eb_last_error() returns a human-readable description of the last failure on
 * the calling thread. All pointers are device pointers unless stated otherwise,
 * `stream` is a cudaStream_t passed as void*. The library never synchronises the
 * device on the hot path.
 *
 * The reference (arpitjain799/elephas) has no native code at all:
 * each entry point names the reference call site whose work it absorbs.
 */
#ifndef ELEPHAS_B200_H_
#define ELEPHAS_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Activation ids (Keras names). Only activations whose derivative can be written in
 * terms of the output are available to the fused epilogues. */
enum {
  EB_ACT_LINEAR = 0,
  EB_ACT_RELU = 1,
  EB_ACT_SIGMOID = 2,
  EB_ACT_TANH = 3,
  EB_ACT_SOFTPLUS = 4,
  EB_ACT_SOFTSIGN = 5,
  EB_ACT_ELU = 6,
  EB_ACT_SELU = 7,
  EB_ACT_EXPONENTIAL = 8,
  EB_ACT_HARD_SIGMOID = 9
};

/* GEMM epilogue modes. */
enum {
  EB_EPI_FWD = 0,    /* out = dropout(act(alpha*acc + bias[n]))                       */
  EB_EPI_ACTGRAD = 1 /* out = alpha*acc * act'(aux[m,n]) * dropout_mask[m,n]/(1-p)    */
};

const char* eb_last_error(void);
int eb_version(void);
/* Number of kernels this library has launched in the calling process (monotonic). */
uint64_t eb_launch_count(void);

/* ------------------------------------------------------------------ K1: dense GEMM
 * C[M,N] = epilogue(A[M,K] x B[N,K]^T), bf16 operands, fp32 accumulation in TMEM
 * (tcgen05.mma fed by TMA, warp-specialised persistent kernel).
 * Absorbs the Dense forward/backward that the reference delegates to Keras
 * (elephas/worker.py:42,108,123; elephas/spark_model.py:248,255,291).
 *
 * Operand storage (row-major, leading dimension in elements, multiple of 8):
 *   a_mn == 0: A stored [M][lda] (K contiguous)     a_mn == 1: A stored [K][lda] (M contiguous)
 *   b_mn == 0: B stored [N][ldb] (K contiguous)     b_mn == 1: B stored [K][ldb] (N contiguous)
 * so the three Dense GEMMs need no transposes in memory:
 *   forward  Y  = X  W^T : A=X  (a_mn=0), B=W [out,in] (b_mn=0)
 *   dgrad    dX = dZ W   : A=dZ (a_mn=0), B=W [out,in] (b_mn=1)
 *   wgrad    dW = dZ^T X : A=dZ (a_mn=1), B=X          (b_mn=1)
 */
typedef struct EbGemmArgs {
  int32_t M, N, K;
  const void* A; int32_t lda; int32_t a_mn;
  const void* B; int32_t ldb; int32_t b_mn;
  float alpha;
  int32_t epi;            /* EB_EPI_*                                            */
  int32_t act;            /* EB_ACT_*                                            */
  const float* bias;      /* [N] fp32 or NULL (EB_EPI_FWD only)                   */
  const void* aux;        /* bf16 [M][ldaux], stored layer output (EB_EPI_ACTGRAD) */
  int32_t ldaux;
  float drop_p;           /* dropout rate, 0 disables                            */
  uint64_t seed;          /* Philox key                                          */
  uint64_t drop_stream;   /* Philox stream id (layer / step)                     */
  const void* step_ptr;   /* optional device int64: stream id += 1024 * (*step_ptr) */
  void* out_bf16;         /* bf16 [M][ldo] or NULL                               */
  int32_t ldo;
  float* out_f32;         /* fp32 [M][ldo32] or NULL; out = beta*out + value      */
  int32_t ldo32;
  float beta;
  int32_t bf16_from_f32;  /* out_bf16 = bf16(final fp32 value) instead of bf16(alpha*acc): fused SGD weight update */
  int32_t block_n;        /* 0 = auto, else 16/32/64/128/256                     */
  int32_t cta_group;      /* 0 = auto, else 1 or 2 (CTA pair, UMMA M=256)        */
  int32_t multicast;      /* 0 = auto, 1 = off, 2 = B tile multicast across two CTA pairs (cluster of 4) */
  int32_t no_tail_split;  /* 1 disables the half-width scheduling of a sparse last wave */
  void* b_copy;           /* K2 (pull fused into the GEMM): when set, B may live in peer memory (a parameter server or
                             another rank's weights, read tile by tile through TMA over NVLink) and every B tile that
                             lands in shared memory is also written to this local bf16 buffer (same [N][ldb] layout,
                             K-major B only). The GEMM is the pull. No counterpart in the reference
                             (client.get_parameters + set_weights + first layer forward, elephas/worker.py:104-119). */
  int32_t operands_local; /* 1: the caller states that A and B are memory of the GPU this kernel runs on; only then may
                             CTA pairs (cta_group 2) and the B-tile multicast be used, by choice or by the automatic
                             pick. 0 (the default of a zeroed struct): operands may be peer / IPC / multi-GPU mappings,
                             so every tile is loaded by a single CTA with a plain TMA load - TMA multicast loads here
                             read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory
                             hung the GPU. Asking for cta_group 2 or multicast 2 without it is an error; b_copy implies 0. */
} EbGemmArgs;

int eb_gemm_bf16(const EbGemmArgs* args, void* stream);

/* ------------------------------------------------------------ elementwise helpers
 * eb_stage_f32_to_bf16: dst[r][0..cols) = bf16(src[r][0..cols)), padding columns zeroed (input staging).
 * eb_cast_f32_to_bf16:  flat cast (refresh of the bf16 compute copy of the weights).
 * eb_colsum_bf16:       out[n] (+)= sum_m x[m][n]  - the bias gradient.
 * eb_act_dropout_fwd/bwd: standalone Activation / Dropout layers (Keras layers of tests/conftest.py:12-18).
 * eb_delta (K5):        out = before - after (reference: subtract_params, elephas/worker.py:44-45). */
int eb_stage_f32_to_bf16(const float* src, int64_t ld_src, void* dst, int64_t ld_dst, int64_t rows, int32_t cols,
                         void* stream);
int eb_cast_f32_to_bf16(const float* src, void* dst, int64_t n, void* stream);
int eb_colsum_bf16(const void* x, int64_t ld, int32_t rows, int32_t cols, float* out, int32_t accumulate,
                   void* stream);
int eb_act_dropout_fwd(const void* x, void* y, int64_t ld, int32_t rows, int32_t cols, int32_t act, float drop_p,
                       uint64_t seed, uint64_t drop_stream, const void* step_ptr, void* stream);
int eb_act_dropout_bwd(const void* dy, const void* y, void* dx, int64_t ld, int32_t rows, int32_t cols, int32_t act,
                       float drop_p, uint64_t seed, uint64_t drop_stream, const void* step_ptr, void* stream);
int eb_delta(const float* before, const float* after, float* out_f32, void* out_bf16, int64_t n, void* stream);

/* ------------------------------------------------------------------ K4: fused loss
 * Output activation + Keras loss + metrics + dL/dz in one pass over the last layer's pre-activations.
 * Absorbs Keras loss / metric evaluation and their backward pass (elephas/worker.py:42,108,123;
 * elephas/spark_model.py:291). */
enum { EB_ACT_SOFTMAX = 100 };
enum { EB_LOSS_MSE = 0, EB_LOSS_MAE = 1, EB_LOSS_MAPE = 2, EB_LOSS_MSLE = 3, EB_LOSS_CCE = 4, EB_LOSS_BCE = 5,
       EB_LOSS_LOGCOSH = 6 };
enum { EB_METRIC_NONE = 0, EB_METRIC_CATEGORICAL_ACC = 1, EB_METRIC_BINARY_ACC = 2, EB_METRIC_MAE = 3,
       EB_METRIC_MSE = 4, EB_METRIC_MAPE = 5 };

typedef struct EbLossArgs {
  const float* z; int64_t ldz;   /* pre-activations fp32 [rows][ldz]                              */
  const float* y; int64_t ldy;   /* targets fp32 [rows][ldy]; ldy == 0: one scalar target per row  */
  float* out; int64_t ldo;       /* optional: activated output fp32                                */
  void* dz; int64_t lddz;        /* optional: gradient wrt z, bf16, scaled by grad_scale           */
  double* stats;                 /* optional: [0] += sum loss, [1..4] += sum metric_k, [7] += rows */
  int32_t rows, cols;
  int32_t loss, act;
  int32_t metrics[4];
  float grad_scale;
} EbLossArgs;

int eb_loss_fwd_bwd(const EbLossArgs* args, void* stream);

/* -------------------------------------------------------------- K10: fused optimizers
 * Keras 2.10 update rules over the flat parameter buffer; `apply` = 0 writes the step as a delta
 * instead of applying it (input of K6 / K7). `step` is a device int64 iteration counter, advanced
 * on the device when `advance` is set (CUDA-graph friendly). Reference: Keras optimizers reached
 * through elephas/worker.py:29-33,88-89. */
enum { EB_OPT_SGD = 1, EB_OPT_RMSPROP = 2, EB_OPT_ADAM = 3, EB_OPT_ADAGRAD = 4, EB_OPT_ADADELTA = 5,
       EB_OPT_ADAMAX = 6 };

typedef struct EbOptimArgs {
  int32_t kind;
  int64_t n;
  float* w32; const float* g32;
  float *slot0, *slot1, *slot2;
  void* w16;                /* optional bf16 copy refreshed when apply != 0 */
  void* delta_bf16; float* delta_f32;
  int32_t apply, advance;
  float lr, decay, momentum, rho, eps, beta1, beta2;
  int32_t nesterov, centered, amsgrad;
  void* step;               /* device int64 */
} EbOptimArgs;

int eb_optimizer_step(const EbOptimArgs* args, void* stream);

/* ------------------------------------------- K6: fused all-reduce + update over NVLink peer memory
 * The synchronous round of the reference (collect deltas, average, subtract:
 * elephas/spark_model.py:220-231) as one kernel per rank. `delta[r]`, `avg[r]`, `pad[r]` are the
 * peer-mapped buffers of rank r (r < world); `local` is 64 bytes of zero-initialised device memory
 * private to the rank. algo: 0 auto, 1 one-shot, 2 two-shot. Must be launched by all ranks; delta
 * buffers must be double buffered across consecutive calls. */
typedef struct EbCommArgs {
  const void* delta[8];
  void* avg[8];
  void* pad[8];
  void* local;
  float* w32; void* w16;
  float* out32;             /* eb_allreduce_sum only */
  int64_t n;
  int32_t rank, world;
  float scale;              /* 1/number_of_sub_models for the update, 1 for a plain sum */
  int32_t dtype_bf16;       /* element type of delta / avg: 1 = bf16, 0 = fp32 */
  int32_t algo, blocks;
} EbCommArgs;

int eb_allreduce_update(const EbCommArgs* args, void* stream);
int eb_allreduce_sum(const EbCommArgs* args, void* stream);
/* Sharded-master form of eb_allreduce_update: rank r reduces and updates only slice r of w32 and stores the
 * refreshed bf16 weights of that slice into every rank's compute copy, whose peer-mapped addresses are passed
 * in avg[]. The fp32 master of the other slices is NOT maintained on this rank (gather the slices afterwards). */
int eb_allreduce_shard_update(const EbCommArgs* args, void* stream);
/* NVLS form of eb_allreduce_shard_update (bf16 deltas): delta[0] and avg[0] are the MULTICAST addresses (cuMulticast*
 * objects over all ranks' buffers, see eb_mc_*) of the delta buffers and of the bf16 weight copies; the sum over ranks is
 * formed inside the NVSwitch by multimem.ld_reduce, the refreshed weights reach every rank with one multimem.st. pad[]
 * are ordinary peer mappings. About half the NVLink bytes of the peer-to-peer kernel, 1/N of its load instructions. */
int eb_allreduce_nvls_update(const EbCommArgs* args, void* stream);

/* Replicate nbytes (multiple of 16) of local memory `src` into every rank's copy of a symmetric buffer through its
 * MULTICAST address `mc_dst` (multimem.st: the sender's NVLink carries the payload once). The kernel is a pure store
 * stream; the caller orders completion (stream synchronize + barrier) before any rank reads its local copy. */
int eb_nvls_store(const void* src, void* mc_dst, int64_t nbytes, int32_t blocks, void* stream);

/* Allocation-free form of eb_allreduce_update for ONE rank (world = 1). This entry point has no
 * counterpart in the reference; it exists for hosts that are not allowed to allocate device memory at the
 * time they need the update: a serving or training process that has already handed all of the GPU's
 * memory to its own pool (a cudaMalloc there fails or fragments the pool), or a loop with a latency
 * budget that cannot absorb the device-wide synchronisation a cudaMalloc / cudaFree implies. The regular
 * entry point needs three caller-allocated buffers besides its operands (signal pad, counters, averaged
 * delta); here they live in one array compiled into the library (EB_NOALLOC_MAX_ELEMS elements of
 * averaged-delta space, about 4 MiB, held by every GPU that loads the library until the process ends).
 * The function itself calls no cudaMalloc / cudaFree and no allocating library routine.
 *
 *   fp32 delta (dtype_bf16 = 0):  w32[i] -= scale * delta[i]
 *   bf16 delta (dtype_bf16 = 1):  w32[i] -= bf16(scale * delta[i])   (the averaged delta is staged in the
 *                                 delta's own type, exactly as the multi-rank kernel exchanges it; the rounding
 *                                 is exact only when scale is a power of two)
 *   and, when w16 is given,       w16[i] = bf16(w32[i])                                            for i < n
 *
 * delta: n elements, bf16 if dtype_bf16 else fp32. n must be a multiple of 8 and at most
 * EB_NOALLOC_MAX_ELEMS; larger n is rejected with an error (use eb_allreduce_update). Results are those of
 * eb_allreduce_update with world = 1, algo = 2. With one rank there is no peer to wait for, so the peer
 * watchdog of the multi-rank kernels cannot fire here; its status word inside the array is cleared by every
 * call so that nothing stale is ever left in it.
 *
 * The first call on each device looks up the array's device address (cudaGetSymbolAddress, that device
 * must be current); make that first call outside of any stream capture. All calls on a device share the
 * one array: it keeps the previous call's contents (the function re-zeroes the counters that must start
 * at zero), and two calls that overlap on the GPU - different streams, or several host threads - race on
 * it. Issue them on one stream, or order them with events. */
#define EB_NOALLOC_MAX_ELEMS (1 << 20)
int eb_allreduce_update_noalloc(const void* delta, int64_t n, int32_t dtype_bf16, float scale, float* w32, void* w16,
                                void* stream);

/* ------------------------------------------------ K7: device-resident parameter server
 * master -= delta via red.global.add.f32 into (peer) HBM, pulls as vectorised (peer) loads fused
 * with the bf16 cast, ticket lock for mode='asynchronous' (elephas/parameter/server.py:107-132,
 * 201-218; elephas/utils/rwlock.py). ctrl: 32 bytes of zero-initialised memory next to the master. */
int eb_ps_lock(void* ctrl, void* stream);
int eb_ps_unlock(void* ctrl, void* stream);
int eb_ps_push(float* master, const float* delta, const float* before, const float* after, int64_t n, void* stream);
int eb_ps_pull(const float* master, float* dst_f32, void* dst_bf16, int64_t n, void* stream);
/* Push into a store that publishes a bf16 mirror of its master (the B operand of the pull-fused GEMM, K2):
 * master -= delta (red.add), then mirror = bf16(master) for the touched elements. Exactly one of delta_f32,
 * delta_bf16 or before/after describes the delta. No counterpart in the reference. */
int eb_ps_push_mirror(float* master, void* mirror_bf16, const float* delta_f32, const void* delta_bf16,
                      const float* before, const float* after, int64_t n, void* stream);

/* ------------------------------------------------ K7 (striped): the parameter server spread over all GPUs
 * The master weights (fp32) and their bf16 mirror are cut into STRIPES, one per rank, each resident in its owner's HBM and
 * peer-mapped by every other rank, so that pushes and pulls of N workers land on N memories, N L2s and N sets of NVLink
 * ports instead of one. `spans` maps the flat parameter index space onto the stripes: span s covers flat elements
 * [flat_off, flat_off + n) and lives at elements [owner_off, owner_off + n) of rank `owner`'s stripe (all three multiples
 * of 64; a Dense kernel is normally one span, so that the pull-fused GEMM, K2, can read it through one tensor map).
 * Every stripe is divided into `chunks` chunks of `chunk_len` elements, each guarded by its own lock record in the owner's
 * `ctrl` array (8 words per chunk: [0] gate and [1] started - one 64-bit word - [2] finished, [3] watchdog status):
 *   push, locked   (mode='asynchronous'): the block that pushes a chunk adds 1 to `started` (the same 64-bit atomic returns
 *                  the gate), applies master -= delta with red.global.add.f32 into the owner's memory, fences and adds 1 to
 *                  `finished` - inside one kernel. Additions are atomic per element, so concurrent writers need no
 *                  exclusion among themselves: every interleaving equals a serial order of the same updates
 *                  (elephas/parameter/server.py:117-132,201-210 serialises them with the write lock). The bf16 mirror is
 *                  NOT maintained by locked pushes; locked pulls convert from the master;
 *   pull, seqlock  (mode='asynchronous'): readers take NO lock: read `finished`, then `started` (equal: nobody inside), copy
 *                  the chunk, read `started` again, retry when it moved. After EB_PS_SEQ_RETRIES failed attempts the reader
 *                  raises the gate (writers that see it step back out and wait), waits for finished == started, copies and
 *                  lowers the gate: its wait is bounded under any write load. Any number of readers proceed at once - the
 *                  reference's "several threads can hold this type of lock" for reads (elephas/utils/rwlock.py:24-33,
 *                  server.py:107-115) - and a pull never observes a half-applied push inside a chunk;
 *   unlocked       (mode='hogwild'): neither (server.py:109,120): relaxed red.add / plain loads; pushes refresh the
 *                  bf16 mirror of every element they touch, pulls and the pull-fused GEMM (K2) read it.
 * The unit of consistency is the chunk: a seqlock pull never observes a half-applied push inside a chunk. */
#define EB_PS_MAX_RANKS 8
#define EB_PS_MAX_SPANS 48
#define EB_PS_MAX_CHUNKS 64
#define EB_PS_LOCK_WORDS 8
#define EB_PS_SEQ_RETRIES 8
typedef struct {
  int64_t flat_off, n, owner_off;
  int32_t owner, reserved;
} EbPsSpan;
typedef struct {
  void* master[EB_PS_MAX_RANKS];        /* fp32 stripe of every rank (peer mapped) */
  void* mirror[EB_PS_MAX_RANKS];        /* bf16 stripe of every rank (peer mapped) */
  void* ctrl[EB_PS_MAX_RANKS];          /* lock records of every rank's stripe: chunks * EB_PS_LOCK_WORDS uint32, zeroed */
  int64_t stripe_len[EB_PS_MAX_RANKS];  /* elements owned by each rank */
  int64_t chunk_len;                    /* elements per chunk (multiple of 64) */
  int32_t chunks;                       /* chunk locks per stripe (<= EB_PS_MAX_CHUNKS) */
  int32_t world, rank, nspans;
  EbPsSpan spans[EB_PS_MAX_SPANS];
} EbPsStore;
/* dst_f32[i] = master[i] for flat i in [f32_lo, f32_hi), dst_bf16[i] = bf16 of it for flat i in [b16_lo, b16_hi)
 * (destinations are indexed by FLAT element; either may be NULL). seqlock != 0: validated reads as described above, bf16
 * converted from the master; seqlock == 0: bf16 copied from the mirror. */
int eb_ps_striped_pull(const EbPsStore* store, float* dst_f32, int64_t f32_lo, int64_t f32_hi, void* dst_bf16,
                       int64_t b16_lo, int64_t b16_hi, int32_t seqlock, void* stream);
/* master -= delta over the whole flat range; exactly one of delta_f32 / delta_bf16 / (before, after) describes the
 * delta (flat indexed). locked != 0: per-chunk ticket + sequence word; locked == 0: the bf16 mirror of every touched
 * element is refreshed. */
int eb_ps_striped_push(const EbPsStore* store, const float* delta_f32, const void* delta_bf16, const float* before,
                       const float* after, int32_t locked, void* stream);
/* The optimizer step (K10: slots updated from the flat gradient opt->g32, Keras formulas) fused with the push: the step
 * of every element goes straight from registers into the owner's master as a red.add - the delta is never written to
 * memory. opt->w32 / w16 / delta_* are ignored; opt->advance advances the iteration counter afterwards. */
int eb_ps_striped_optimizer_push(const EbPsStore* store, const EbOptimArgs* opt, int32_t locked, void* stream);

/* ------------------------------------------------------------ K8: whole-MLP forward in one persistent kernel
 * z = W_{L-1} act(... act(W_0 x + b_0) ...) + b_{L-1} for a chain of 2-4 Dense layers whose hidden widths are multiples
 * of 16 up to 128 and whose output has at most 128 units: a CTA carries a tile of 128 rows through every layer, layer 0
 * streams through a TMA ring, later weights are resident in shared memory, activations go TMEM -> registers -> shared
 * memory (swizzled A operand of the next tcgen05.mma) and never touch HBM. z is the fp32 pre-activation of the last
 * layer (feed it to eb_loss_fwd_bwd for the output activation / loss / metrics). Reference: model.predict / evaluate
 * on the workers (elephas/spark_model.py:244-255,284-308). */
#define EB_MLP_MAX_LAYERS 4
typedef struct {
  int32_t rows, layers;
  int32_t dims[EB_MLP_MAX_LAYERS + 1];   /* dims[0] = input width, dims[l + 1] = units of layer l */
  const void* x;  int64_t ldx;           /* bf16 input [rows][ldx] */
  const void* w[EB_MLP_MAX_LAYERS];      /* bf16 kernels [units][ldw], K-major */
  int64_t ldw[EB_MLP_MAX_LAYERS];
  const float* bias[EB_MLP_MAX_LAYERS];  /* fp32 [units] or NULL */
  int32_t act[EB_MLP_MAX_LAYERS];        /* EB_ACT_* of the hidden layers (the last entry is ignored) */
  float* z;       int64_t ldz;           /* fp32 output [rows][ldz] */
} EbMlpArgs;
int eb_mlp_forward(const EbMlpArgs* args, void* stream);

/* ------------------------------------------------------------ narrow Dense layers (at most 8 output units)
 * Regression heads and tiny classifiers: one pass over the activations at HBM speed instead of a nearly empty
 * tensor-core tile. Layouts as in EbGemmArgs (bf16, K-major, leading dimensions in elements).
 *   eb_head_fwd    z[b][o]   = bias[o] + sum_k A[b][k] W[o][k]                       (z fp32)
 *   eb_head_wgrad  dw[o][k]  = sum_b dz[b][o] A[b][k]                                (dw fp32, overwritten)
 *   eb_head_dgrad  dzp[b][k] = (sum_o dz[b][o] W[o][k]) * act'(aux[b][k])            (aux = stored layer output)
 * Reference: the last Dense layer of the regression examples (Keras fit/predict, elephas/worker.py:42,108,123). */
typedef struct {
  int32_t rows, K, out;
  const void* A;  int64_t lda;     /* activations [rows][lda] bf16 (fwd, wgrad) */
  const void* W;  int64_t ldw;     /* weights [out][ldw] bf16 (fwd, dgrad) */
  const float* bias;               /* [out] or NULL (fwd) */
  float* z;       int64_t ldz;     /* fwd output */
  const void* dz; int64_t lddz;    /* upstream gradient [rows][lddz] bf16 (wgrad, dgrad) */
  float* dw;      int64_t lddw;    /* wgrad output [out][lddw] */
  const void* aux; int64_t ldaux;  /* previous layer's stored output (dgrad) */
  void* dzp;      int64_t lddzp;   /* dgrad output [rows][lddzp] bf16 */
  int32_t act;                     /* activation of the previous layer (EB_ACT_*) */
} EbHeadArgs;
int eb_head_fwd(const EbHeadArgs* args, void* stream);
int eb_head_wgrad(const EbHeadArgs* args, void* stream);
int eb_head_dgrad(const EbHeadArgs* args, void* stream);

/* ------------------------------------------------------------ K9: convolutions as implicit GEMMs on tcgen05
 * Channels-last bf16 activations [N][H][W][C], kernels [Cout][R][S][Cin] (OHWI, bf16 copies of the fp32 master), fp32
 * accumulation in TMEM. No im2col matrix is ever built: the A operand of the GEMM is fetched patch by patch by TMA from
 * a 4-D tensor map over the activation (a box of pn x ph x pw output pixels x 64 channels per filter tap, shifted by the
 * tap offset, element strides = convolution stride, out-of-bounds = zero padding), straight into the 128B-swizzled
 * layout tcgen05.mma reads. The reference reaches these ops through Keras' model.fit on any Keras model
 * (elephas/worker.py:42,108); BASELINE.json names a ResNet-50-shaped CNN.
 *
 *   eb_conv_igemm   out[n][p*os][q*os][k] = bias[k] + sum_{r,s,c} in[n][p*stride - pad_h + r][q*stride - pad_w + s][c]
 *                                                                  * w[k][tap(r,s)][c]
 *       forward:        in = x, out = y, weights K-major ([Cout][R*S][Cin] = the OHWI kernel itself)
 *       input gradient: in = dy, out = dx, stride 1, pad = R-1-pad, flip_taps = 1, w_mn = 1: the SAME kernel array is read
 *                       "N-major" (Cout becomes the reduction dimension), so no transposed copy of the weights exists;
 *                       for a strided 1x1 convolution os = stride scatters the result (the caller zeroes dx first)
 *   eb_conv_wgrad   dw[k][tap][c] += sum_{n,p,q} dy[n][p][q][k] * x[n][p*stride - pad_h + r][q*stride - pad_w + s][c]
 *                   (fp32, accumulated with red.add: the pixel reduction is split over CTAs; zero dw before the call)
 * Requirements: Cin and Cout multiples of 8, 16-byte aligned tensors; pw * stride <= 256. */
typedef struct {
  const void* in;  int32_t N, H, W, C;       /* bf16 [N][H][W][C]: the tensor the patches are cut from */
  const void* w;   int32_t K, R, S;          /* bf16 kernel [K][R][S][Cw]; forward: Cw = C, K = output channels;
                                                input gradient (w_mn): the forward kernel [C][R][S][K], i.e. its first
                                                dimension is the reduction and its last the output channels */
  int32_t w_mn, flip_taps;
  int32_t stride, pad_h, pad_w;
  int32_t P, Q;                              /* patch-origin grid (= output pixels of the GEMM) */
  int32_t pn, ph, pw;                        /* output pixels per tile: pn*ph*pw <= 128 */
  void* out;       int32_t OH, OW, os;       /* bf16 [N][OH][OW][Kout]; GEMM pixel (n,p,q) lands at (n, p*os, q*os) */
  const float* bias;                         /* [Kout] or NULL */
  int32_t block_n;                           /* 0 = auto, else 64 / 128 */
} EbConvArgs;
int eb_conv_igemm(const EbConvArgs* args, void* stream);

typedef struct {
  const void* x;   int32_t N, H, W, C;       /* bf16 forward input */
  const void* dy;  int32_t P, Q, K;          /* bf16 output gradient [N][P][Q][K] */
  int32_t R, S, stride, pad_h, pad_w;
  int32_t pn, ph, pw;                        /* pixels per reduction block: pn*ph*pw <= 64 */
  float* dw;                                 /* fp32 [K][R][S][C], accumulated */
  int32_t splits;                            /* 0 = auto: CTAs that share one output tile's pixel reduction */
} EbConvWgradArgs;
int eb_conv_wgrad(const EbConvWgradArgs* args, void* stream);

/* Patch matrix of the first layer (3 input channels do not fill a TMA row): out[n*P*Q + p*Q + q][(r*S + s)*C + c] =
 * x[n][p*stride - pad + r][q*stride - pad + s][c], zero padded to `ld` columns; x is fp32 or bf16 (x_is_f32). */
int eb_conv_im2col(const void* x, int32_t x_is_f32, int32_t N, int32_t H, int32_t W, int32_t C, int32_t R, int32_t S,
                   int32_t stride, int32_t pad, int32_t P, int32_t Q, void* out_bf16, int32_t ld, void* stream);

/* Pooling over channels-last bf16 [N][H][W][C] (C multiple of 8).
 *   eb_pool_max_fwd   y[n][p][q][c] = max over the R x S window at (p*stride - pad_t, q*stride - pad_l); elements outside
 *                     the image count as -inf (pad_is_zero = 0, Keras padding='same') or as zeros (pad_is_zero = 1, a
 *                     ZeroPadding2D folded into the pool). idx (bytes, same shape as y) receives the winning tap
 *                     r*S + s (255: the padding won); the first maximum in row-major tap order wins.
 *   eb_pool_max_bwd   dx[n][h][w][c] = sum of dy over the windows that contain (h, w) and were won by it (a gather over the
 *                     input pixels: no atomics)
 *   eb_pool_gap_fwd   y[n][c] = mean over the HW pixels (global average pooling), fp32 accumulation
 *   eb_pool_gap_bwd   dx[n][m][c] = dy[n][c] / HW */
int eb_pool_max_fwd(const void* x, int32_t N, int32_t H, int32_t W, int32_t C, int32_t R, int32_t S, int32_t stride,
                    int32_t pad_t, int32_t pad_l, int32_t P, int32_t Q, int32_t pad_is_zero, void* y, void* idx, void* stream);
int eb_pool_max_bwd(const void* dy, const void* idx, int32_t N, int32_t H, int32_t W, int32_t C, int32_t R, int32_t S,
                    int32_t stride, int32_t pad_t, int32_t pad_l, int32_t P, int32_t Q, void* dx, void* stream);
int eb_pool_gap_fwd(const void* x, int32_t N, int32_t HW, int32_t C, void* y, void* stream);
int eb_pool_gap_bwd(const void* dy, int32_t N, int32_t HW, int32_t C, void* dx, void* stream);

/* BatchNormalization over channels-last bf16 [M][C] (C multiple of 8), statistics in fp32. The two reductions write one
 * row of partial sums per thread block into `partials` (fp32, room for max_blocks * 2 * C values, no need to clear it)
 * and a second small kernel adds the rows: no atomics, fixed summation order.
 *   eb_bn_stats       mean, rstd = 1/sqrt(var + eps) of x; running statistics updated with `momentum` (Keras:
 *                     running = running * momentum + batch * (1 - momentum); unbiased batch variance) when given
 *   eb_bn_apply       y = act((x - mean) * rstd * gamma + beta [+ residual]),  act = ReLU when relu != 0
 *   eb_bn_bwd_reduce  dz = dy * (y > 0 if relu);  sum_dz[c] (+)= sum dz (= d beta), sum_dzx[c] (+)= sum dz * xhat
 *                     (= d gamma); with accumulate != 0 the totals are ADDED to the two vectors, which may then be the
 *                     gradient buffers of beta and gamma themselves (zeroed at the start of the step)
 *   eb_bn_bwd_apply   dx = gamma * rstd * (dz - sum_dz/M - xhat * sum_dzx/M); dres = dz when requested
 * Reference: Keras BatchNormalization / Activation / Add layers inside model.fit (elephas/worker.py:42). */
int eb_bn_stats(const void* x, int64_t M, int32_t C, float* partials, int32_t max_blocks, float eps, float momentum,
                float* mean, float* rstd, float* running_mean, float* running_var, void* stream);
int eb_bn_apply(const void* x, int64_t M, int32_t C, const float* mean, const float* rstd, const float* gamma,
                const float* beta, const void* residual, int32_t relu, void* y, void* stream);
int eb_bn_bwd_reduce(const void* dy, const void* y, const void* x, int64_t M, int32_t C, const float* mean,
                     const float* rstd, int32_t relu, float* partials, int32_t max_blocks, float* sum_dz, float* sum_dzx,
                     int32_t accumulate, void* stream);
int eb_bn_bwd_apply(const void* dy, const void* y, const void* x, int64_t M, int32_t C, const float* mean,
                    const float* rstd, const float* gamma, const float* sum_dz, const float* sum_dzx, int32_t relu,
                    void* dx, void* dres, void* stream);

/* ------------------------------------------------------------ host side of the input pipeline
 * dst[r][0..cols) = bf16(src[r][0..cols)) (round to nearest even, bit-identical to the device cast), padding columns
 * [cols, ld_dst) zeroed; runs on `threads` CPU threads (persistent pool) and returns when the batch is converted.
 * src/dst are host pointers (dst usually pinned). Halves the bytes a step sends over PCIe. No counterpart in the
 * reference. */
int eb_host_f32_to_bf16(const float* src, int64_t ld_src, void* dst, int64_t ld_dst, int64_t rows, int32_t cols,
                        int32_t threads);

/* ------------------------------------------------------------ symmetric memory on the VMM API + multicast objects
 * eb_vmm_*: physical allocations (cuMemCreate) exportable as POSIX file descriptors, import + map of a peer's allocation
 * (NVLink peer access without cudaIpc*), eb_mc_*: an NVSwitch multicast object over the same allocation of all ranks -
 * mapping ITS handle yields the address multimem.ld_reduce / multimem.st use. Protocol: every rank eb_vmm_create +
 * eb_vmm_map + export; one rank eb_mc_create + export; every rank imports, eb_mc_add_device, (barrier), eb_mc_bind of its
 * own allocation, (barrier), eb_vmm_map of the multicast handle. File descriptors travel over a Unix socket
 * (parallel/symm.py). No counterpart in the reference (Spark broadcast / collect, elephas/spark_model.py:206-227). */
int eb_vmm_supported(int32_t* vmm_out, int32_t* multicast_out);
int eb_vmm_create(int64_t bytes, int32_t multicast_ranks, uint64_t* handle_out, int64_t* size_out);
int eb_vmm_export_fd(uint64_t handle, int32_t* fd_out);
int eb_vmm_import_fd(int32_t fd, uint64_t* handle_out);
int eb_vmm_map(uint64_t handle, int64_t size, void** ptr_out);
int eb_vmm_unmap(void* ptr, int64_t size);
int eb_vmm_release(uint64_t handle);
int eb_mc_create(int32_t ranks, int64_t size, uint64_t* handle_out);
int eb_mc_add_device(uint64_t mc_handle);
int eb_mc_bind(uint64_t mc_handle, uint64_t mem_handle, int64_t size);

/* ------------------------------------------------------------ memory / IPC plumbing */
int eb_device_malloc(void** ptr, int64_t bytes);   /* cudaMalloc + zero fill (IPC-exportable) */
int eb_device_free(void* ptr);
int eb_ipc_export(void* ptr, void* handle64);
int eb_ipc_import(const void* handle64, void** ptr);
int eb_ipc_close(void* ptr);
int eb_enable_peer_access(int32_t peer_device);

#ifdef __cplusplus
}
#endif
#endif /* ELEPHAS_B200_H_ */
