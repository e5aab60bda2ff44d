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
  void* out_bf16;         /* bf16 [M][ldo] or NULL                               */
  int32_t ldo;
  float* out_f32;         /* fp32 [M][ldo32] or NULL; out = beta*out + value      */
  int32_t ldo32;
  float beta;
  int32_t block_n;        /* 0 = auto, else 16/32/64/128/256                     */
  int32_t cta_group;      /* 0 = auto, else 1 or 2 (CTA pair, UMMA M=256)        */
} EbGemmArgs;

int eb_gemm_bf16(const EbGemmArgs* args, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ELEPHAS_B200_H_ */
