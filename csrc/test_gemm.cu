// Standalone numerics + speed harness for eb_gemm_bf16 (no torch, quick start-up).
//   test_gemm check M N K a_mn b_mn block_n cta_group epi act drop_p [f32out]
//       -> prints "PASS"/"FAIL" with max error vs a naive fp32 GPU reference
//   test_gemm bench M N K a_mn b_mn block_n cta_group iters
//       -> prints achieved TFLOP/s (CUDA events, L2 flushed between iterations)
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../include/elephas_b200.h"
#include "math.cuh"

#define CK(x)                                                                        \
  do {                                                                               \
    cudaError_t e = (x);                                                             \
    if (e != cudaSuccess) {                                                          \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
      exit(2);                                                                       \
    }                                                                                \
  } while (0)

__global__ void ref_gemm(int M, int N, int K, const __nv_bfloat16* A, int lda, int a_mn, const __nv_bfloat16* B,
                         int ldb, int b_mn, float alpha, int epi, int act, const float* bias,
                         const __nv_bfloat16* aux, int ldaux, float drop_p, uint64_t seed, uint64_t stream,
                         float* out, const float* prev, float beta) {
  int n = blockIdx.x * blockDim.x + threadIdx.x;
  int m = blockIdx.y;
  if (n >= N || m >= M) return;
  float acc = 0.f;
  for (int k = 0; k < K; ++k) {
    float a = __bfloat162float(a_mn ? A[(size_t)k * lda + m] : A[(size_t)m * lda + k]);
    float b = __bfloat162float(b_mn ? B[(size_t)k * ldb + n] : B[(size_t)n * ldb + k]);
    acc += a * b;
  }
  float x = alpha * acc;
  bool keep = true;
  float scale = 1.f;
  if (drop_p > 0.f) {
    bool k4[4];
    uint64_t idx = (uint64_t)m * N + n;
    eb::dropout_keep4(seed, stream, idx & ~3ull, eb::dropout_threshold(drop_p), k4);
    keep = k4[idx & 3];
    scale = 1.f / (1.f - drop_p);
  }
  if (epi == EB_EPI_FWD) {
    if (bias) x += bias[n];
    x = eb::act_apply(act, x);
    x = keep ? x * scale : 0.f;
  } else {
    float y = __bfloat162float(aux[(size_t)m * ldaux + n]);
    if (drop_p > 0.f) y *= (1.f - drop_p);
    x = keep ? x * scale * eb::act_grad_from_out(act, y) : 0.f;
  }
  if (prev) x += beta * prev[(size_t)m * N + n];
  out[(size_t)m * N + n] = x;
}

static float frand() { return (float)rand() / RAND_MAX * 2.f - 1.f; }
static int round8(int x) { return (x + 7) / 8 * 8; }

int main(int argc, char** argv) {
  if (argc < 9) {
    printf("usage\n");
    return 2;
  }
  const char* mode = argv[1];
  int M = atoi(argv[2]), N = atoi(argv[3]), K = atoi(argv[4]);
  int a_mn = atoi(argv[5]), b_mn = atoi(argv[6]), bn = atoi(argv[7]), cg = atoi(argv[8]);
  srand(1234);
  // padded leading dimensions (+8 so that ld != logical extent is exercised)
  int lda = a_mn ? round8(M) + 8 : round8(K) + 8;
  int ldb = b_mn ? round8(N) + 8 : round8(K) + 8;
  size_t a_elems = (size_t)(a_mn ? K : M) * lda, b_elems = (size_t)(b_mn ? K : N) * ldb;
  std::vector<__nv_bfloat16> hA(a_elems), hB(b_elems);
  for (auto& v : hA) v = __float2bfloat16(frand());
  for (auto& v : hB) v = __float2bfloat16(frand());
  __nv_bfloat16 *dA, *dB;
  CK(cudaMalloc(&dA, a_elems * 2));
  CK(cudaMalloc(&dB, b_elems * 2));
  CK(cudaMemcpy(dA, hA.data(), a_elems * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, hB.data(), b_elems * 2, cudaMemcpyHostToDevice));
  int ldo = round8(N) + 8;
  __nv_bfloat16* dO;
  CK(cudaMalloc(&dO, (size_t)M * ldo * 2));
  CK(cudaMemset(dO, 0, (size_t)M * ldo * 2));

  EbGemmArgs g;
  memset(&g, 0, sizeof(g));
  g.M = M; g.N = N; g.K = K;
  g.A = dA; g.lda = lda; g.a_mn = a_mn;
  g.B = dB; g.ldb = ldb; g.b_mn = b_mn;
  g.alpha = 1.f;
  g.out_bf16 = dO; g.ldo = ldo;
  g.block_n = bn; g.cta_group = cg;
  g.no_tail_split = getenv("EB_NO_TAIL_SPLIT") ? 1 : 0;
  g.operands_local = 1;  // A and B are cudaMalloc'ed on the one device this program uses
  g.multicast = getenv("EB_MC") ? atoi(getenv("EB_MC")) : 0;

  if (!strcmp(mode, "bench")) {
    int iters = argc > 9 ? atoi(argv[9]) : 10;
    size_t flush_bytes = 256ull << 20;
    void* flush;
    CK(cudaMalloc(&flush, flush_bytes));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int i = 0; i < 3; ++i)
      if (eb_gemm_bf16(&g, 0)) { printf("ERR %s\n", eb_last_error()); return 1; }
    CK(cudaDeviceSynchronize());
    float best = 1e30f, total = 0.f;
    for (int i = 0; i < iters; ++i) {
      CK(cudaMemsetAsync(flush, i, flush_bytes, 0));
      CK(cudaEventRecord(e0, 0));
      if (eb_gemm_bf16(&g, 0)) { printf("ERR %s\n", eb_last_error()); return 1; }
      CK(cudaEventRecord(e1, 0));
      CK(cudaEventSynchronize(e1));
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      best = fminf(best, ms);
      total += ms;
    }
    double fl = 2.0 * M * N * K;
    printf("BENCH M=%d N=%d K=%d a_mn=%d b_mn=%d bn=%d cg=%d best_ms=%.4f avg_ms=%.4f best_tflops=%.1f avg_tflops=%.1f\n",
           M, N, K, a_mn, b_mn, bn, cg, best, total / iters, fl / best * 1e-9, fl / (total / iters) * 1e-9);
    return 0;
  }

  int epi = argc > 9 ? atoi(argv[9]) : 0;
  int act = argc > 10 ? atoi(argv[10]) : 0;
  float drop_p = argc > 11 ? atof(argv[11]) : 0.f;
  int f32out = argc > 12 ? atoi(argv[12]) : 0;
  std::vector<float> hbias(N);
  for (auto& v : hbias) v = frand();
  float* dbias;
  CK(cudaMalloc(&dbias, N * 4));
  CK(cudaMemcpy(dbias, hbias.data(), N * 4, cudaMemcpyHostToDevice));
  int ldaux = round8(N) + 16;
  std::vector<__nv_bfloat16> haux((size_t)M * ldaux);
  for (auto& v : haux) v = __float2bfloat16(frand() * 0.5f + 0.5f);
  __nv_bfloat16* daux;
  CK(cudaMalloc(&daux, haux.size() * 2));
  CK(cudaMemcpy(daux, haux.data(), haux.size() * 2, cudaMemcpyHostToDevice));
  int ldo32 = (N + 3) / 4 * 4 + 4;
  float *dO32 = nullptr, *dPrev = nullptr;
  std::vector<float> hprev((size_t)M * N);
  g.epi = epi; g.act = act; g.drop_p = drop_p; g.seed = 0x1234abcd5678ull; g.drop_stream = 7;
  g.alpha = 0.5f;
  if (epi == EB_EPI_FWD) g.bias = dbias; else { g.aux = daux; g.ldaux = ldaux; }
  if (f32out) {
    CK(cudaMalloc(&dO32, (size_t)M * ldo32 * 4));
    std::vector<float> init((size_t)M * ldo32);
    for (int m = 0; m < M; ++m)
      for (int n = 0; n < N; ++n) { float v = frand(); init[(size_t)m * ldo32 + n] = v; hprev[(size_t)m * N + n] = v; }
    CK(cudaMemcpy(dO32, init.data(), init.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&dPrev, hprev.size() * 4));
    CK(cudaMemcpy(dPrev, hprev.data(), hprev.size() * 4, cudaMemcpyHostToDevice));
    g.out_f32 = dO32; g.ldo32 = ldo32; g.beta = 0.75f;
    if (f32out == 2) { g.out_bf16 = nullptr; }
  }
  float* dRef;
  CK(cudaMalloc(&dRef, (size_t)M * N * 4));
  dim3 blk(128), grd((N + 127) / 128, M);
  ref_gemm<<<grd, blk>>>(M, N, K, dA, lda, a_mn, dB, ldb, b_mn, g.alpha, epi, act, g.bias, daux, ldaux, drop_p, g.seed,
                         g.drop_stream, dRef, nullptr, 0.f);
  CK(cudaDeviceSynchronize());
  if (eb_gemm_bf16(&g, 0)) {
    printf("FAIL launch: %s\n", eb_last_error());
    return 1;
  }
  CK(cudaDeviceSynchronize());
  std::vector<float> href((size_t)M * N);
  CK(cudaMemcpy(href.data(), dRef, href.size() * 4, cudaMemcpyDeviceToHost));
  double max_err = 0, max_ref = 0;
  long bad = 0;
  if (g.out_bf16) {
    std::vector<__nv_bfloat16> hout((size_t)M * ldo);
    CK(cudaMemcpy(hout.data(), dO, hout.size() * 2, cudaMemcpyDeviceToHost));
    for (int m = 0; m < M; ++m)
      for (int n = 0; n < N; ++n) {
        double r = href[(size_t)m * N + n], o = __bfloat162float(hout[(size_t)m * ldo + n]);
        double err = fabs(r - o), tol = 0.02 * fabs(r) + 0.02 * sqrt((double)K) * 0.05 + 1e-2;
        if (err > tol || o != o) ++bad;
        max_err = fmax(max_err, err);
        max_ref = fmax(max_ref, fabs(r));
      }
    // padding columns must stay untouched (zero)
    for (int m = 0; m < M; ++m)
      for (int n = N; n < ldo; ++n)
        if (__bfloat162float(hout[(size_t)m * ldo + n]) != 0.f) ++bad;
  }
  if (g.out_f32) {
    std::vector<float> hout((size_t)M * ldo32);
    CK(cudaMemcpy(hout.data(), dO32, hout.size() * 4, cudaMemcpyDeviceToHost));
    for (int m = 0; m < M; ++m)
      for (int n = 0; n < N; ++n) {
        double r = href[(size_t)m * N + n] + 0.75 * hprev[(size_t)m * N + n], o = hout[(size_t)m * ldo32 + n];
        double err = fabs(r - o), tol = 1e-3 * fabs(r) + 1e-3 * sqrt((double)K) + 1e-3;
        if (err > tol || o != o) ++bad;
        max_err = fmax(max_err, err);
        max_ref = fmax(max_ref, fabs(r));
      }
  }
  printf("%s M=%d N=%d K=%d a_mn=%d b_mn=%d bn=%d cg=%d epi=%d act=%d drop=%.2f f32=%d max_err=%.5f max_ref=%.3f bad=%ld\n",
         bad == 0 ? "PASS" : "FAIL", M, N, K, a_mn, b_mn, bn, cg, epi, act, drop_p, f32out, max_err, max_ref, bad);
  return bad == 0 ? 0 : 1;
}
