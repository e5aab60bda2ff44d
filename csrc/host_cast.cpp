// fp32 -> bf16 conversion of host rows (round to nearest even, NaN -> 0x7FFF: bit-identical to __float2bfloat16_rn on
// the device), written with explicit AVX-512 / AVX2 intrinsics and non-temporal stores, chosen at run time from what the
// CPU supports. Compiled by the host compiler alone (no nvcc front end in the way of the target attributes).
// Used by the input pipeline (host_stage.cu): features that are stored or streamed as bf16 cost half the PCIe bytes.
// No counterpart in the reference (pickled NumPy rows through Spark, elephas/utils/rdd_utils.py:12-27).
#include <immintrin.h>
#include <stdint.h>
#include <string.h>

namespace eb {

static inline uint16_t cast_one(uint32_t u) {
  const uint32_t rounded = u + 0x7FFFu + ((u >> 16) & 1u);
  return (u & 0x7FFFFFFFu) > 0x7F800000u ? static_cast<uint16_t>(0x7FFFu) : static_cast<uint16_t>(rounded >> 16);
}

static void convert_row_scalar(const uint32_t* s, uint16_t* d, int32_t cols) {
  for (int32_t c = 0; c < cols; ++c) d[c] = cast_one(s[c]);
}

__attribute__((target("avx512f,avx512bw"))) static void convert_row_avx512(const uint32_t* s, uint16_t* d, int32_t cols) {
  const __m512i bias = _mm512_set1_epi32(0x7FFF), one = _mm512_set1_epi32(1), absmask = _mm512_set1_epi32(0x7FFFFFFF);
  const __m512i inf = _mm512_set1_epi32(0x7F800000), qnan = _mm512_set1_epi32(0x7FFF);
  int32_t c = 0;
  // head: scalar until the destination is 64-byte aligned (non-temporal stores need it)
  while (c < cols && (reinterpret_cast<uintptr_t>(d + c) & 63)) { d[c] = cast_one(s[c]); ++c; }
  for (; c + 32 <= cols; c += 32) {
    const __m512i u0 = _mm512_loadu_si512(s + c), u1 = _mm512_loadu_si512(s + c + 16);
    __m512i r0 = _mm512_add_epi32(_mm512_add_epi32(u0, bias), _mm512_and_si512(_mm512_srli_epi32(u0, 16), one));
    __m512i r1 = _mm512_add_epi32(_mm512_add_epi32(u1, bias), _mm512_and_si512(_mm512_srli_epi32(u1, 16), one));
    r0 = _mm512_srli_epi32(r0, 16);
    r1 = _mm512_srli_epi32(r1, 16);
    const __mmask16 n0 = _mm512_cmpgt_epu32_mask(_mm512_and_si512(u0, absmask), inf);
    const __mmask16 n1 = _mm512_cmpgt_epu32_mask(_mm512_and_si512(u1, absmask), inf);
    r0 = _mm512_mask_mov_epi32(r0, n0, qnan);
    r1 = _mm512_mask_mov_epi32(r1, n1, qnan);
    const __m256i lo = _mm512_cvtepi32_epi16(r0), hi = _mm512_cvtepi32_epi16(r1);
    const __m512i out = _mm512_inserti64x4(_mm512_castsi256_si512(lo), hi, 1);
    _mm512_stream_si512(reinterpret_cast<__m512i*>(d + c), out);
  }
  for (; c < cols; ++c) d[c] = cast_one(s[c]);
}

__attribute__((target("avx2"))) static void convert_row_avx2(const uint32_t* s, uint16_t* d, int32_t cols) {
  const __m256i bias = _mm256_set1_epi32(0x7FFF), one = _mm256_set1_epi32(1), absmask = _mm256_set1_epi32(0x7FFFFFFF);
  const __m256i inf = _mm256_set1_epi32(0x7F800000), qnan = _mm256_set1_epi32(0x7FFF);
  int32_t c = 0;
  for (; c + 16 <= cols; c += 16) {
    __m256i r[2];
    for (int k = 0; k < 2; ++k) {
      const __m256i u = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(s + c + 8 * k));
      __m256i v = _mm256_add_epi32(_mm256_add_epi32(u, bias), _mm256_and_si256(_mm256_srli_epi32(u, 16), one));
      v = _mm256_srli_epi32(v, 16);
      const __m256i isnan = _mm256_cmpgt_epi32(_mm256_and_si256(u, absmask), inf);  // both operands non-negative
      r[k] = _mm256_blendv_epi8(v, qnan, isnan);
    }
    __m256i packed = _mm256_packus_epi32(r[0], r[1]);          // lanes interleaved: a0 b0 a1 b1 (64-bit quarters)
    packed = _mm256_permute4x64_epi64(packed, 0xD8);           // -> a0 a1 b0 b1
    _mm256_storeu_si256(reinterpret_cast<__m256i*>(d + c), packed);
  }
  for (; c < cols; ++c) d[c] = cast_one(s[c]);
}

typedef void (*RowFn)(const uint32_t*, uint16_t*, int32_t);

static RowFn pick() {
  __builtin_cpu_init();
  if (__builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw")) return convert_row_avx512;
  if (__builtin_cpu_supports("avx2")) return convert_row_avx2;
  return convert_row_scalar;
}

const char* host_cast_isa() {
  __builtin_cpu_init();
  if (__builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw")) return "avx512";
  if (__builtin_cpu_supports("avx2")) return "avx2";
  return "scalar";
}

// dst[r][0..cols) = bf16(src[r][0..cols)), dst[r][cols..ld_dst) = 0, for rows [r0, r1)
void convert_rows(const float* src, int64_t ld_src, uint16_t* dst, int64_t ld_dst, int64_t r0, int64_t r1, int32_t cols) {
  static const RowFn fn = pick();
  for (int64_t r = r0; r < r1; ++r) {
    uint16_t* d = dst + r * ld_dst;
    fn(reinterpret_cast<const uint32_t*>(src + r * ld_src), d, cols);
    if (ld_dst > cols) memset(d + cols, 0, static_cast<size_t>(ld_dst - cols) * 2);  // padding the TMA boxes may touch
  }
  _mm_sfence();  // non-temporal stores are visible before the caller hands the buffer to a DMA engine
}

}  // namespace eb
