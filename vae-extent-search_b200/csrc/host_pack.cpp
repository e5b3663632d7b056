// Host-side narrowing of fp32 feature rows to fp16 ahead of the PCIe copy (vaex_segsum_mlp_predict_host).
// The tcgen05 row encoder rounds every feature to fp16 (cvt.rn.satfinite) before its first MMA anyway, so doing
// that rounding on the host halves the bytes on the bus without changing one bit of the result: VCVTPS2PH with
// round-to-nearest-even after a clamp to +-65504 is the same function (NaN stays NaN; subnormal halves are kept,
// MXCSR.FTZ does not apply to this instruction).  Compiled by g++ with -mavx2 -mf16c; callers test
// host_pack_supported() first.
#include <immintrin.h>
#include <stddef.h>
#include <stdint.h>

namespace vaex {

bool host_pack_supported() { return __builtin_cpu_supports("avx2") && __builtin_cpu_supports("f16c"); }

static inline uint16_t narrow_one(float f) {
  __m128 v = _mm_set_ss(f);
  v = _mm_min_ss(_mm_set_ss(65504.f), _mm_max_ss(_mm_set_ss(-65504.f), v));   // NaN in the 2nd operand propagates
  return (uint16_t)_mm_extract_epi16(_mm_cvtps_ph(v, _MM_FROUND_TO_NEAREST_INT), 0);
}

static inline bool tiny_one(float f) {   // non-zero, but rounds to an fp16 zero (|f| <= 2^-25)
  uint32_t u;
  __builtin_memcpy(&u, &f, 4);
  u &= 0x7fffffffu;
  return u != 0 && u <= 0x33000000u;
}

// out[i] = fp16(in[i]), i < n.  `out` must be 2-byte aligned; the streaming-store body starts at the first
// 64-byte aligned output element (AVX-512 when the CPU has it, AVX2 otherwise and for the tail).  Returns true if some non-zero input became a zero: the "program with all-zero
// features" rule (mlp_model.py:842-844) is decided from the narrowed rows on the device, so the caller sends such a
// chunk as fp32 instead.
// AVX-512 body: 16 floats per iteration with 9 vector uops instead of 20 (one masked unsigned compare finds the
// "non-zero but rounds to zero" values: (|x| bits) - 1 < 0x33000000).
__attribute__((target("avx512f,avx512bw,avx512vl")))
static size_t pack_body_avx512(const float* in, uint16_t* out, size_t i, size_t n, bool* tiny) {
  const __m512 hi = _mm512_set1_ps(65504.f), lo = _mm512_set1_ps(-65504.f);
  const __m512i absm = _mm512_set1_epi32(0x7fffffff), one = _mm512_set1_epi32(1), lim = _mm512_set1_epi32(0x33000000);
  __mmask16 acc = 0;
  for (; i + 32 <= n; i += 32) {
    const __m512 xa = _mm512_loadu_ps(in + i), xb = _mm512_loadu_ps(in + i + 16);
    acc |= _mm512_cmplt_epu32_mask(_mm512_sub_epi32(_mm512_and_si512(_mm512_castps_si512(xa), absm), one), lim);
    acc |= _mm512_cmplt_epu32_mask(_mm512_sub_epi32(_mm512_and_si512(_mm512_castps_si512(xb), absm), one), lim);
    const __m256i ha = _mm512_cvtps_ph(_mm512_min_ps(hi, _mm512_max_ps(lo, xa)), _MM_FROUND_TO_NEAREST_INT);
    const __m256i hb = _mm512_cvtps_ph(_mm512_min_ps(hi, _mm512_max_ps(lo, xb)), _MM_FROUND_TO_NEAREST_INT);
    _mm512_stream_si512(reinterpret_cast<__m512i*>(out + i), _mm512_inserti64x4(_mm512_castsi256_si512(ha), hb, 1));
  }
  if (acc) *tiny = true;
  return i;
}

static const bool kHasAvx512 = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") && __builtin_cpu_supports("avx512vl");

bool host_pack_f16(const float* in, uint16_t* out, size_t n) {
  size_t i = 0;
  bool tiny = false;
  while (i < n && (reinterpret_cast<uintptr_t>(out + i) & 63)) { out[i] = narrow_one(in[i]); tiny |= tiny_one(in[i]); ++i; }
  if (kHasAvx512) i = pack_body_avx512(in, out, i, n, &tiny);
  const __m256 hi = _mm256_set1_ps(65504.f), lo = _mm256_set1_ps(-65504.f);
  const __m256i absm = _mm256_set1_epi32(0x7fffffff), lim = _mm256_set1_epi32(0x33000001), zero = _mm256_setzero_si256();
  __m256i acc = zero;
  for (; i + 16 <= n; i += 16) {
    const __m256 xa = _mm256_loadu_ps(in + i), xb = _mm256_loadu_ps(in + i + 8);
    const __m256i ta = _mm256_and_si256(_mm256_castps_si256(xa), absm), tb = _mm256_and_si256(_mm256_castps_si256(xb), absm);
    acc = _mm256_or_si256(acc, _mm256_and_si256(_mm256_cmpgt_epi32(lim, ta), _mm256_cmpgt_epi32(ta, zero)));
    acc = _mm256_or_si256(acc, _mm256_and_si256(_mm256_cmpgt_epi32(lim, tb), _mm256_cmpgt_epi32(tb, zero)));
    const __m256 a = _mm256_min_ps(hi, _mm256_max_ps(lo, xa));
    const __m256 b = _mm256_min_ps(hi, _mm256_max_ps(lo, xb));
    const __m128i ha = _mm256_cvtps_ph(a, _MM_FROUND_TO_NEAREST_INT), hb = _mm256_cvtps_ph(b, _MM_FROUND_TO_NEAREST_INT);
    _mm256_stream_si256(reinterpret_cast<__m256i*>(out + i), _mm256_set_m128i(hb, ha));
  }
  for (; i < n; ++i) { out[i] = narrow_one(in[i]); tiny |= tiny_one(in[i]); }
  _mm_sfence();
  return tiny || !_mm256_testz_si256(acc, acc);
}

}  // namespace vaex
