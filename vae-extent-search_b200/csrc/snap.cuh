// postprocess + clamp of one decoded value (vae_sch_extent.ipynb:1374-1383,1457), shared by the
// stand-alone snap kernel and the fused decoder epilogue.
#pragma once
#include "common.cuh"

namespace vaex {

// torch: clamp(round(exp(v)).int(), min=1); round = half-to-even; out-of-range float->int32
// follows x86 cvttss2si (INT_MIN), which the clamp then maps to 1.
// unfused mul/add (no FMA contraction) keeps v bit-identical with torch's x*(std+1e-8)+mean.
//
// Only round(e) is kept, so a fast exp decides it whenever e is further from a rounding boundary k + 0.5 than the fast
// exp can be wrong: __expf (ex2.approx of v * log2 e, two instructions) is within 2 + floor(1.173 |v|) ulp (CUDA
// programming guide, intrinsic error bounds), i.e. <= 18 ulp <= 2.2e-6 relative for e < 2^20, and the band below is
// 3e-6 e + 1e-6.  The exact (double) path runs for the rare values inside the band (~6e-6 e of them) and for e >= 2^20,
// where the integer is the float itself.  Bit-identical with the correctly rounded result on every input
// (tests: every boundary k + 0.5 up to 2e5, +-3 ulp, both modes); the accurate expf + fp64 version cost the decoder's
// output epilogue 0.23 ms per 10^6 latents.
__device__ __forceinline__ float exp_for_round(float v, bool minus_one) {
  const float ef = __expf(v);
  const float e = minus_one ? ef - 1.f : ef;
  const float fr = e - floorf(e);
  if (ef < 1048576.f && fabsf(fr - 0.5f) > fmaf(ef, 3e-6f, 1e-6f)) return e;
  return minus_one ? (float)expm1((double)v) : (float)exp((double)v);
}

__device__ __forceinline__ int32_t snap_one(float x, float mean, float scale, int mode) {
  float e;
  if ((mode & 3) == VAEX_SNAP_EXPM1) {
    const float v = __fadd_rn(__fmul_rn(x, scale), mean);
    e = exp_for_round(v, true);
  } else {
    const float v = __fadd_rn(__fmul_rn(x, __fadd_rn(scale, 1e-8f)), mean);
    e = exp_for_round(v, false);
  }
  const float r = rintf(e);
  int32_t q = (r >= -2147483648.f && r < 2147483648.f) ? (int32_t)r : INT32_MIN;
  return q < 1 ? 1 : q;
}

// Eight values at once for the decoder's output epilogue: the fast decisions of all eight are computed without a branch
// between them (their latencies overlap; with a per-value branch the eight dependent chains of ld -> mul -> add -> ex2 ->
// floor -> compare ran one after the other and cost the fused kernel 0.18 ms per 10^6 latents), then ONE branch handles
// the rare values that need the exact path.  Same results as snap_one, bit for bit.  mean / scale: 16-byte aligned.
__device__ __forceinline__ void snap_vec8(const float (&x)[8], const float* __restrict__ mean, const float* __restrict__ scale,
                                          int mode, int (&q)[8]) {
  const bool m1 = (mode & 3) == VAEX_SNAP_EXPM1;
  const float4 mu0 = __ldg(reinterpret_cast<const float4*>(mean)), mu1 = __ldg(reinterpret_cast<const float4*>(mean) + 1);
  const float4 sc0 = __ldg(reinterpret_cast<const float4*>(scale)), sc1 = __ldg(reinterpret_cast<const float4*>(scale) + 1);
  const float mu[8] = {mu0.x, mu0.y, mu0.z, mu0.w, mu1.x, mu1.y, mu1.z, mu1.w};
  const float sc[8] = {sc0.x, sc0.y, sc0.z, sc0.w, sc1.x, sc1.y, sc1.z, sc1.w};
  float v[8], e[8];
  uint32_t slow = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    v[i] = m1 ? __fadd_rn(__fmul_rn(x[i], sc[i]), mu[i]) : __fadd_rn(__fmul_rn(x[i], __fadd_rn(sc[i], 1e-8f)), mu[i]);
    const float ef = __expf(v[i]);
    e[i] = m1 ? ef - 1.f : ef;
    const float fr = e[i] - floorf(e[i]);
    slow |= (uint32_t)!(ef < 1048576.f && fabsf(fr - 0.5f) > fmaf(ef, 3e-6f, 1e-6f)) << i;
  }
  if (slow) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if ((slow >> i) & 1u) e[i] = m1 ? (float)expm1((double)v[i]) : (float)exp((double)v[i]);
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float r = rintf(e[i]);
    const int32_t t = (r >= -2147483648.f && r < 2147483648.f) ? (int32_t)r : INT32_MIN;
    q[i] = t < 1 ? 1 : t;
  }
}

}  // namespace vaex
