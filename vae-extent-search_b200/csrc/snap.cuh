// postprocess + clamp of one decoded value (vae_sch_extent.ipynb:1374-1383,1457), shared by the
// stand-alone snap kernel and the fused decoder epilogue.
#pragma once
#include "common.cuh"

namespace vaex {

// torch: clamp(round(exp(v)).int(), min=1); round = half-to-even; out-of-range float->int32
// follows x86 cvttss2si (INT_MIN), which the clamp then maps to 1.
// unfused mul/add (no FMA contraction) keeps v bit-identical with torch's x*(std+1e-8)+mean.
//
// Only round(e) is kept, so the fp32 expf (<= 2 ulp) decides it whenever e is further than 8 ulp from a
// rounding boundary k + 0.5; the exact (double) path runs for the rare values that are not, and for
// e >= 2^20 where the integer is the float itself.  Bit-identical results, ~10x fewer FP64 instructions.
__device__ __forceinline__ float exp_for_round(float v, bool minus_one) {
  const float ef = expf(v);
  const float e = minus_one ? ef - 1.f : ef;
  const float fr = e - floorf(e);
  if (ef < 1048576.f && fabsf(fr - 0.5f) > fmaf(ef, 1e-6f, 1e-6f)) return e;
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

}  // namespace vaex
