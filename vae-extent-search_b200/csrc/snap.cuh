// postprocess + clamp of one decoded value (vae_sch_extent.ipynb:1374-1383,1457), shared by the
// stand-alone snap kernel and the fused decoder epilogue.
#pragma once
#include "common.cuh"

namespace vaex {

// torch: clamp(round(exp(v)).int(), min=1); round = half-to-even; out-of-range float->int32
// follows x86 cvttss2si (INT_MIN), which the clamp then maps to 1.
// unfused mul/add (no FMA contraction) keeps v bit-identical with torch's x*(std+1e-8)+mean.
__device__ __forceinline__ int32_t snap_one(float x, float mean, float scale, int mode) {
  float e;
  if ((mode & 3) == VAEX_SNAP_EXPM1) {
    const float v = __fadd_rn(__fmul_rn(x, scale), mean);
    e = (float)expm1((double)v);
  } else {
    const float v = __fadd_rn(__fmul_rn(x, __fadd_rn(scale, 1e-8f)), mean);
    e = (float)exp((double)v);
  }
  const float r = rintf(e);
  int32_t q = (r >= -2147483648.f && r < 2147483648.f) ? (int32_t)r : INT32_MIN;
  return q < 1 ? 1 : q;
}

}  // namespace vaex
