// Philox4x32-10 counter-based generator shared by the reparameterisation, the stand-alone dropout kernel and the
// dropout epilogue of the tcgen05 predictor chain (the same (seed, counter) -> the same draw on every path).
#pragma once
#include <stdint.h>

namespace vaex {

// ROUNDS = 10 is Random123's default; 7 is the smallest count its authors report as passing BigCrush ("Parallel random
// numbers: as easy as 1, 2, 3", SC'11, table 2).  The reparameterisation noise keeps 10; Dropout keep-bits -- whose
// generation is what bounds the MC-dropout passes of the tensor-core predictor -- use 7 (PhiloxDrop).
template <int ROUNDS>
struct PhiloxR {
  uint32_t k0, k1;
  __device__ PhiloxR(uint64_t seed) : k0((uint32_t)seed), k1((uint32_t)(seed >> 32)) {}
  __device__ uint4 operator()(uint64_t ctr_lo, uint64_t ctr_hi) const {
    uint32_t c0 = (uint32_t)ctr_lo, c1 = (uint32_t)(ctr_lo >> 32), c2 = (uint32_t)ctr_hi,
             c3 = (uint32_t)(ctr_hi >> 32);
    uint32_t a = k0, b = k1;
#pragma unroll
    for (int r = 0; r < ROUNDS; ++r) {
      const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
      const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
      c0 = hi1 ^ c1 ^ a; c1 = lo1; c2 = hi0 ^ c3 ^ b; c3 = lo0;
      a += 0x9E3779B9u; b += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
  }
};
using Philox = PhiloxR<10>;
using PhiloxDrop = PhiloxR<7>;
__device__ __forceinline__ float u01(uint32_t x) { return (x >> 8) * (1.0f / 16777216.0f) + (0.5f / 16777216.0f); }
// Dropout keeps an element iff its 16-bit draw is >= this threshold (drop probability round(p * 65536) / 65536)
__host__ __device__ __forceinline__ uint32_t dropout_threshold(float p) { return (uint32_t)(p * 65536.f + 0.5f); }


}  // namespace vaex
