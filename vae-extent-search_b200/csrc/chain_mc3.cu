// MC-dropout passes of the extent cost predictor on tcgen05 tensor cores, version 3.
//
//   z [n, L<=64] -> h = relu(W0 z + b0) / (1 - p) -> for t < T: cost_t = wd . relu(W1 (h * keep_t) + b1) + bd
//   out = mean_t cost_t, var_out = unbiased variance over t     (T == 1: out = cost_0)
//   (VAECostPredictor.mc_predict, utils/model.py:241-273; Dropout(0.1) after the first ReLU, :128-141; T = 20 at
//    vae_extent_search.py:1675)
//
// chain_tc.cu ran these passes with the layer-0 activations of a tile in registers (32 per thread) and N = 128 MMA
// pieces: 0.31 ms per pass and 10^6 latents.  Here layer 0's ACCUMULATOR simply stays in tensor memory for all T passes
// (TMEM columns [0, 256)): reading it back is cheap (tcgen05.ld moves 3.9 KB/clk per SM, tools/tmem_rate.cu), so every
// pass re-reads it, applies bias + ReLU + 1/(1-p) + the pass's keep bits, and writes the fp16 operand of the 256 x 256
// GEMM -- into one of TWO operand buffers, so that the mask epilogue of pass t + 1 runs under the MMAs of pass t.  The
// GEMM is 16 N = 256 MMAs (2 048 clk instead of 32 x 93.6 = 3 000) into TMEM columns [256, 512); its scalar-head
// epilogue (pass t) follows the mask epilogue of pass t + 1 in every worker.  Both weights are resident (W0 16 KB, W1
// 64 KB per CTA): with one tile per CTA pair there is room.  Worker thread = (row, 64-column group = one k-quarter atom).
// Keep bits: recorded masks uint8 [T][n][256], or the 16-bit Philox draws of dropout_kernel (elementwise.cu) for seed +
// t -- the same units are dropped on the fp32 SIMT path for a given seed.
#include <algorithm>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "model.cuh"
#include "philox.cuh"
#include "tc_ptx.cuh"

namespace vaex {

namespace m3 {
constexpr int TILE = 128, PAIR = 256, ATOM = 128 * 128, HID = 256;
constexpr int WORKER_WARPS = 16, WORKERS = 32 * WORKER_WARPS, THREADS = WORKERS + 32;
constexpr int W0_OFF = 0, W1_OFF = ATOM, A_OFF = 5 * ATOM, BIAS_OFF = 13 * ATOM;   // W0 | W1 (4 atoms) | A[2] (8 atoms) | b0 b1 wd bd
constexpr int BIAS_COUNT = 3 * HID + 4;
constexpr int BAR_OFF = BIAS_OFF + BIAS_COUNT * 4;          // multiple of 16
enum { B_Z = 0, B_ACC0, B_IN0, B_IN1, B_ACC1, B_DRAIN, B_COUNT };
constexpr int DOT_OFF = BAR_OFF + 64;                        // [2][3][128] partial dots
constexpr int SMEM = DOT_OFF + 2 * 3 * TILE * 4;
constexpr int SMEM_MAX = 232448;
static_assert(SMEM <= SMEM_MAX, "shared memory budget");

__host__ __device__ inline uint32_t sw_off(int u, int k) {   // (row u, column k < 64) inside one SWIZZLE_128B atom of 128 rows
  return (uint32_t)((u >> 3) * 1024 + (u & 7) * 128 + (((k >> 3) ^ (u & 7)) << 4) + (k & 7) * 2);
}
}  // namespace m3

struct Mc3Params {
  const float* X;            // [n, in_dim] fp32 latents
  const __half* w0;          // [2][ATOM / 2]: per-rank image of W0 (128 units x 64 k)
  const __half* w1;          // [2][4 ATOM / 2]: per-rank image of W1 (128 units x 256 k, four k-quarter atoms)
  const float* bias;         // b0[256] b1[256] wd[256] bd
  float* out;                // [n] mean (T == 1: the pass's cost)
  float* var_out;            // [n] unbiased variance or null
  const uint8_t* drop_mask;  // [T][n][256] keep masks or null
  uint64_t drop_seed;
  int64_t drop_chunk;        // rows per Philox chunk (the SIMT path re-seeds every VAE_CHUNK rows)
  int64_t n;
  float drop_p, drop_inv;
  int32_t T_tiles, in_dim, passes;
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(m3::THREADS, 1)
chain_mc3_kernel(const Mc3Params p) {
  using namespace m3;
  extern __shared__ uint8_t smem[];                 // no static shared memory: the base is 1024-aligned (checked at pack time)
  const uint32_t sbase = ptx::smem_u32(smem);
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const int cluster = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + BAR_OFF);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + BAR_OFF + 8 * B_COUNT);
  float* s_dot = reinterpret_cast<float*>(smem + DOT_OFF);
  const float* s_bias = reinterpret_cast<const float*>(smem + BIAS_OFF);
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto bar = [&](int i) { return bar0 + 8u * i; };

  if (threadIdx.x == 0) {
    ptx::mbar_init(bar(B_Z), 2 * WORKER_WARPS);
    ptx::mbar_init(bar(B_ACC0), 1);
    ptx::mbar_init(bar(B_IN0), 2 * WORKER_WARPS);
    ptx::mbar_init(bar(B_IN1), 2 * WORKER_WARPS);
    ptx::mbar_init(bar(B_ACC1), 1);
    ptx::mbar_init(bar(B_DRAIN), 2 * WORKER_WARPS);
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc_2cta(ptx::smem_u32(tmem_slot), 512);
  {
    const uint4* g0 = reinterpret_cast<const uint4*>(p.w0) + (size_t)rank * (ATOM / 16);
    const uint4* g1 = reinterpret_cast<const uint4*>(p.w1) + (size_t)rank * (4 * ATOM / 16);
    uint4* s0 = reinterpret_cast<uint4*>(smem + W0_OFF);
    uint4* s1 = reinterpret_cast<uint4*>(smem + W1_OFF);
    for (int i = threadIdx.x; i < ATOM / 16; i += THREADS) s0[i] = __ldg(g0 + i);
    for (int i = threadIdx.x; i < 4 * ATOM / 16; i += THREADS) s1[i] = __ldg(g1 + i);
    float* sb = reinterpret_cast<float*>(smem + BIAS_OFF);
    for (int i = threadIdx.x; i < BIAS_COUNT; i += THREADS) sb[i] = p.bias[i];
    uint4* a = reinterpret_cast<uint4*>(smem + A_OFF);      // K padding of the z atom stays 0
    for (int i = threadIdx.x; i < 8 * ATOM / 16; i += THREADS) a[i] = make_uint4(0, 0, 0, 0);
  }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp < WORKER_WARPS) {
#include "chain_mc3_worker.inc"
  } else if (rank == 0 && ptx::elect_one()) {
    // ===== MMA issuer: one thread of the leader CTA =====
    const uint32_t idesc = ptx::umma_idesc_f16(256, 256);
    const uint64_t dW0 = ptx::umma_desc_sw128(sbase + W0_OFF);
    uint32_t ph_z = 0, ph_in = 0, ph_drain = 0;            // parity of the next wait (bit b of ph_in: buffer b)
    bool first = true;
    for (int t = cluster; t < p.T_tiles; t += n_clusters) {
      ptx::mbar_wait(bar(B_Z), ph_z);
      ph_z ^= 1u;
      ptx::tc_fence_after();
      {
        const uint64_t dA = ptx::umma_desc_sw128(sbase + A_OFF + 7 * ATOM);        // z: atom 3 of operand buffer 1
#pragma unroll
        for (int k = 0; k < 4; ++k) ptx::umma_f16_2cta(tmem, dA + (uint64_t)((k * 32) >> 4), dW0 + (uint64_t)((k * 32) >> 4), idesc, k > 0);
        ptx::umma_commit_2cta(bar(B_ACC0), 3);
      }
      for (int pass = 0; pass < p.passes; ++pass) {
        const int b = pass & 1;
        ptx::mbar_wait(bar(B_IN0 + b), (ph_in >> b) & 1u);
        ph_in ^= 1u << b;
        if (!first) {                                        // the previous pass's head has read the accumulator
          ptx::mbar_wait(bar(B_DRAIN), ph_drain);
          ph_drain ^= 1u;
        }
        first = false;
        ptx::tc_fence_after();
        const uint32_t a_base = sbase + A_OFF + b * 4 * ATOM;
#pragma unroll
        for (int kq = 0; kq < 4; ++kq) {
          const uint64_t dA = ptx::umma_desc_sw128(a_base + kq * ATOM);
          const uint64_t dW = ptx::umma_desc_sw128(sbase + W1_OFF + kq * ATOM);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            ptx::umma_f16_2cta(tmem + 256u, dA + (uint64_t)((k * 32) >> 4), dW + (uint64_t)((k * 32) >> 4), idesc, (kq > 0 || k > 0) ? 1u : 0u);
        }
        ptx::umma_commit_2cta(bar(B_ACC1), 3);
      }
    }
  }
  __syncwarp();

  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  if (warp == 1) ptx::tmem_dealloc_2cta(tmem, 512);
}

}  // namespace vaex

#include "chain_mc3_host.inc"
