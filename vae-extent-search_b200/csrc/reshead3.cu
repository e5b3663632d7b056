// Latent heads of the segment-VAE family in ONE launch on tcgen05 tensor cores:
//
//   y0 = norm(S)                              (eval-mode BatchNorm folded to a per-column affine, or identity)
//   y1 = relu(l0 y0 + b) + y0
//   y2 = relu(l1 y1 + b) + y1
//   [mean | logvar] = [fc_mean; fc_logvar] y2 + b
//   (SegmentVAE.encode, scripts/vae_experiments/models/vae.py:87-101; VAECostPredictor.encode, regression.py:82-101)
//
// Until now this was three single-layer launches of head_tc.cu, each reading and writing an fp32 [P, 256] matrix
// (1.8 ms per 10^6 programs for 0.39 MFLOP each).  Here a CTA pair keeps a tile of 256 programs on chip through all
// three layers: the fp16 operand (64 KB) is rewritten in place by every epilogue, the fp32 residual of a row lives in
// tensor memory (columns [0, 256): written by tcgen05.st, read back by the next epilogue -- residuals stay fp32, an
// fp16 skip path costs 5e-4 of score error), the accumulator in columns [256, 512), and the three 256 x 256 weight
// matrices stream through a 5-slot shared-memory ring as k-quarter pieces (cp.async.bulk.tensor, the scheme of
// chain_tc3.cu).  Layers are N = 256 MMAs.  S is read with fully coalesced 128-bit loads and transposed to the
// lane-per-row layout of TMEM / the operand through a per-warp staging area; the outputs leave the same way.
// Worker thread = (row, 64-column group = one k-quarter atom); 16 worker warps, warp 16 issues, warp 17 streams weights.
#include <cuda.h>

#include <algorithm>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "model.cuh"
#include "tc_ptx.cuh"

namespace vaex {

namespace r3 {
constexpr int TILE = 128, PAIR = 256, ATOM = 128 * 128, HID = 256;
constexpr int WORKER_WARPS = 16, ISSUER_WARP = 16, PRODUCER_WARP = 17, THREADS = 32 * 18;
constexpr int RING = 5, SLOT = ATOM, BOX_ROWS = 32;
constexpr int A_OFF = 0, RING_OFF = 4 * ATOM, STAGE_OFF = RING_OFF + RING * SLOT, BIAS_OFF = STAGE_OFF + WORKER_WARPS * 4096;
constexpr int BIAS_COUNT = 3 * HID + 2 * HID;       // three bias vectors, then the input affine: scale[256], shift[256]
constexpr int BAR_OFF = BIAS_OFF + BIAS_COUNT * 4;
enum { B_IN = 0, B_ACC = 4, B_DRAIN = 5, B_WFULL = 6, B_WEMPTY = 6 + RING, B_COUNT = 6 + 2 * RING };
constexpr int SMEM = BAR_OFF + 8 * B_COUNT + 16;
static_assert(SMEM <= 232448, "shared memory budget");

__host__ __device__ inline uint32_t sw_off(int u, int k) {   // (row u, column k < 64) inside one SWIZZLE_128B atom
  return (uint32_t)((u >> 3) * 1024 + (u & 7) * 128 + (((k >> 3) ^ (u & 7)) << 4) + (k & 7) * 2);
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
        "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]),
        "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]),
        "r"(v[30]), "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
        "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
}
}  // namespace r3

struct R3Params {
  const float* S;            // [n, 256] fp32 segment sums
  const float* nscale;       // [256] or null: y0 = S * nscale + nshift
  const float* nshift;
  const float* bias;         // b_l0[256] b_l1[256] b_fc[256] (zero padded)
  float* out0;               // mean [n, L]
  float* out1;               // logvar [n, L] or null
  int64_t n;
  int32_t T_tiles, L, n_fc;  // n_fc = 2 L (128 or 256): width of the stacked output layer
  int32_t w_row[3];          // first 128-byte image row of each layer (per rank)
  int32_t rank_rows;
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(r3::THREADS, 1)
latent_heads3_kernel(const R3Params p, const __grid_constant__ CUtensorMap tmap) {
  using namespace r3;
  extern __shared__ uint8_t smem[];                 // no static shared memory: the base is 1024-aligned (checked at pack time)
  const uint32_t sbase = ptx::smem_u32(smem);
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const int cluster = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + BAR_OFF);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + BAR_OFF + 8 * B_COUNT);
  const float* s_bias = reinterpret_cast<const float*>(smem + BIAS_OFF);
  const float* s_aff = s_bias + 3 * HID;
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto bar = [&](int i) { return bar0 + 8u * i; };

  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) ptx::mbar_init(bar(B_IN + i), 2 * 4);                  // the four warps of a column group x 2 CTAs
    ptx::mbar_init(bar(B_ACC), 1);
    ptx::mbar_init(bar(B_DRAIN), 2 * WORKER_WARPS);
    for (int i = 0; i < RING; ++i) { ptx::mbar_init(bar(B_WFULL + i), 1); ptx::mbar_init(bar(B_WEMPTY + i), 1); }
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc_2cta(ptx::smem_u32(tmem_slot), 512);
  {
    float* sb = reinterpret_cast<float*>(smem + BIAS_OFF);
    for (int i = threadIdx.x; i < 3 * HID; i += THREADS) sb[i] = p.bias[i];
    for (int i = threadIdx.x; i < HID; i += THREADS) {
      sb[3 * HID + i] = p.nscale ? p.nscale[i] : 1.f;
      sb[4 * HID + i] = p.nshift ? p.nshift[i] : 0.f;
    }
  }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp < WORKER_WARPS) {
#include "reshead3_worker.inc"
  } else if (warp == ISSUER_WARP) {
    if (rank == 0 && ptx::elect_one()) {
      // ===== MMA issuer: one thread of the leader CTA =====
      uint32_t in_phase = 0, drain_phase = 0, pc = 0;
      bool first = true;
      for (int t = cluster; t < p.T_tiles; t += n_clusters) {
        for (int l = 0; l < 3; ++l) {
          const uint32_t idesc = ptx::umma_idesc_f16(256, l < 2 ? HID : p.n_fc);
          for (int kq = 0; kq < 4; ++kq, ++pc) {
            const uint32_t slot = pc % RING;
            ptx::mbar_wait(bar(B_WFULL + slot), (pc / RING) & 1);
            ptx::mbar_wait(bar(B_IN + kq), (in_phase >> kq) & 1u);
            in_phase ^= 1u << kq;
            if (kq == 0 && !first) {                    // the previous layer's epilogue has read the accumulator
              ptx::mbar_wait(bar(B_DRAIN), drain_phase);
              drain_phase ^= 1u;
            }
            first = false;
            ptx::tc_fence_after();
            const uint64_t dA = ptx::umma_desc_sw128(sbase + A_OFF + kq * ATOM);
            const uint64_t dW = ptx::umma_desc_sw128(sbase + RING_OFF + slot * SLOT);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ptx::umma_f16_2cta(tmem + 256u, dA + (uint64_t)((k * 32) >> 4), dW + (uint64_t)((k * 32) >> 4), idesc, (kq > 0 || k > 0) ? 1u : 0u);
            ptx::umma_commit_2cta(bar(B_WEMPTY + slot), 3);
          }
          ptx::umma_commit_2cta(bar(B_ACC), 3);
        }
      }
    }
    __syncwarp();
  } else if (warp == PRODUCER_WARP) {
    if (ptx::elect_one()) {
      // ===== weight producer: one thread per CTA streams this CTA's half of every piece =====
      ptx::tma_prefetch_desc(&tmap);
      uint32_t pc = 0;
      for (int t = cluster; t < p.T_tiles; t += n_clusters) {
        for (int l = 0; l < 3; ++l) {
          const int rows = (l < 2 ? HID : p.n_fc) >> 1;
          for (int kq = 0; kq < 4; ++kq, ++pc) {
            const uint32_t slot = pc % RING;
            if (pc >= (uint32_t)RING) ptx::mbar_wait(bar(B_WEMPTY + slot), ((pc / RING) - 1) & 1);
            const uint32_t full_leader = ptx::mapa_rank(bar(B_WFULL + slot), 0);
            const int row0 = (int)rank * p.rank_rows + p.w_row[l] + kq * rows;
            for (int b = 0; b < rows; b += BOX_ROWS)
              ptx::tma_load_2d_2cta(sbase + RING_OFF + slot * SLOT + b * 128, &tmap, full_leader, 0, row0 + b);
            if (rank == 0) ptx::mbar_arrive_expect_tx(bar(B_WFULL + slot), 2u * (uint32_t)rows * 128u);
          }
        }
      }
    }
    __syncwarp();
  }

  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  if (warp == 1) ptx::tmem_dealloc_2cta(tmem, 512);
}

}  // namespace vaex

#include "reshead3_host.inc"
