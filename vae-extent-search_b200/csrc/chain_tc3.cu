// Fused MLP chains of the extent-VAE family on tcgen05 tensor cores, version 3: TWO tiles in flight per CTA pair,
// weights streamed from L2 through a shared-memory ring by TMA, 256-wide layers as N = 256 MMAs.
//
//   decoder   z [n, L<=128] -> relu(W0) -> relu(W1) -> relu(W2) -> W3 -> x [n, D<=64]  (+ snap to int32 extents)
//   encoder   x [n, D<=128] -> relu(W0) -> relu(W1) -> relu(W2) -> [fc_mu; fc_logvar] -> mu, logvar [n, 64]
//   predictor z [n, L<=128] -> relu(W0) -> relu(W1) [-> relu(W2)] -> w . h + b -> cost [n]
//   (VAE_feature_head.decode / encode, utils/model.py:31-39,55-64; VAECostPredictor.predict_cost, :128-141,168-170;
//    the segment family's BN-folded cost_predictor, scripts/vae_experiments/models/vae.py)
//
// Why a third version (profiles/r02_chain_cycles.md): with ONE tile per CTA pair (chain_tc.cu) every layer pays the
// hand-off latencies commit -> worker wake-up -> epilogue -> publish -> issuer wake-up with an empty tensor pipe, and
// the resident weights (160 KB) leave no room for a second operand; and an M256 N128 MMA issues every 93.6 clk where an
// M256 N256 MMA issues every 128.1 clk (tools/mma_rate.cu), so the two N = 128 halves that made the in-tile pipeline
// possible cost 1.46 x the tensor time of one N = 256 instruction.  Here:
//   * a CTA pair works on tiles X and Y (256 rows each) at once: operands A_X, A_Y (2 x 64 KB), accumulators in TMEM
//     columns [0, 256) and [256, 512); while the eight worker warps of X run an epilogue, the tensor pipe runs Y's MMAs;
//   * weights are NOT resident: each layer is cut into k-quarter pieces ([N/2 rows x 64 halves] per CTA, <= 16 KB, the
//     exact SWIZZLE_128B operand image), and one producer thread per CTA streams them through a 5-slot ring with
//     cp.async.bulk.tensor (cta_group::2: both CTAs' boxes complete on the leader's mbarrier); tcgen05.commit releases
//     a slot.  Issue order per layer: all of X's pieces (k0 k2 k1 k3), commit ACC_X, all of Y's, commit ACC_Y -- every
//     tile gets its own copy of a piece (160 KB per tile per CTA from L2).  Sharing a piece between X and Y halves that
//     traffic but couples the tiles: each then waits for the other's epilogue inside every layer (measured slower);
//   * worker thread = (row, half of the 256 columns): four 32-column TMEM loads, bias + ReLU + fp16 pack into two
//     k-quarter atoms of the next layer's operand, one IN[kq] arrival per atom; DRAIN after the last TMEM load lets
//     the next layer overwrite the accumulator;
//   * no local memory: with 227 KB of shared memory carved out the L1 is a few KB and every local access is an L2
//     round trip (a dynamically indexed phase array and 172 B of spills made the first version 1.5 x slower than v2).
// MC-dropout passes: chain_mc3.cu (one tile per CTA pair, layer 0's accumulator kept in TMEM across the passes).
#include <cuda.h>

#include <algorithm>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "model.cuh"
#include "snap.cuh"
#include "tc_ptx.cuh"

namespace vaex {

namespace c3 {
constexpr int TILE = 128, PAIR = 256, ATOM = 128 * 128, HID = 256;
constexpr int GROUP_WARPS = 8, GROUP = 32 * GROUP_WARPS, WORKER_WARPS = 16;
constexpr int ISSUER_WARP = 16, PRODUCER_WARP = 17, THREADS = 32 * 18;
constexpr int MAX_LAYERS = 6;
constexpr int RING = 5, SLOT = ATOM;                    // five 16 KB slots
constexpr int BOX_ROWS = 32;                            // one TMA box = 32 image rows of 128 B
constexpr int A_OFF = 0, RING_OFF = 2 * 4 * ATOM, FIXED = RING_OFF + RING * SLOT;   // 212 992 B
constexpr int SMEM_MAX = 232448;
enum { MODE_VECTOR = 0, MODE_DOT = 1 };
// barriers: IN[tile][kq] (8), ACC[tile] (2), DRAIN[tile] (2), W_FULL[RING], W_EMPTY[RING]
enum { B_IN = 0, B_ACC = 8, B_DRAIN = 10, B_WFULL = 12, B_WEMPTY = 12 + RING, B_COUNT = 12 + 2 * RING };

__host__ __device__ inline uint32_t sw_off(int u, int k) {   // (row u, column k < 64) inside one SWIZZLE_128B atom
  return (uint32_t)((u >> 3) * 1024 + (u & 7) * 128 + (((k >> 3) ^ (u & 7)) << 4) + (k & 7) * 2);
}
// ring order of a layer's k-quarters: the two quarters an epilogue delivers first come first
__host__ __device__ inline int kq_of(int nq, int i) { return nq == 4 ? ((i & 1) << 1) | (i >> 1) : i; }
}  // namespace c3

struct Chain3Params {
  const float* X;            // [n, in_dim] fp32
  const float* bias;         // fp32 biases of the MMA layers, concatenated (zero padded to N) [+ MODE_DOT: wd[256], bd]
  float* out;                // MODE_VECTOR: [n, out_dim] fp32 (nullable); MODE_DOT: [n]
  float* out2;               // two stacked heads (N = 2 * out_dim): columns [out_dim, 2 out_dim)
  int32_t* extents;          // MODE_VECTOR: [n, out_dim] int32 (nullable): fused snap
  const float* snap_mean;    // [out_dim]
  const float* snap_scale;
  int64_t n;
  int32_t T, in_dim, out_dim, mode, snap_mode, n_mma;
  int32_t K[c3::MAX_LAYERS], N[c3::MAX_LAYERS], b_off[c3::MAX_LAYERS];
  int32_t w_row[c3::MAX_LAYERS];      // first 128-byte image row of the layer (per rank); pieces follow in ring order
  int32_t rank_rows;                  // image rows per rank
  int32_t bias_off, bias_count, wd_off, bar_off;
};

// -DCHAIN3_PROF: cycle counters of cluster 0 (issuer waits per kind, worker waits / epilogues of tile X's warp 0),
// accumulated in shared memory and dumped to c3_prof at exit; read with vaex_debug_chain3_prof (tools/prof_chain3.py)
#ifdef CHAIN3_PROF
__device__ unsigned long long c3_prof[64];
#define C3P(i, v) do { if (blockIdx.x == 0) sp[i] += (unsigned long long)(v); } while (0)
#define C3T() clock64()
#else
#define C3P(i, v) do { (void)(v); } while (0)
#define C3T() 0ll
#endif

template <int KZ>     // atoms of the layer-0 operand: in_dim <= 64 KZ
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(c3::THREADS, 1)
chain3_kernel(const Chain3Params p, const __grid_constant__ CUtensorMap tmap) {
  using namespace c3;
  extern __shared__ uint8_t smem[];                 // no static shared memory: the base is 1024-aligned (checked at pack time)
  const uint32_t sbase = ptx::smem_u32(smem);
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const int cluster = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const int T2 = (p.T + 1) >> 1;                     // tile pairs (X = 2 d, Y = 2 d + 1)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.bar_off);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + p.bar_off + 8 * B_COUNT);
  float* s_dot = reinterpret_cast<float*>(smem + p.bar_off + 8 * B_COUNT + 16);   // MODE_DOT: [2][128] partial dots
  const float* s_bias = reinterpret_cast<const float*>(smem + p.bias_off);
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto bar = [&](int i) { return bar0 + 8u * i; };
#ifdef CHAIN3_PROF
  unsigned long long* sp = reinterpret_cast<unsigned long long*>(smem + p.bar_off + 8 * B_COUNT + 16 + 2 * TILE * 4);
  if (threadIdx.x < 64) sp[threadIdx.x] = 0;
#endif

  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) ptx::mbar_init(bar(B_IN + i), 2 * GROUP_WARPS);        // the tile's worker warps x 2 CTAs
    for (int i = 0; i < 2; ++i) ptx::mbar_init(bar(B_ACC + i), 1);
    for (int i = 0; i < 2; ++i) ptx::mbar_init(bar(B_DRAIN + i), 2 * GROUP_WARPS);
    for (int i = 0; i < RING; ++i) { ptx::mbar_init(bar(B_WFULL + i), 1); ptx::mbar_init(bar(B_WEMPTY + i), 1); }
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc_2cta(ptx::smem_u32(tmem_slot), 512);
  {
    float* sb = reinterpret_cast<float*>(smem + p.bias_off);
    for (int i = threadIdx.x; i < p.bias_count; i += THREADS) sb[i] = p.bias[i];
  }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp < WORKER_WARPS) {
#include "chain_tc3_worker.inc"
  } else if (warp == ISSUER_WARP) {
    if (rank == 0 && ptx::elect_one()) {
      // ===== MMA issuer: one thread of the leader CTA =====
      // phase bits in registers: with 227 KB of shared memory carved out the L1 is a few KB, and a dynamically indexed
      // local array costs an L2 round trip per access (that alone made the first version 1.5 x slower than v2)
      uint32_t in_phase = 0, drain_phase = 0;           // bit 4 g + kq / bit g: parity of the next wait
      uint32_t pc = 0;                                  // pieces consumed so far (ring position)
      bool first = true;
      const long long tt0 = C3T();
      for (int d = cluster; d < T2; d += n_clusters) {
        C3P(1, 1);
        for (int l = 0; l < p.n_mma; ++l) {
          const int K = p.K[l], N = p.N[l], nq = K >> 6;
          const int h0 = (nq + 1) >> 1;                 // pieces of the first half
          const uint32_t idesc = ptx::umma_idesc_f16(256, N);
          // tile-major: all of X's pieces, then all of Y's (each tile gets its own copy of a piece through the ring), so
          // that X's epilogue runs under Y's MMAs and the other way round.  Sharing a piece between the tiles
          // (X(k0) X(k2) Y(k0) Y(k2) X(k1) X(k3) Y(k1) Y(k3)) halves the L2 traffic but couples the tiles: each then waits
          // for the other's epilogue inside every layer (measured: 0.55 ms per 10^6 decoder rows against 0.37 for v2).
          (void)h0;
          for (int g = 0; g < 2; ++g) {
            if (!(first && l == 0)) {                   // the previous layer's epilogue has read the accumulator
              const long long c0 = C3T();
              ptx::mbar_wait(bar(B_DRAIN + g), (drain_phase >> g) & 1u);
              drain_phase ^= 1u << g;
              C3P(2 + g, C3T() - c0);
            }
            for (int i = 0; i < nq; ++i, ++pc) {
              const uint32_t slot = pc % RING;
              const long long c0 = C3T();
              ptx::mbar_wait(bar(B_WFULL + slot), (pc / RING) & 1);
              const long long c1 = C3T();
              const int kq = kq_of(nq, i);
              ptx::mbar_wait(bar(B_IN + 4 * g + kq), (in_phase >> (4 * g + kq)) & 1u);
              in_phase ^= 1u << (4 * g + kq);
              C3P(4, c1 - c0);
              C3P(5 + g, C3T() - c1);
              C3P(8 + l, C3T() - c0);
              ptx::tc_fence_after();
              const uint64_t dA = ptx::umma_desc_sw128(sbase + A_OFF + g * 4 * ATOM + kq * ATOM);
              const uint64_t dW = ptx::umma_desc_sw128(sbase + RING_OFF + slot * SLOT);
#pragma unroll
              for (int s = 0; s < 4; ++s)
                ptx::umma_f16_2cta(tmem + 256u * g, dA + (uint64_t)((s * 32) >> 4), dW + (uint64_t)((s * 32) >> 4), idesc, (i > 0 || s > 0) ? 1u : 0u);
              ptx::umma_commit_2cta(bar(B_WEMPTY + slot), 3);
            }
            ptx::umma_commit_2cta(bar(B_ACC + g), 3);
          }
        }
        first = false;
      }
      C3P(0, C3T() - tt0);
    }
    __syncwarp();
  } else if (warp == PRODUCER_WARP) {
    if (ptx::elect_one()) {
      // ===== weight producer: one thread per CTA streams this CTA's half of every piece =====
      ptx::tma_prefetch_desc(&tmap);
      uint32_t pc = 0;
      for (int d = cluster; d < T2; d += n_clusters) {
        for (int l = 0; l < p.n_mma; ++l) {
          const int nq = p.K[l] >> 6;
          const int rows = p.N[l] >> 1;                 // image rows of one piece (per CTA)
          for (int gi = 0; gi < 2 * nq; ++gi, ++pc) {
            const int i = gi >= nq ? gi - nq : gi;     // X's copy of every piece, then Y's
            const uint32_t slot = pc % RING;
            if (pc >= (uint32_t)RING) ptx::mbar_wait(bar(B_WEMPTY + slot), ((pc / RING) - 1) & 1);
            const uint32_t full_leader = ptx::mapa_rank(bar(B_WFULL + slot), 0);
            const int row0 = (int)rank * p.rank_rows + p.w_row[l] + i * rows;
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
#ifdef CHAIN3_PROF
  if (blockIdx.x == 0 && threadIdx.x < 64) c3_prof[threadIdx.x] = sp[threadIdx.x];
#endif
}

}  // namespace vaex

#include "chain_tc3_host.inc"
