// Fused SegmentSumMLP row encoder + per-program segment sum on tcgen05 tensor cores.
//
//   S[p, :] = sum_{r in program p} relu(W1 relu(W0 x_r + b0) + b1)
//   (SegmentSumMLPModule.forward, python/tvm/auto_scheduler/cost_model/mlp_model.py:189-208;
//    SegmentVAE.segment_sum, scripts/vae_experiments/models/vae.py:68-85)
//
// One persistent CTA pair (cluster of 2, tcgen05 cta_group::2) per two SMs; a pair tile is 256
// feature rows (128 per CTA).  fp16 operands (11-bit significand), fp32 accumulation in TMEM.
//
//   layer 0:  D0[256 rows x 256 hid]  = X[256 x K0] * W0^T        A = X tile (this CTA's 128 rows)
//             (TMEM cols 0..255)                                    B = W0 (this CTA's 128 hidden units)
//   epi 0  :  TMEM -> +b0, ReLU, fp16 -> A1 smem (K-major, 128B swizzle)       lane = row
//   layer 1:  D1[256 hid x 256 rows] = W1[256 x 256] * A1^T       A = W1 (this CTA's 128 hidden units)
//             (TMEM cols 256..511)                                  B = A1 (this CTA's 128 rows)
//   epi 1  :  TMEM -> +b1, ReLU, running sum over *columns* (= rows of the tile) cut at the
//             program boundaries -> S.  lane = hidden unit, so the segment sum needs no
//             shuffles or shared memory and boundaries are uniform across the CTA.
//
// Layer 1 is computed transposed on purpose: it puts the segment dimension on TMEM columns.
// Weights stay resident in shared memory (each CTA holds half of W0 and half of W1: 112 KB),
// X is streamed with 128-bit loads, converted to fp16 and written straight into the swizzled
// operand layout (the fp32 -> fp16 cast rules out a direct TMA load of X).
//
// Programs cut by a tile boundary receive two partial sums through atomicAdd on zeroed rows
// (two addends: order-independent, so results are deterministic for S_i <= 256).
#include <algorithm>
#include <vector>

#include "common.cuh"
#include "model.cuh"
#include "tc_ptx.cuh"

namespace vaex {

namespace tc {
constexpr int HID = 256;
constexpr int TILE = 128;            // rows per CTA per tile
constexpr int PAIR = 256;            // rows per CTA pair per tile
constexpr int K0P = 192;             // layer-0 K padded to 3 swizzle atoms of 64 fp16
constexpr int ATOM = 128 * 128;      // bytes of one [128 rows x 64 fp16] swizzle atom
constexpr int W0_OFF = 0, W0_BYTES = 3 * ATOM;
constexpr int W1_OFF = W0_OFF + W0_BYTES, W1_BYTES = 4 * ATOM;
constexpr int A0_OFF = W1_OFF + W1_BYTES, A0_BYTES = 3 * ATOM;
constexpr int A1_OFF = A0_OFF + A0_BYTES, A1_BYTES = 4 * ATOM;
constexpr int SMEM_TILES = A1_OFF + A1_BYTES;          // 229376
constexpr int SMEM_DYN = SMEM_TILES + 1024;            // + slack to align the base to 1024 B
constexpr int LOAD_WARPS = 4, MMA_WARP = 4, EPI_WARP0 = 5, EPI_WARPS = 4;
constexpr int THREADS = 32 * (LOAD_WARPS + 1 + EPI_WARPS);   // 288
enum { B_A0_FULL = 0, B_A0_EMPTY, B_ACC0_FULL, B_A1_FULL, B_ACC1_FULL, B_ACC1_EMPTY, B_COUNT };

__host__ __device__ inline uint32_t sw128_off(int row, int k) {
  return (uint32_t)((k >> 6) * ATOM + (row >> 3) * 1024 + (row & 7) * 128 + ((((k & 63) >> 3) ^ (row & 7)) << 4) +
                    (k & 7) * 2);
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
  a = fminf(fmaxf(a, -65504.f), 65504.f);
  b = fminf(fmaxf(b, -65504.f), 65504.f);
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
}  // namespace tc

struct TcParams {
  const float* feats;          // [n_rows, F]
  const int32_t* off;          // [P + 1]
  const int32_t* first_prog;   // [T] program holding the first row of pair tile t
  const __half* w0;            // [2][W0_BYTES/2] swizzled images per CTA rank
  const __half* w1;            // [2][W1_BYTES/2]
  const float* b0;             // [256]
  const float* b1;             // [256]
  float* out;                  // [P, 256]
  uint32_t* rowmask;           // [T * 8] bit r: feature row has a non-zero
  int64_t n_rows;
  int32_t P, T, F, k0_steps;
};

// ---------------------------------------------------------------------------------------------
template <int VEC>
__device__ __forceinline__ void load_convert_tile(const TcParams& p, uint8_t* A0, int64_t row0, int warp, int lane,
                                                  uint32_t& nzmask) {
  using namespace tc;
  constexpr int RB = 8;                       // rows in flight per lane
  const int U = p.F / VEC;                    // vector units per row
  constexpr int UU = (K0P / VEC + 31) / 32;   // units per lane per row
  nzmask = 0;
  for (int rb = 0; rb < 32; rb += RB) {
    float v[RB][UU][VEC];
#pragma unroll
    for (int i = 0; i < RB; ++i) {
      const int64_t row = row0 + warp * 32 + rb + i;
      const bool ok = row < p.n_rows;
      const float* src = p.feats + row * p.F;
#pragma unroll
      for (int uu = 0; uu < UU; ++uu) {
        const int u = lane + 32 * uu;
        if (ok && u < U) {
          if (VEC == 4) { float4 t = ptx::ldg_stream_f4(src + u * 4); v[i][uu][0] = t.x; v[i][uu][1] = t.y; v[i][uu][2 % VEC] = t.z; v[i][uu][3 % VEC] = t.w; }
          else if (VEC == 2) { float2 t = ptx::ldg_stream_f2(src + u * 2); v[i][uu][0] = t.x; v[i][uu][1 % VEC] = t.y; }
          else { v[i][uu][0] = __ldg(src + u); }
        } else {
#pragma unroll
          for (int e = 0; e < VEC; ++e) v[i][uu][e] = 0.f;
        }
      }
    }
#pragma unroll
    for (int i = 0; i < RB; ++i) {
      const int r = warp * 32 + rb + i;       // row inside this CTA's half tile
      bool nz = false;
#pragma unroll
      for (int uu = 0; uu < UU; ++uu) {
        const int u = lane + 32 * uu;
#pragma unroll
        for (int e = 0; e < VEC; ++e) nz |= (v[i][uu][e] != 0.f);
        if (u < U) {
          uint8_t* dst = A0 + sw128_off(r, u * VEC);
          if (VEC == 4) *reinterpret_cast<uint2*>(dst) = make_uint2(pack_h2(v[i][uu][0], v[i][uu][1]), pack_h2(v[i][uu][2 % VEC], v[i][uu][3 % VEC]));
          else if (VEC == 2) *reinterpret_cast<uint32_t*>(dst) = pack_h2(v[i][uu][0], v[i][uu][1 % VEC]);
          else *reinterpret_cast<__half*>(dst) = __float2half_rn(fminf(fmaxf(v[i][uu][0], -65504.f), 65504.f));
        }
      }
      if (__any_sync(0xffffffffu, nz)) nzmask |= 1u << (rb + i);
    }
  }
}

// ---------------------------------------------------------------------------------------------
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(tc::THREADS, 1)
segsum_tc_kernel(const TcParams p) {
  using namespace tc;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[B_COUNT];
  __shared__ uint32_t tmem_base_slot;
  __shared__ float b0s[HID];

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t sbase = ptx::smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const int cluster = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const uint32_t bar0 = ptx::smem_u32(&bars[0]);
  auto bar = [&](int i) { return bar0 + 8u * i; };

  // ---- prologue -------------------------------------------------------------------------------
  if (warp == MMA_WARP && lane == 0) {
    ptx::mbar_init(bar(B_A0_FULL), 2 * LOAD_WARPS);
    ptx::mbar_init(bar(B_A0_EMPTY), 1);
    ptx::mbar_init(bar(B_ACC0_FULL), 1);
    ptx::mbar_init(bar(B_A1_FULL), 2 * EPI_WARPS);
    ptx::mbar_init(bar(B_ACC1_FULL), 1);
    ptx::mbar_init(bar(B_ACC1_EMPTY), 2 * EPI_WARPS);
    ptx::fence_mbar_init();
  }
  if (warp == EPI_WARP0) ptx::tmem_alloc_2cta(ptx::smem_u32(&tmem_base_slot), 512);
  {
    // resident weights: already in the exact swizzled byte layout -> straight 16-byte copies
    const uint4* g0 = reinterpret_cast<const uint4*>(p.w0) + (size_t)rank * (W0_BYTES / 16);
    const uint4* g1 = reinterpret_cast<const uint4*>(p.w1) + (size_t)rank * (W1_BYTES / 16);
    uint4* s0 = reinterpret_cast<uint4*>(smem + W0_OFF);
    uint4* s1 = reinterpret_cast<uint4*>(smem + W1_OFF);
    for (int i = threadIdx.x; i < W0_BYTES / 16; i += THREADS) s0[i] = __ldg(g0 + i);
    for (int i = threadIdx.x; i < W1_BYTES / 16; i += THREADS) s1[i] = __ldg(g1 + i);
    uint4* a0 = reinterpret_cast<uint4*>(smem + A0_OFF);   // zero once: K padding columns stay 0
    for (int i = threadIdx.x; i < A0_BYTES / 16; i += THREADS) a0[i] = make_uint4(0, 0, 0, 0);
    for (int i = threadIdx.x; i < HID; i += THREADS) b0s[i] = p.b0[i];
  }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  const uint32_t tmem = tmem_base_slot;

  // ---- roles ----------------------------------------------------------------------------------
  if (warp < LOAD_WARPS) {
    // ===== loader: global fp32 rows -> fp16 swizzled A0 =====
    uint8_t* A0 = smem + A0_OFF;
    const bool al16 = ((reinterpret_cast<uintptr_t>(p.feats) & 15) == 0) && (p.F % 4 == 0);
    const bool al8 = ((reinterpret_cast<uintptr_t>(p.feats) & 7) == 0) && (p.F % 2 == 0);
    int it = 0;
    for (int t = cluster; t < p.T; t += n_clusters, ++it) {
      if (it > 0) ptx::mbar_wait(bar(B_A0_EMPTY), (it - 1) & 1);
      const int64_t row0 = (int64_t)t * PAIR + rank * TILE;
      uint32_t nzmask;
      if (al16) load_convert_tile<4>(p, A0, row0, warp, lane, nzmask);
      else if (al8) load_convert_tile<2>(p, A0, row0, warp, lane, nzmask);
      else load_convert_tile<1>(p, A0, row0, warp, lane, nzmask);
      if (lane == 0) p.rowmask[(size_t)t * 8 + rank * 4 + warp] = nzmask;
      ptx::fence_proxy_async();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(bar(B_A0_FULL), 0);
    }
  } else if (warp == MMA_WARP) {
    // ===== MMA issuer: one thread of the leader CTA =====
    if (rank == 0) {
      const uint32_t idesc = ptx::umma_idesc_f16(256, 256);
      const uint64_t dW0 = ptx::umma_desc_sw128(sbase + W0_OFF), dW1 = ptx::umma_desc_sw128(sbase + W1_OFF);
      const uint64_t dA0 = ptx::umma_desc_sw128(sbase + A0_OFF), dA1 = ptx::umma_desc_sw128(sbase + A1_OFF);
      // +32 B per K=16 step inside an atom (encoded >>4 -> +2), +ATOM per 64 columns
      auto kadv = [](int k) -> uint64_t { return (uint64_t)(((k >> 2) * ATOM + (k & 3) * 32) >> 4); };
      int it = 0;
      for (int t = cluster; t < p.T; t += n_clusters, ++it) {
        ptx::mbar_wait(bar(B_A0_FULL), it & 1);
        ptx::tc_fence_after();
        if (lane == 0) {
          for (int k = 0; k < p.k0_steps; ++k)
            ptx::umma_f16_2cta(tmem + 0, dA0 + kadv(k), dW0 + kadv(k), idesc, k > 0);
          ptx::umma_commit_2cta(bar(B_A0_EMPTY), 3);
          ptx::umma_commit_2cta(bar(B_ACC0_FULL), 3);
        }
        __syncwarp();
        ptx::mbar_wait(bar(B_A1_FULL), it & 1);
        if (it > 0) ptx::mbar_wait(bar(B_ACC1_EMPTY), (it - 1) & 1);
        ptx::tc_fence_after();
        if (lane == 0) {
#pragma unroll
          for (int k = 0; k < HID / 16; ++k)
            ptx::umma_f16_2cta(tmem + 256, dW1 + kadv(k), dA1 + kadv(k), idesc, k > 0);
          ptx::umma_commit_2cta(bar(B_ACC1_FULL), 3);
        }
        __syncwarp();
      }
    }
  } else {
    // ===== epilogue warps =====
    const int q = warp & 3;                       // TMEM lane quarter this warp may access
    const uint32_t lane_base = (uint32_t)(32 * q) << 16;
    const int row = 32 * q + lane;                // epi 0: row of this CTA's half tile
    const int hu = rank * TILE + 32 * q + lane;   // epi 1: hidden unit
    const float b1 = p.b1[hu];
    uint8_t* A1 = smem + A1_OFF;
    const uint32_t a1_row = (uint32_t)((row >> 3) * 1024 + (row & 7) * 128);
    int it = 0;
    for (int t = cluster; t < p.T; t += n_clusters, ++it) {
      // ---- epi 0 ----
      ptx::mbar_wait(bar(B_ACC0_FULL), it & 1);
      ptx::tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < HID / 32; ++c) {
        uint32_t v[32];
        ptx::tmem_ld32(tmem + lane_base + 32 * c, v);
        ptx::tmem_ld_wait();
        uint32_t h[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float x0 = fmaxf(__uint_as_float(v[2 * j]) + b0s[32 * c + 2 * j], 0.f);
          const float x1 = fmaxf(__uint_as_float(v[2 * j + 1]) + b0s[32 * c + 2 * j + 1], 0.f);
          h[j] = pack_h2(x0, x1);
        }
        // columns 32c .. 32c+31 = 4 chunks of 8 fp16 inside atom (32c / 64)
        uint8_t* base = A1 + (c >> 1) * ATOM + a1_row;
#pragma unroll
        for (int m = 0; m < 4; ++m) {
          const int chunk = (c & 1) * 4 + m;
          *reinterpret_cast<uint4*>(base + ((chunk ^ (row & 7)) << 4)) = make_uint4(h[4 * m], h[4 * m + 1], h[4 * m + 2], h[4 * m + 3]);
        }
      }
      ptx::fence_proxy_async();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(bar(B_A1_FULL), 0);

      // ---- epi 1: segment sum over the tile's rows (TMEM columns) ----
      const int64_t R0 = (int64_t)t * PAIR;
      const int64_t Rend = min(R0 + (int64_t)PAIR, p.n_rows);
      int prog = p.first_prog[t];
      int batch_base = prog;
      int boff = p.off[min(batch_base + lane, p.P)];
      auto off_at = [&](int idx) { return __shfl_sync(0xffffffffu, boff, idx - batch_base); };
      int seg_begin = off_at(prog);
      int seg_end = off_at(prog + 1);
      ptx::mbar_wait(bar(B_ACC1_FULL), it & 1);
      ptx::tc_fence_after();
      float acc = 0.f;
#pragma unroll 1
      for (int c = 0; c < PAIR / 32; ++c) {
        uint32_t v[32];
        ptx::tmem_ld32(tmem + lane_base + 256 + 32 * c, v);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          acc += fmaxf(__uint_as_float(v[j]) + b1, 0.f);
          const int r1 = (int)(R0 + 32 * c + j + 1);
          if (r1 == seg_end) {                       // uniform across the CTA
            float* dst = p.out + (size_t)prog * HID + hu;
            if (seg_begin < R0) atomicAdd(dst, acc); else *dst = acc;
            acc = 0.f;
            do {                                     // next program, skipping empty ones
              ++prog;
              if (prog + 1 - batch_base > 31) { batch_base = prog; boff = p.off[min(batch_base + lane, p.P)]; }
              seg_begin = seg_end;
              seg_end = prog < p.P ? off_at(prog + 1) : INT32_MAX;
            } while (seg_end == r1);
          }
        }
      }
      if (prog < p.P && seg_begin < Rend) atomicAdd(p.out + (size_t)prog * HID + hu, acc);   // continues in the next tile
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(bar(B_ACC1_EMPTY), 0);
    }
  }

  // ---- teardown -------------------------------------------------------------------------------
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  if (warp == EPI_WARP0) ptx::tmem_dealloc_2cta(tmem, 512);
}

// ---------------------------------------------------------------------------------------------
// pre-pass: first program of every pair tile; zero the output rows that receive atomics
// (programs cut by a tile boundary) or nothing at all (empty programs).
__global__ void tc_prepare_kernel(const int32_t* __restrict__ off, int P, int T, int32_t* __restrict__ first_prog,
                                  float* __restrict__ out) {
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid < T) {
    const int target = gid * tc::PAIR;            // last p in [0, P) with off[p] <= target
    int lo = 0, hi = P;                           // invariant: off[lo] <= target, answer in [lo, hi)
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (off[mid] <= target) lo = mid; else hi = mid;
    }
    first_prog[gid] = lo;
  }
  const int lane = threadIdx.x & 31;
  const int pbase = gid - lane;                   // 32 programs per warp
  int need = 0;
  if (gid < P) {
    const int s = off[gid], e = off[gid + 1];
    need = (e == s) || (s / tc::PAIR != (e - 1) / tc::PAIR);
  }
  unsigned m = __ballot_sync(0xffffffffu, need);
  while (m) {
    const int b = __ffs(m) - 1;
    m &= m - 1;
    float4* dst = reinterpret_cast<float4*>(out + (size_t)(pbase + b) * tc::HID);
    dst[lane] = make_float4(0.f, 0.f, 0.f, 0.f);
    dst[lane + 32] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// valid[p] = any row of p has a non-zero feature (MLPModel.predict's rule, mlp_model.py:842-844)
__global__ void tc_valid_kernel(const int32_t* __restrict__ off, int P, const uint32_t* __restrict__ rowmask,
                                uint8_t* __restrict__ valid) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  int any = 0;
  for (int r = off[p]; r < off[p + 1] && !any; ++r) any = (rowmask[r >> 5] >> (r & 31)) & 1u;
  valid[p] = (uint8_t)any;
}

// ---------------------------------------------------------------------------------------------
static __half h_clamp(float x) { return __float2half_rn(std::min(std::max(x, -65504.f), 65504.f)); }

int tc_pack(vaex_model* m, const std::vector<float>& w0, cudaStream_t st) {
  using namespace tc;
  m->tc.ready = false;
  if (m->H != HID || m->F > K0P || m->F < 1) return VAEX_OK;   // shape outside the tcgen05 kernel: SIMT path
  const int F = m->F;
  std::vector<__half> i0(2 * W0_BYTES / 2, __float2half_rn(0.f)), i1(2 * W1_BYTES / 2);
  const std::vector<float>& w1 = m->host["segment_encoder.2.weight"].v;
  for (int r = 0; r < 2; ++r)
    for (int row = 0; row < TILE; ++row) {
      const int h = r * TILE + row;
      for (int k = 0; k < F; ++k) i0[(size_t)r * (W0_BYTES / 2) + sw128_off(row, k) / 2] = h_clamp(w0[(size_t)h * F + k]);
      for (int k = 0; k < HID; ++k) i1[(size_t)r * (W1_BYTES / 2) + sw128_off(row, k) / 2] = h_clamp(w1[(size_t)h * HID + k]);
    }
  TcPack& t = m->tc;
  if (!t.w0) VAEX_CUDA(cudaMalloc(&t.w0, 2 * W0_BYTES));
  if (!t.w1) VAEX_CUDA(cudaMalloc(&t.w1, 2 * W1_BYTES));
  if (!t.b0) VAEX_CUDA(cudaMalloc(&t.b0, HID * sizeof(float)));
  if (!t.b1) VAEX_CUDA(cudaMalloc(&t.b1, HID * sizeof(float)));
  VAEX_CUDA(cudaMemcpyAsync(t.w0, i0.data(), 2 * W0_BYTES, cudaMemcpyHostToDevice, st));
  VAEX_CUDA(cudaMemcpyAsync(t.w1, i1.data(), 2 * W1_BYTES, cudaMemcpyHostToDevice, st));
  VAEX_CUDA(cudaMemcpyAsync(t.b0, m->host["segment_encoder.0.bias"].v.data(), HID * sizeof(float), cudaMemcpyHostToDevice, st));
  VAEX_CUDA(cudaMemcpyAsync(t.b1, m->host["segment_encoder.2.bias"].v.data(), HID * sizeof(float), cudaMemcpyHostToDevice, st));
  VAEX_CUDA(cudaStreamSynchronize(st));
  VAEX_CUDA(cudaFuncSetAttribute(segsum_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_DYN));
  t.F = F;
  t.K0P = K0P;
  t.ready = true;
  return VAEX_OK;
}

int segsum_tc_encode(vaex_model* m, const float* feats, const int32_t* off, int64_t P, int64_t R, float* segsum,
                     uint8_t* valid, cudaStream_t st) {
  using namespace tc;
  VAEX_CHECK(m->tc.ready, "tcgen05 weights not packed");
  VAEX_CHECK(P < (1ll << 31) && R < (1ll << 31) - PAIR, "sizes exceed int32 row indexing");
  const int T = (int)cdiv(R, PAIR);
  const int nprep = (int)std::max<int64_t>(P, T);
  { int rc = m->tc_ws.ensure(sizeof(int32_t) * (size_t)std::max(T, 1) * 9); if (rc != VAEX_OK) return rc; }
  int32_t* first_prog = m->tc_ws.as<int32_t>();
  uint32_t* rowmask = reinterpret_cast<uint32_t*>(first_prog + std::max(T, 1));
  tc_prepare_kernel<<<(unsigned)cdiv(nprep, 256), 256, 0, st>>>(off, (int)P, T, first_prog, segsum);
  VAEX_LAUNCH_CHECK();
  if (T > 0) {
    TcParams prm;
    prm.feats = feats; prm.off = off; prm.first_prog = first_prog;
    prm.w0 = m->tc.w0; prm.w1 = m->tc.w1; prm.b0 = m->tc.b0; prm.b1 = m->tc.b1;
    prm.out = segsum; prm.rowmask = rowmask; prm.n_rows = R; prm.P = (int)P; prm.T = T; prm.F = m->F;
    prm.k0_steps = (m->F + 15) / 16;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int clusters = std::max(1, std::min(sms / 2, T));
    segsum_tc_kernel<<<2 * clusters, THREADS, SMEM_DYN, st>>>(prm);
    VAEX_LAUNCH_CHECK();
  }
  if (valid) {
    tc_valid_kernel<<<(unsigned)cdiv(P, 256), 256, 0, st>>>(off, (int)P, rowmask, valid);
    VAEX_LAUNCH_CHECK();
  }
  return VAEX_OK;
}

}  // namespace vaex
