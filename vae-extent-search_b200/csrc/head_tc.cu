// Residual head of the SegmentSumMLP on tcgen05 tensor cores:
//
//   y0 = norm(S)                       (Identity, or the folded eval-mode BatchNorm affine)
//   y1 = relu(y0 Wl0^T + bl0) + y0
//   y2 = relu(y1 Wl1^T + bl1) + y1
//   score = y2 . wd + bd               (+ sigmoid, + -inf for invalid programs)
//   (SegmentSumMLPModule.forward, python/tvm/auto_scheduler/cost_model/mlp_model.py:210-221)
//
// One persistent CTA pair (cta_group::2) per 256 programs per step: each CTA owns 128 programs
// (TMEM lanes) and half of the output units of both 256x256 weights (resident, 128 KB fp16); the
// 64 KB A operand holds fp16(y0) and is rewritten in place with fp16(y1).  TMEM: columns [0,256)
// = fp32 y0 (the residual of layer 0), [256,512) = the accumulator.  Nothing intermediate touches
// HBM and S is read exactly once.
//
// v3 (same piece pipeline as chain_tc.cu): every GEMM is issued as four MMA pieces (n-half, k-half),
// k-half major, with one commit per n-half; the epilogue of n-half h writes exactly k-half h of the
// next GEMM's operand, so pieces start as soon as their half is published (mbarriers IN0/IN1) and
// their accumulator half has been drained.  score = wd.relu(acc1 + bl1) + wd.y1 + bd, and wd.y1 is
// accumulated while y1 is formed, so y1 never goes back to TMEM and the last epilogue reads only
// the accumulator.  16 worker warps: thread = (program row, 32-column group of the half).  The
// residual columns double as the staging buffer of the next step: they are free once the layer-0
// epilogue has read them, so the next rows are loaded and stored there (fp32) behind GEMM 1, and
// the step boundary only converts them into the operand (tcgen05.ld -> fp16 -> st.shared).
// Warp 16 of the leader CTA issues the MMAs.
#include <algorithm>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "model.cuh"
#include "tc_ptx.cuh"

namespace vaex {

namespace hd {
constexpr int HID = 256, TILE = 128, PAIR = 256, ATOM = 128 * 128;
constexpr int W_BYTES = 4 * ATOM;                 // [128 units x 256 k] fp16, K-major SW128
constexpr int WL0_OFF = 0, WL1_OFF = W_BYTES, A_OFF = 2 * W_BYTES;
constexpr int SMEM_TILES = 3 * W_BYTES;           // 196608
constexpr int SMEM_DYN = SMEM_TILES + 1024;
constexpr int WORKER_WARPS = 16, WORKERS = 32 * WORKER_WARPS, THREADS = WORKERS + 32;
enum { B_IN0 = 0, B_IN1, B_ACC0, B_ACC1, B_COUNT };

__host__ __device__ inline uint32_t sw128_off(int row, int k) {
  return (uint32_t)((k >> 6) * ATOM + (row >> 3) * 1024 + (row & 7) * 128 + ((((k & 63) >> 3) ^ (row & 7)) << 4) +
                    (k & 7) * 2);
}
}  // namespace hd

struct HeadParams {
  const float* S;            // [P, 256] segment sums
  const __half* wl0;         // [2][W_BYTES/2] swizzled images per CTA rank
  const __half* wl1;
  const float* bl0;          // [256]
  const float* bl1;          // [256]
  const float* wd;           // [256]
  const float* bd;           // [1]
  const float* nscale;       // [256] or null (BatchNorm fold)
  const float* nshift;
  const uint8_t* valid;      // [P] or null
  float* scores;             // [P]
  int32_t P, T, sigmoid;
  // single-layer mode (HF_STOP0): out = act(X Wl0^T + bl0) [+ X], as a vector or reduced with wd -- the building block of
  // the segment-VAE family's latent heads, decoder and cost predictor (vae.py:87-111, regression.py:82-111)
  int32_t flags, k_in;       // k_in: valid input columns (a multiple of 32; the rest of the 256-wide operand is zero)
  int64_t pitch;             // floats between input rows
  float* out0;               // HF_OUT_VEC: columns [0, n0) -> out0 [P, n0]; columns [n0, n0 + n1) -> out1 [P, n1]
  float* out1;
  int32_t n0, n1;
};
enum { HF_RELU0 = 1, HF_RES0 = 2, HF_STOP0 = 16, HF_OUT_VEC = 32 };

__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
        "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
        "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
        "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(hd::THREADS, 1)
head_tc_kernel(const HeadParams p) {
  using namespace hd;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[B_COUNT];
  __shared__ uint32_t tmem_base_slot;
  __shared__ __align__(16) float s_bl0[HID], s_bl1[HID], s_wd[HID], s_ns[HID], s_nb[HID];
  __shared__ float s_dot[3 * TILE];

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t sbase = ptx::smem_u32(smem);
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const int cluster = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const uint32_t bar0 = ptx::smem_u32(&bars[0]);
  auto bar = [&](int i) { return bar0 + 8u * i; };
  const bool has_norm = p.nscale != nullptr;

  if (threadIdx.x == 0) {
    ptx::mbar_init(bar(B_IN0), 2 * WORKER_WARPS);      // worker warps x 2 CTAs
    ptx::mbar_init(bar(B_IN1), 2 * WORKER_WARPS);
    ptx::mbar_init(bar(B_ACC0), 1);
    ptx::mbar_init(bar(B_ACC1), 1);
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc_2cta(ptx::smem_u32(&tmem_base_slot), 512);
  {
    const uint4* g0 = reinterpret_cast<const uint4*>(p.wl0) + (size_t)rank * (W_BYTES / 16);
    const uint4* g1 = reinterpret_cast<const uint4*>(p.wl1) + (size_t)rank * (W_BYTES / 16);
    uint4* s0 = reinterpret_cast<uint4*>(smem + WL0_OFF);
    uint4* s1 = reinterpret_cast<uint4*>(smem + WL1_OFF);
    for (int i = threadIdx.x; i < W_BYTES / 16; i += THREADS) { s0[i] = __ldg(g0 + i); s1[i] = __ldg(g1 + i); }
    for (int i = threadIdx.x; i < HID; i += THREADS) {
      s_bl0[i] = p.bl0[i]; s_bl1[i] = p.bl1 ? p.bl1[i] : 0.f; s_wd[i] = p.wd ? p.wd[i] : 0.f;
      s_ns[i] = has_norm ? p.nscale[i] : 1.f;
      s_nb[i] = has_norm ? p.nshift[i] : 0.f;
    }
  }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  const uint32_t tmem = tmem_base_slot;

  if (warp < WORKER_WARPS) {
    // ===== workers =====
    const int q = warp & 3, cg = warp >> 2;          // TMEM lane quarter, 32-column group of a half
    const int row = 32 * q + lane;                   // program row (TMEM lane) of this thread
    const uint32_t a_row = sbase + A_OFF + (row >> 3) * 1024 + (row & 7) * 128;
    const int sw = row & 7;
    const uint32_t t_res = tmem + ((uint32_t)(32 * q) << 16), t_acc = t_res + 256;
    const float bd = p.bd ? p.bd[0] : 0.f;
    float4 x[8];                                     // 32 floats of the next step's row (one half at a time)

    auto load_half = [&](int t, int h) {
      const int64_t prog = (int64_t)t * PAIR + rank * TILE + row;
      if (t < p.T && prog < p.P && 128 * h + 32 * cg < p.k_in) {
        const float* src = p.S + prog * p.pitch + 128 * h + 32 * cg;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const ptx::F8 f = ptx::ldg_stream_f8(src + 8 * j);
          x[2 * j] = make_float4(f.v[0], f.v[1], f.v[2], f.v[3]);
          x[2 * j + 1] = make_float4(f.v[4], f.v[5], f.v[6], f.v[7]);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) x[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    };
    auto prefetch_l2 = [&](int t) {                  // this warp's 8 KB slice of a later step
      if (lane != 0 || t >= p.T) return;
      const int64_t r0 = (int64_t)t * PAIR + rank * TILE + warp * 8;
      const int64_t nr = min((int64_t)8, (int64_t)p.P - r0);
      if (nr > 0 && p.pitch == p.k_in) ptx::prefetch_l2_bulk(p.S + r0 * p.pitch, (uint32_t)(nr * p.pitch * 4));
    };
    // 32 fp32 values -> 64 B of the fp16 A operand, 32-column chunk cc of the row
    auto pack32 = [&](const uint32_t (&v)[32], int cc) {
      const uint32_t base = a_row + (cc >> 1) * ATOM;
#pragma unroll
      for (int m = 0; m < 4; ++m)
        ptx::sts_v4(base + ((((cc & 1) * 4 + m) ^ sw) << 4),
                    ptx::cvt_h2_sat(__uint_as_float(v[8 * m]), __uint_as_float(v[8 * m + 1])),
                    ptx::cvt_h2_sat(__uint_as_float(v[8 * m + 2]), __uint_as_float(v[8 * m + 3])),
                    ptx::cvt_h2_sat(__uint_as_float(v[8 * m + 4]), __uint_as_float(v[8 * m + 5])),
                    ptx::cvt_h2_sat(__uint_as_float(v[8 * m + 6]), __uint_as_float(v[8 * m + 7])));
    };
    auto publish = [&](int b) {                      // operand + TMEM writes of this warp -> issuer
      ptx::fence_proxy_async();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster_relaxed(bar(b), 0);
    };
    // y0 half h of the loaded row: norm -> fp32 residual columns of TMEM.  The residual doubles as the staging
    // buffer of the next step: it is free as soon as the layer-0 epilogue has read it, long before the operand is.
    auto commit_half = [&](int h) {
      const int col = 128 * h + 32 * cg;
      uint32_t v[32];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float4 a = x[j];
        if (has_norm) {
          const float4 sc = *reinterpret_cast<const float4*>(s_ns + col + 4 * j), sh = *reinterpret_cast<const float4*>(s_nb + col + 4 * j);
          a.x = a.x * sc.x + sh.x; a.y = a.y * sc.y + sh.y; a.z = a.z * sc.z + sh.z; a.w = a.w * sc.w + sh.w;
        }
        v[4 * j] = __float_as_uint(a.x); v[4 * j + 1] = __float_as_uint(a.y);
        v[4 * j + 2] = __float_as_uint(a.z); v[4 * j + 3] = __float_as_uint(a.w);
      }
      tmem_st32(t_res + col, v);
    };
    // residual half h -> fp16 operand k-half h, publish
    auto pack_half = [&](int h) {
      uint32_t v[32];
      ptx::tmem_ld32(t_res + 128 * h + 32 * cg, v);
      ptx::tmem_ld_wait();
      pack32(v, 4 * h + cg);
      publish(h == 0 ? B_IN0 : B_IN1);
    };

    uint32_t acc_use0 = 0, acc_use1 = 0;
    prefetch_l2(cluster + n_clusters);
    load_half(cluster, 0);
    commit_half(0);
    load_half(cluster, 1);
    commit_half(1);
    tmem_st_wait();
    pack_half(0);
    pack_half(1);
    if (p.flags & HF_STOP0) {
      // ===== single layer per tile: epilogue of half h, then the next tile's half h goes into the freed residual columns;
      // the operand is rewritten once both accumulator halves have been committed (every MMA of the tile is complete) =====
      const bool relu = p.flags & HF_RELU0, res = p.flags & HF_RES0, vec = p.flags & HF_OUT_VEC;
      for (int t = cluster; t < p.T; t += n_clusters) {
        prefetch_l2(t + 2 * n_clusters);
        const int64_t prog = (int64_t)t * PAIR + rank * TILE + row;
        float dot0 = 0.f, dot1 = 0.f;
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
          const int col = 128 * h + 32 * cg;
          if (h == 0) { ptx::mbar_wait(bar(B_ACC0), acc_use0 & 1); ++acc_use0; }
          else        { ptx::mbar_wait(bar(B_ACC1), acc_use1 & 1); ++acc_use1; }
          ptx::tc_fence_after();
          {
            uint32_t v[32];
            ptx::tmem_ld32(t_acc + col, v);
            ptx::tmem_ld_wait();
            const float* bp = s_bl0 + col;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float a = __uint_as_float(v[j]) + bp[j];
              v[j] = __float_as_uint(relu ? fmaxf(a, 0.f) : a);
            }
            if (res) {
              uint32_t r[32];
              ptx::tmem_ld32(t_res + col, r);
              ptx::tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(r[j]));
            }
            if (vec) {
              if (prog < p.P && col < p.n0 + p.n1) {
                float* dst = col < p.n0 ? p.out0 + prog * p.n0 + col : p.out1 + prog * p.n1 + (col - p.n0);
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  ptx::stg_f8(dst + 8 * j, __uint_as_float(v[8 * j]), __uint_as_float(v[8 * j + 1]), __uint_as_float(v[8 * j + 2]),
                              __uint_as_float(v[8 * j + 3]), __uint_as_float(v[8 * j + 4]), __uint_as_float(v[8 * j + 5]),
                              __uint_as_float(v[8 * j + 6]), __uint_as_float(v[8 * j + 7]));
              }
            } else {
              const float4* w4 = reinterpret_cast<const float4*>(s_wd + col);
#pragma unroll
              for (int j4 = 0; j4 < 8; ++j4) {
                const float4 w = w4[j4];
                dot0 = fmaf(__uint_as_float(v[4 * j4]), w.x, dot0); dot1 = fmaf(__uint_as_float(v[4 * j4 + 1]), w.y, dot1);
                dot0 = fmaf(__uint_as_float(v[4 * j4 + 2]), w.z, dot0); dot1 = fmaf(__uint_as_float(v[4 * j4 + 3]), w.w, dot1);
              }
            }
          }
          load_half(t + n_clusters, h);                    // the next tile's rows (L2 hits: bulk-prefetched two tiles ahead)
          commit_half(h);                                  // residual half h is free: the next tile's rows go in
        }
        tmem_st_wait();
        ptx::tc_fence_before();
        if (t + n_clusters < p.T) {
          pack_half(0);
          pack_half(1);
        }
        if (!vec) {
          const float dot = dot0 + dot1;
          if (cg) s_dot[(cg - 1) * TILE + row] = dot;
          ptx::bar_sync(1, WORKERS);
          if (cg == 0 && prog < p.P) {
            float sc = ((dot + s_dot[row]) + (s_dot[TILE + row] + s_dot[2 * TILE + row])) + bd;
            if (p.sigmoid) sc = 1.f / (1.f + expf(-sc));
            if (p.valid && !p.valid[prog]) sc = -INFINITY;
            p.scores[prog] = sc;
          }
          ptx::bar_sync(1, WORKERS);                       // s_dot is reused by the next tile
        }
      }
    } else
    for (int t = cluster; t < p.T; t += n_clusters) {
      prefetch_l2(t + 2 * n_clusters);
      float dot0 = 0.f, dot1 = 0.f;
      // ---- layer 0: y1 = relu(acc + bl0) + y0 -> operand (fp16); wd . y1 accumulated ----
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        const int col = 128 * h + 32 * cg;
        if (h == 0) { ptx::mbar_wait(bar(B_ACC0), acc_use0 & 1); ++acc_use0; }
        else        { ptx::mbar_wait(bar(B_ACC1), acc_use1 & 1); ++acc_use1; }
        ptx::tc_fence_after();
        uint32_t v[32], r[32];
        ptx::tmem_ld32(t_acc + col, v);
        ptx::tmem_ld32(t_res + col, r);
        ptx::tmem_ld_wait();
        const float4* b4 = reinterpret_cast<const float4*>(s_bl0 + col);
        const float4* w4 = reinterpret_cast<const float4*>(s_wd + col);
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 b = b4[j4], w = w4[j4];
          const float y0 = fmaxf(__uint_as_float(v[4 * j4 + 0]) + b.x, 0.f) + __uint_as_float(r[4 * j4 + 0]);
          const float y1 = fmaxf(__uint_as_float(v[4 * j4 + 1]) + b.y, 0.f) + __uint_as_float(r[4 * j4 + 1]);
          const float y2 = fmaxf(__uint_as_float(v[4 * j4 + 2]) + b.z, 0.f) + __uint_as_float(r[4 * j4 + 2]);
          const float y3 = fmaxf(__uint_as_float(v[4 * j4 + 3]) + b.w, 0.f) + __uint_as_float(r[4 * j4 + 3]);
          dot0 = fmaf(y0, w.x, dot0); dot1 = fmaf(y1, w.y, dot1); dot0 = fmaf(y2, w.z, dot0); dot1 = fmaf(y3, w.w, dot1);
          v[4 * j4 + 0] = __float_as_uint(y0); v[4 * j4 + 1] = __float_as_uint(y1);
          v[4 * j4 + 2] = __float_as_uint(y2); v[4 * j4 + 3] = __float_as_uint(y3);
        }
        pack32(v, 4 * h + cg);
        publish(h == 0 ? B_IN0 : B_IN1);
      }
      // the next step's rows stream into the (now free) residual columns behind GEMM 1
      load_half(t + n_clusters, 0);
      commit_half(0);
      load_half(t + n_clusters, 1);
      // ---- layer 1: wd . relu(acc + bl1) ----
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        const int col = 128 * h + 32 * cg;
        if (h == 0) { ptx::mbar_wait(bar(B_ACC0), acc_use0 & 1); ++acc_use0; }
        else        { ptx::mbar_wait(bar(B_ACC1), acc_use1 & 1); ++acc_use1; }
        ptx::tc_fence_after();
        uint32_t v[32];
        ptx::tmem_ld32(t_acc + col, v);
        ptx::tmem_ld_wait();
        const float4* b4 = reinterpret_cast<const float4*>(s_bl1 + col);
        const float4* w4 = reinterpret_cast<const float4*>(s_wd + col);
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 b = b4[j4], w = w4[j4];
          dot0 = fmaf(fmaxf(__uint_as_float(v[4 * j4 + 0]) + b.x, 0.f), w.x, dot0);
          dot1 = fmaf(fmaxf(__uint_as_float(v[4 * j4 + 1]) + b.y, 0.f), w.y, dot1);
          dot0 = fmaf(fmaxf(__uint_as_float(v[4 * j4 + 2]) + b.z, 0.f), w.z, dot0);
          dot1 = fmaf(fmaxf(__uint_as_float(v[4 * j4 + 3]) + b.w, 0.f), w.w, dot1);
        }
        if (h == 0) commit_half(1);
      }
      // every MMA of this step has completed: the operand is free, the next y0 is already in the residual columns
      tmem_st_wait();
      ptx::tc_fence_before();
      if (t + n_clusters < p.T) {
        pack_half(0);
        pack_half(1);
      }
      const float dot = dot0 + dot1;
      if (cg) s_dot[(cg - 1) * TILE + row] = dot;
      ptx::bar_sync(1, WORKERS);
      const int64_t prog = (int64_t)t * PAIR + rank * TILE + row;
      if (cg == 0 && prog < p.P) {
        float s = ((dot + s_dot[row]) + (s_dot[TILE + row] + s_dot[2 * TILE + row])) + bd;
        if (p.sigmoid) s = 1.f / (1.f + expf(-s));
        if (p.valid && !p.valid[prog]) s = -INFINITY;
        p.scores[prog] = s;
      }
    }
  } else if (rank == 0) {
    // ===== MMA issuer (leader CTA) =====
    const uint32_t idesc = ptx::umma_idesc_f16(256, 128);
    const uint64_t dA = ptx::umma_desc_sw128(sbase + A_OFF);
    auto kadv = [](int k) -> uint64_t { return (uint64_t)(((k >> 2) * ATOM + (k & 3) * 32) >> 4); };
    uint32_t in_use0 = 0, in_use1 = 0;
    // piece (n-half nh, k-half kh) of the GEMM with weights at w_off
    auto piece = [&](uint32_t w_off, int nh, int kh) {
      const uint64_t dW = ptx::umma_desc_sw128(sbase + w_off + nh * 64 * 128);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int k = 8 * kh + i;
        ptx::umma_f16_2cta(tmem + 256 + 128 * nh, dA + kadv(k), dW + kadv(k), idesc, k > 0);
      }
    };
    auto wait_in = [&](int h) {
      if (h == 0) { ptx::mbar_wait(bar(B_IN0), in_use0 & 1); ++in_use0; }
      else        { ptx::mbar_wait(bar(B_IN1), in_use1 & 1); ++in_use1; }
      ptx::tc_fence_after();
    };
    for (int t = cluster; t < p.T; t += n_clusters) {
      // GEMM 0: both accumulator halves are free once the operand is there
      wait_in(0);
      if (ptx::elect_one()) { piece(WL0_OFF, 0, 0); piece(WL0_OFF, 1, 0); }
      __syncwarp();
      wait_in(1);
      if (ptx::elect_one()) {
        piece(WL0_OFF, 0, 1); ptx::umma_commit_2cta(bar(B_ACC0), 3);
        piece(WL0_OFF, 1, 1); ptx::umma_commit_2cta(bar(B_ACC1), 3);
      }
      __syncwarp();
      if (p.flags & HF_STOP0) continue;                    // single-layer mode: the next tile's operand is the next wait
      // GEMM 1: accumulator half h is free when the layer-0 epilogue of half h has published
      wait_in(0);
      if (ptx::elect_one()) piece(WL1_OFF, 0, 0);
      __syncwarp();
      wait_in(1);
      if (ptx::elect_one()) {
        piece(WL1_OFF, 1, 0);
        piece(WL1_OFF, 0, 1); ptx::umma_commit_2cta(bar(B_ACC0), 3);
        piece(WL1_OFF, 1, 1); ptx::umma_commit_2cta(bar(B_ACC1), 3);
      }
      __syncwarp();
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  if (warp == 1) ptx::tmem_dealloc_2cta(tmem, 512);
}

// ---------------------------------------------------------------------------------------------
static __half hh_clamp(float x) { return __float2half_rn(std::min(std::max(x, -65504.f), 65504.f)); }

// [out_rows <= 256, in_cols <= 256] fp32 weight, zero-padded to 256 x 256, as the per-rank swizzled fp16 images of one
// head_tc layer (each GEMM is issued as two 128-column MMAs: image rows [64 h, +64) of rank r = units 128 h + 64 r + [0, 64))
int head_tc_pack_layer(const std::vector<float>& w, int out_rows, int in_cols, __half** dev, cudaStream_t st) {
  using namespace hd;
  VAEX_CHECK(out_rows >= 1 && out_rows <= HID && in_cols >= 1 && in_cols <= HID, "layer outside the 256 x 256 head kernel");
  std::vector<__half> img(2 * W_BYTES / 2, __float2half_rn(0.f));
  for (int r = 0; r < 2; ++r)
    for (int row = 0; row < TILE; ++row) {
      const int unit = 128 * (row >> 6) + 64 * r + (row & 63);
      if (unit >= out_rows) continue;
      for (int k = 0; k < in_cols; ++k)
        img[(size_t)r * (W_BYTES / 2) + sw128_off(row, k) / 2] = hh_clamp(w[(size_t)unit * in_cols + k]);
    }
  if (!*dev) VAEX_CUDA(cudaMalloc(dev, 2 * W_BYTES));
  VAEX_CUDA(cudaMemcpyAsync(*dev, img.data(), 2 * W_BYTES, cudaMemcpyHostToDevice, st));
  VAEX_CUDA(cudaStreamSynchronize(st));
  VAEX_CUDA(cudaFuncSetAttribute(head_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_DYN));
  return VAEX_OK;
}

// One layer on tensor cores: Y = act(X W^T + b) [+ X]  ->  vector outputs (out0 [P, n0], out1 [P, n1]) or the scalar
// wd . Y + bd.  X fp32 [P, k_in] with `pitch` floats between rows (k_in, n0, n1 multiples of 32); w_img from
// head_tc_pack_layer; bias256 / wd256: fp32 [256], zero padded.
int head_tc_linear(const HeadLinear& a, cudaStream_t st) {
  using namespace hd;
  if (a.P == 0) return VAEX_OK;
  VAEX_CHECK(a.k_in % 32 == 0 && a.k_in <= HID && a.pitch % 8 == 0 && a.pitch >= a.k_in, "bad input shape for the head kernel");
  VAEX_CHECK((reinterpret_cast<uintptr_t>(a.X) & 31) == 0, "input must be 32-byte aligned");
  HeadParams prm;
  memset(&prm, 0, sizeof(prm));
  prm.S = a.X; prm.wl0 = a.w_img; prm.wl1 = a.w_img; prm.bl0 = a.bias256; prm.bl1 = nullptr;
  prm.wd = a.wd256; prm.bd = a.bd; prm.nscale = a.nscale; prm.nshift = a.nshift; prm.valid = nullptr; prm.scores = a.score;
  prm.P = (int)a.P; prm.T = (int)cdiv(a.P, PAIR); prm.sigmoid = 0;
  prm.flags = HF_STOP0 | (a.relu ? HF_RELU0 : 0) | (a.residual ? HF_RES0 : 0) | (a.score ? 0 : HF_OUT_VEC);
  prm.k_in = a.k_in; prm.pitch = a.pitch; prm.out0 = a.out0; prm.n0 = a.n0; prm.out1 = a.out1; prm.n1 = a.n1;
  if (!a.score) {
    VAEX_CHECK(a.out0 && a.n0 % 32 == 0 && a.n1 % 32 == 0 && a.n0 + a.n1 <= HID && (a.n1 == 0 || a.out1), "bad vector output");
    VAEX_CHECK(((reinterpret_cast<uintptr_t>(a.out0) | reinterpret_cast<uintptr_t>(a.out1)) & 31) == 0, "outputs must be 32-byte aligned");
  } else {
    VAEX_CHECK(a.wd256 && a.bd, "the scalar output needs wd and bd");
  }
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int clusters = std::max(1, std::min(sms / 2, prm.T));
  head_tc_kernel<<<2 * clusters, THREADS, SMEM_DYN, st>>>(prm);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

int head_tc_pack(vaex_model* m, cudaStream_t st) {
  using namespace hd;
  m->tc.head_ready = false;
  if (m->H != HID) return VAEX_OK;
  std::vector<__half> img[2];
  const char* names[2] = {"l0.0.weight", "l1.0.weight"};
  for (int l = 0; l < 2; ++l) {
    img[l].assign(2 * W_BYTES / 2, __float2half_rn(0.f));
    const std::vector<float>& w = m->host[names[l]].v;
    for (int r = 0; r < 2; ++r)
      for (int row = 0; row < TILE; ++row)
        for (int k = 0; k < HID; ++k)
          // each GEMM is issued as two 128-column MMAs: image rows [64 h, +64) of rank r = units 128 h + 64 r + [0, 64)
          img[l][(size_t)r * (W_BYTES / 2) + sw128_off(row, k) / 2] = hh_clamp(w[(size_t)(128 * (row >> 6) + 64 * r + (row & 63)) * HID + k]);
  }
  TcPack& t = m->tc;
  if (!t.wl0) VAEX_CUDA(cudaMalloc(&t.wl0, 2 * W_BYTES));
  if (!t.wl1) VAEX_CUDA(cudaMalloc(&t.wl1, 2 * W_BYTES));
  VAEX_CUDA(cudaMemcpyAsync(t.wl0, img[0].data(), 2 * W_BYTES, cudaMemcpyHostToDevice, st));
  VAEX_CUDA(cudaMemcpyAsync(t.wl1, img[1].data(), 2 * W_BYTES, cudaMemcpyHostToDevice, st));
  VAEX_CUDA(cudaStreamSynchronize(st));
  VAEX_CUDA(cudaFuncSetAttribute(head_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_DYN));
  t.head_ready = true;
  return VAEX_OK;
}

int head_tc_forward(vaex_model* m, const float* S, int64_t P, const uint8_t* valid, float* scores, cudaStream_t st) {
  using namespace hd;
  VAEX_CHECK(m->tc.head_ready, "tcgen05 head weights not packed");
  if (P == 0) return VAEX_OK;
  HeadParams prm;
  prm.S = S; prm.wl0 = m->tc.wl0; prm.wl1 = m->tc.wl1;
  prm.bl0 = m->w("l0.0.bias"); prm.bl1 = m->w("l1.0.bias");
  prm.wd = m->w("decoder.weight"); prm.bd = m->w("decoder.bias");
  prm.nscale = m->use_norm ? m->w("norm.scale") : nullptr;
  prm.nshift = m->use_norm ? m->w("norm.shift") : nullptr;
  prm.valid = valid; prm.scores = scores; prm.P = (int)P; prm.T = (int)cdiv(P, PAIR); prm.sigmoid = m->add_sigmoid;
  prm.flags = 0; prm.k_in = HID; prm.pitch = HID; prm.out0 = prm.out1 = nullptr; prm.n0 = prm.n1 = 0;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int clusters = std::max(1, std::min(sms / 2, prm.T));
  head_tc_kernel<<<2 * clusters, THREADS, SMEM_DYN, st>>>(prm);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

}  // namespace vaex
