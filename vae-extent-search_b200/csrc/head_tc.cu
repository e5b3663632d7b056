// Residual head of the SegmentSumMLP on tcgen05 tensor cores:
//
//   y0 = norm(S)                       (Identity, or the folded eval-mode BatchNorm affine)
//   y1 = relu(y0 Wl0^T + bl0) + y0
//   y2 = relu(y1 Wl1^T + bl1) + y1
//   score = y2 . wd + bd               (+ sigmoid, + -inf for invalid programs)
//   (SegmentSumMLPModule.forward, python/tvm/auto_scheduler/cost_model/mlp_model.py:210-221)
//
// One persistent CTA pair (cta_group::2) per 256 programs per step: each CTA owns 128 programs
// (TMEM lanes) and half of the output units of both 256x256 weights (resident, 128 KB fp16).
// A worker thread owns one program row and one 128-column half of it end to end:
//   A. its 512 B of S arrive as 16 LDG.256 (issued during the previous step's second GEMM),
//      go as fp32 into TMEM columns [0,256) (the residual) and as fp16 into the A operand (64 KB)
//   B. GEMM 0 -> TMEM columns [256,512)
//   C. y1 = relu(acc + bl0) + residual: fp32 back over the residual, fp16 over the A operand
//   D. GEMM 1 -> the same accumulator columns
//   E. y2 = relu(acc + bl1) + residual, dotted with wd in registers
// so no intermediate touches HBM and S is read exactly once.  Warps 0-7 are workers (lane
// quarter = warp % 4, column half = warp / 4), warp 8 of the leader CTA issues the MMAs; hand-offs
// are mbarriers (warp arrivals -> issuer, tcgen05.commit multicast -> workers), no cluster barrier
// inside the loop.
#include <algorithm>
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
constexpr int WORKERS = 256, THREADS = 288;     // 8 worker warps + the MMA issuer
enum { B_Y0 = 0, B_ACC0, B_Y1, B_ACC1, B_COUNT };

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
};

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
struct F8 { float v[8]; };
__device__ __forceinline__ F8 ldg_stream_f8(const float* p) {   // one full 32 B sector per lane
  F8 r;
  asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7])
               : "l"(p));
  return r;
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(hd::THREADS, 1)
head_tc_kernel(const HeadParams p) {
  using namespace hd;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[B_COUNT];
  __shared__ uint32_t tmem_base_slot;
  __shared__ __align__(16) float s_bl0[HID], s_bl1[HID], s_wd[HID], s_ns[HID], s_nb[HID];
  __shared__ float s_dot[TILE];

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t sbase = ptx::smem_u32(smem);
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const int cluster = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const uint32_t bar0 = ptx::smem_u32(&bars[0]);
  auto bar = [&](int i) { return bar0 + 8u * i; };
  const bool has_norm = p.nscale != nullptr;

  if (threadIdx.x == 0) {
    ptx::mbar_init(bar(B_Y0), 16);      // 8 worker warps x 2 CTAs
    ptx::mbar_init(bar(B_ACC0), 1);
    ptx::mbar_init(bar(B_Y1), 16);
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
      s_bl0[i] = p.bl0[i]; s_bl1[i] = p.bl1[i]; s_wd[i] = p.wd[i];
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

  if (warp < 8) {
    // ===== workers =====
    const int q = warp & 3, ch = warp >> 2;          // TMEM lane quarter, column half
    const uint32_t lane_base = (uint32_t)(32 * q) << 16;
    const int row = 32 * q + lane;                   // program row (TMEM lane) of this thread
    const uint32_t a_row = sbase + A_OFF + (row >> 3) * 1024 + (row & 7) * 128;
    const int sw = row & 7;
    const int col0 = 128 * ch;
    const uint32_t t_res = tmem + lane_base + col0, t_acc = tmem + lane_base + 256 + col0;
    const float bd = p.bd[0];
    F8 pf[16];                                       // this thread's 128 floats of the next step

    auto fetch = [&](int t) {
      const int64_t prog = (int64_t)t * PAIR + rank * TILE + row;
      if (t < p.T && prog < p.P) {
        const float* src = p.S + prog * HID + col0;
#pragma unroll
        for (int j = 0; j < 16; ++j) pf[j] = ldg_stream_f8(src + 8 * j);
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j)
#pragma unroll
          for (int e = 0; e < 8; ++e) pf[j].v[e] = 0.f;
      }
    };
    auto prefetch_l2 = [&](int t) {                  // this warp's 16 KB slice of a later step
      if (lane != 0 || t >= p.T) return;
      const int64_t r0 = (int64_t)t * PAIR + rank * TILE + warp * 16;
      const int64_t nr = min((int64_t)16, (int64_t)p.P - r0);
      if (nr > 0) ptx::prefetch_l2_bulk(p.S + r0 * HID, (uint32_t)(nr * HID * 4));
    };
    // fp32 values of 32 columns -> 64 B of the fp16 A operand (columns [col0 + 32c, +32))
    auto pack32 = [&](const uint32_t (&v)[32], int c) {
      const int cc = 4 * ch + c;                     // 32-column chunk of the row
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
      tmem_st_wait();
      ptx::fence_proxy_async();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster_relaxed(bar(b), 0);
    };

    prefetch_l2(cluster + n_clusters);
    fetch(cluster);
    int it = 0;
    for (int t = cluster; t < p.T; t += n_clusters, ++it) {
      prefetch_l2(t + 2 * n_clusters);
      // ---- A. y0 = norm(S): fp32 -> TMEM residual, fp16 -> A operand ----
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float x = pf[4 * c + (j >> 3)].v[j & 7];
          if (has_norm) x = x * s_ns[col0 + 32 * c + j] + s_nb[col0 + 32 * c + j];
          v[j] = __float_as_uint(x);
        }
        tmem_st32(t_res + 32 * c, v);
        pack32(v, c);
      }
      publish(B_Y0);
      // ---- C. y1 = relu(acc + bl0) + y0 -> residual (fp32) and A operand (fp16) ----
      ptx::mbar_wait(bar(B_ACC0), it & 1);
      ptx::tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32], r[32];
        ptx::tmem_ld32(t_acc + 32 * c, v);
        ptx::tmem_ld32(t_res + 32 * c, r);
        ptx::tmem_ld_wait();
        const float4* b4 = reinterpret_cast<const float4*>(s_bl0 + col0 + 32 * c);
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 b = b4[j4];
          v[4 * j4 + 0] = __float_as_uint(fmaxf(__uint_as_float(v[4 * j4 + 0]) + b.x, 0.f) + __uint_as_float(r[4 * j4 + 0]));
          v[4 * j4 + 1] = __float_as_uint(fmaxf(__uint_as_float(v[4 * j4 + 1]) + b.y, 0.f) + __uint_as_float(r[4 * j4 + 1]));
          v[4 * j4 + 2] = __float_as_uint(fmaxf(__uint_as_float(v[4 * j4 + 2]) + b.z, 0.f) + __uint_as_float(r[4 * j4 + 2]));
          v[4 * j4 + 3] = __float_as_uint(fmaxf(__uint_as_float(v[4 * j4 + 3]) + b.w, 0.f) + __uint_as_float(r[4 * j4 + 3]));
        }
        tmem_st32(t_res + 32 * c, v);
        pack32(v, c);
      }
      publish(B_Y1);
      // the next step's rows stream in while GEMM 1 runs
      fetch(t + n_clusters);
      // ---- E. y2 = relu(acc + bl1) + y1; score = y2 . wd + bd ----
      ptx::mbar_wait(bar(B_ACC1), it & 1);
      ptx::tc_fence_after();
      float dot0 = 0.f, dot1 = 0.f;
#pragma unroll 1
      for (int c = 0; c < 8; ++c) {
        uint32_t v[16], r[16];
        ptx::tmem_ld16(t_acc + 16 * c, v);
        ptx::tmem_ld16(t_res + 16 * c, r);
        ptx::tmem_ld_wait();
        const float4* b4 = reinterpret_cast<const float4*>(s_bl1 + col0 + 16 * c);
        const float4* w4 = reinterpret_cast<const float4*>(s_wd + col0 + 16 * c);
#pragma unroll
        for (int j4 = 0; j4 < 4; ++j4) {
          const float4 b = b4[j4], w = w4[j4];
          dot0 = fmaf(fmaxf(__uint_as_float(v[4 * j4 + 0]) + b.x, 0.f) + __uint_as_float(r[4 * j4 + 0]), w.x, dot0);
          dot1 = fmaf(fmaxf(__uint_as_float(v[4 * j4 + 1]) + b.y, 0.f) + __uint_as_float(r[4 * j4 + 1]), w.y, dot1);
          dot0 = fmaf(fmaxf(__uint_as_float(v[4 * j4 + 2]) + b.z, 0.f) + __uint_as_float(r[4 * j4 + 2]), w.z, dot0);
          dot1 = fmaf(fmaxf(__uint_as_float(v[4 * j4 + 3]) + b.w, 0.f) + __uint_as_float(r[4 * j4 + 3]), w.w, dot1);
        }
      }
      const float dot = dot0 + dot1;
      if (ch == 1) s_dot[row] = dot;
      ptx::tc_fence_before();
      ptx::bar_sync(1, WORKERS);
      const int64_t prog = (int64_t)t * PAIR + rank * TILE + row;
      if (ch == 0 && prog < p.P) {
        float s = dot + s_dot[row] + bd;
        if (p.sigmoid) s = 1.f / (1.f + expf(-s));
        if (p.valid && !p.valid[prog]) s = -INFINITY;
        p.scores[prog] = s;
      }
    }
  } else if (rank == 0) {
    // ===== MMA issuer (leader CTA) =====
    const uint32_t idesc = ptx::umma_idesc_f16(256, 256);
    const uint64_t dWl0 = ptx::umma_desc_sw128(sbase + WL0_OFF), dWl1 = ptx::umma_desc_sw128(sbase + WL1_OFF);
    const uint64_t dA = ptx::umma_desc_sw128(sbase + A_OFF);
    auto kadv = [](int k) -> uint64_t { return (uint64_t)(((k >> 2) * ATOM + (k & 3) * 32) >> 4); };
    int it = 0;
    for (int t = cluster; t < p.T; t += n_clusters, ++it) {
      ptx::mbar_wait(bar(B_Y0), it & 1);
      ptx::tc_fence_after();
      if (ptx::elect_one()) {
#pragma unroll
        for (int k = 0; k < 16; ++k) ptx::umma_f16_2cta(tmem + 256, dA + kadv(k), dWl0 + kadv(k), idesc, k > 0);
        ptx::umma_commit_2cta(bar(B_ACC0), 3);
      }
      __syncwarp();
      ptx::mbar_wait(bar(B_Y1), it & 1);
      ptx::tc_fence_after();
      if (ptx::elect_one()) {
#pragma unroll
        for (int k = 0; k < 16; ++k) ptx::umma_f16_2cta(tmem + 256, dA + kadv(k), dWl1 + kadv(k), idesc, k > 0);
        ptx::umma_commit_2cta(bar(B_ACC1), 3);
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
          img[l][(size_t)r * (W_BYTES / 2) + sw128_off(row, k) / 2] = hh_clamp(w[(size_t)(r * TILE + row) * HID + k]);
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
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int clusters = std::max(1, std::min(sms / 2, prm.T));
  head_tc_kernel<<<2 * clusters, THREADS, SMEM_DYN, st>>>(prm);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

}  // namespace vaex
