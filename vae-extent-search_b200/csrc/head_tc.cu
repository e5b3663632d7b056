// Residual head of the SegmentSumMLP on tcgen05 tensor cores:
//
//   y0 = norm(S)                       (Identity, or the folded eval-mode BatchNorm affine)
//   y1 = relu(y0 Wl0^T + bl0) + y0
//   y2 = relu(y1 Wl1^T + bl1) + y1
//   score = y2 . wd + bd               (+ sigmoid, + -inf for invalid programs)
//   (SegmentSumMLPModule.forward, python/tvm/auto_scheduler/cost_model/mlp_model.py:210-221)
//
// One CTA pair (cta_group::2) per 256 programs: each CTA owns 128 programs (TMEM lanes) and half
// of the output units of both 256x256 weights (resident, 128 KB fp16).  The A operand buffer
// (64 KB) first holds fp16(y0), then fp16(y1); fp32 y1 is parked in TMEM (over the dead layer-0
// accumulator) for the second residual, so no intermediate touches HBM.  This kernel is ~7 % of
// the path's FLOPs; phases are separated by cluster barriers instead of a warp-specialised pipe.
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
constexpr int THREADS = 256;

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

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(hd::THREADS, 1)
head_tc_kernel(const HeadParams p) {
  using namespace hd;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[2];
  __shared__ uint32_t tmem_base_slot;
  __shared__ float s_bl0[HID], s_bl1[HID], s_wd[HID], s_ns[HID], s_nb[HID];
  __shared__ float s_dot[TILE];

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t sbase = ptx::smem_u32(smem);
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const int cluster = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const uint32_t bar_l0 = ptx::smem_u32(&bars[0]), bar_l1 = ptx::smem_u32(&bars[1]);
  const bool has_norm = p.nscale != nullptr;

  if (threadIdx.x == 0) {
    ptx::mbar_init(bar_l0, 1);
    ptx::mbar_init(bar_l1, 1);
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

  const uint32_t idesc = ptx::umma_idesc_f16(256, 256);
  const uint64_t dWl0 = ptx::umma_desc_sw128(sbase + WL0_OFF), dWl1 = ptx::umma_desc_sw128(sbase + WL1_OFF);
  const uint64_t dA = ptx::umma_desc_sw128(sbase + A_OFF);
  auto kadv = [](int k) -> uint64_t { return (uint64_t)(((k >> 2) * ATOM + (k & 3) * 32) >> 4); };

  const int q = warp & 3, ch = warp >> 2;          // TMEM lane quarter, column half of the epilogues
  const uint32_t lane_base = (uint32_t)(32 * q) << 16;
  const int erow = 32 * q + lane;                  // program row (TMEM lane) of this thread in the epilogues
  const uint32_t a_row = sbase + A_OFF + (erow >> 3) * 1024 + (erow & 7) * 128;
  const int sw = erow & 7;
  const float bd = p.bd[0];

  int it = 0;
  for (int t = cluster; t < p.T; t += n_clusters, ++it) {
    const int64_t prog0 = (int64_t)t * PAIR + rank * TILE;
    // ---- 1. y0 = norm(S) -> fp16 A operand: warp w converts rows [16w, 16w+16), 8 B stores ----
#pragma unroll 4
    for (int r = 0; r < 16; ++r) {
      const int row = warp * 16 + r;
      const int64_t prog = prog0 + row;
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        const int c = (lane + 32 * hh) * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (prog < p.P) v = __ldg(reinterpret_cast<const float4*>(p.S + prog * HID + c));
        if (has_norm) {
          v.x = v.x * s_ns[c] + s_nb[c]; v.y = v.y * s_ns[c + 1] + s_nb[c + 1];
          v.z = v.z * s_ns[c + 2] + s_nb[c + 2]; v.w = v.w * s_ns[c + 3] + s_nb[c + 3];
        }
        ptx::sts_v2(sbase + A_OFF + sw128_off(row, c), ptx::cvt_h2_sat(v.x, v.y), ptx::cvt_h2_sat(v.z, v.w));
      }
    }
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    ptx::cluster_sync();
    // ---- 2. layer l0 ----
    if (rank == 0 && threadIdx.x == 0) {
      ptx::tc_fence_after();
#pragma unroll
      for (int k = 0; k < 16; ++k) ptx::umma_f16_2cta(tmem, dA + kadv(k), dWl0 + kadv(k), idesc, k > 0);
      ptx::umma_commit_2cta(bar_l0, 3);
    }
    ptx::mbar_wait(bar_l0, it & 1);
    ptx::tc_fence_after();
    // ---- 3. y1 = relu(acc0 + bl0) + y0 -> fp32 back into TMEM, fp16 into A ----
    const int64_t eprog = prog0 + erow;
    const float* srow = p.S + (eprog < p.P ? eprog : 0) * HID;
#pragma unroll 1
    for (int c4 = 0; c4 < 4; ++c4) {
      const int c = ch * 4 + c4;                    // 32-column chunk
      uint32_t v[32];
      ptx::tmem_ld32(tmem + lane_base + 32 * c, v);
      ptx::tmem_ld_wait();
#pragma unroll
      for (int j4 = 0; j4 < 8; ++j4) {
        float4 y0 = __ldg(reinterpret_cast<const float4*>(srow + 32 * c + 4 * j4));
        const float yy[4] = {y0.x, y0.y, y0.z, y0.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int col = 32 * c + 4 * j4 + e;
          float r0 = yy[e];
          if (has_norm) r0 = r0 * s_ns[col] + s_nb[col];
          v[4 * j4 + e] = __float_as_uint(fmaxf(__uint_as_float(v[4 * j4 + e]) + s_bl0[col], 0.f) + r0);
        }
      }
      tmem_st32(tmem + lane_base + 32 * c, v);
      const uint32_t base = a_row + (c >> 1) * ATOM;
#pragma unroll
      for (int m = 0; m < 4; ++m)
        ptx::sts_v4(base + ((((c & 1) * 4 + m) ^ sw) << 4),
                    ptx::cvt_h2_sat(__uint_as_float(v[8 * m]), __uint_as_float(v[8 * m + 1])),
                    ptx::cvt_h2_sat(__uint_as_float(v[8 * m + 2]), __uint_as_float(v[8 * m + 3])),
                    ptx::cvt_h2_sat(__uint_as_float(v[8 * m + 4]), __uint_as_float(v[8 * m + 5])),
                    ptx::cvt_h2_sat(__uint_as_float(v[8 * m + 6]), __uint_as_float(v[8 * m + 7])));
    }
    tmem_st_wait();
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    ptx::cluster_sync();
    // ---- 4. layer l1 ----
    if (rank == 0 && threadIdx.x == 0) {
      ptx::tc_fence_after();
#pragma unroll
      for (int k = 0; k < 16; ++k) ptx::umma_f16_2cta(tmem + 256, dA + kadv(k), dWl1 + kadv(k), idesc, k > 0);
      ptx::umma_commit_2cta(bar_l1, 3);
    }
    ptx::mbar_wait(bar_l1, it & 1);
    ptx::tc_fence_after();
    // ---- 5. y2 = relu(acc1 + bl1) + y1; score = y2 . wd + bd ----
    float dot = 0.f;
#pragma unroll 1
    for (int c4 = 0; c4 < 4; ++c4) {
      const int c = ch * 4 + c4;
      uint32_t v[32], y1[32];
      ptx::tmem_ld32(tmem + lane_base + 256 + 32 * c, v);
      ptx::tmem_ld32(tmem + lane_base + 32 * c, y1);
      ptx::tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float y2 = fmaxf(__uint_as_float(v[j]) + s_bl1[32 * c + j], 0.f) + __uint_as_float(y1[j]);
        dot = fmaf(y2, s_wd[32 * c + j], dot);
      }
    }
    if (ch == 1) s_dot[erow] = dot;
    ptx::tc_fence_before();
    __syncthreads();
    if (ch == 0 && eprog < p.P) {
      float s = dot + s_dot[erow] + bd;
      if (p.sigmoid) s = 1.f / (1.f + expf(-s));
      if (p.valid && !p.valid[eprog]) s = -INFINITY;
      p.scores[eprog] = s;
    }
    // the next tile's first cluster barrier orders these TMEM / smem reads before its MMAs
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
