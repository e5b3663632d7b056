// Fused MLP chains of the extent-VAE family on tcgen05 tensor cores:
//
//   decoder   z [n, L<=64] -> relu(W0) -> relu(W1) -> relu(W2) -> W3 -> x [n, D<=64]  (+ snap to int32 extents)
//             (VAE_feature_head.decode, utils/model.py:31-39,63-64; postprocess, vae_sch_extent.ipynb:1374-1383)
//   predictor z [n, L<=64] -> relu(W0) -> relu(W1) -> w2 . h + b2 -> cost [n]
//             (VAECostPredictor.predict_cost, utils/model.py:128-141,168-170; eval mode, Dropout = identity)
//
// One persistent CTA pair (cta_group::2) per 256 rows per step; all layer weights stay resident
// in shared memory (each CTA holds the half of every layer's output units: 160 KB for the decoder),
// the 64 KB A operand is rewritten in place by each layer's epilogue (bias + ReLU + fp16 pack),
// the accumulator lives in TMEM: activations never touch HBM -- per row the kernel reads 4*L
// bytes and writes 4*D (or 4).  Same worker layout as head_tc.cu: a thread owns one row and one
// half of the columns; warp 8 of the leader CTA issues the MMAs; hand-offs are two mbarriers.
#include <algorithm>
#include <vector>

#include "common.cuh"
#include "model.cuh"
#include "snap.cuh"
#include "tc_ptx.cuh"

namespace vaex {

namespace ch {
constexpr int TILE = 128, PAIR = 256, ATOM = 128 * 128, HID = 256, K0 = 64;
constexpr int WORKERS = 256, THREADS = 288;
constexpr int MAX_LAYERS = 4;
enum { B_IN = 0, B_ACC = 1 };
enum { MODE_VECTOR = 0, MODE_DOT = 1 };
constexpr int SMEM_MAX = 232448;

// byte offset of (row u, column k) in a K-major SWIZZLE_128B image with `rows` rows per 64-column atom
__host__ __device__ inline uint32_t sw_off(int rows, int u, int k) {
  return (uint32_t)((k >> 6) * rows * 128 + (u >> 3) * 1024 + (u & 7) * 128 + ((((k & 63) >> 3) ^ (u & 7)) << 4) + (k & 7) * 2);
}
}  // namespace ch

struct ChainParams {
  const float* X;            // [n, in_dim] fp32
  const __half* w;           // all layer images, rank-major: [2][w_bytes / 2]
  const __half* bias;        // fp16 biases of the MMA layers, concatenated
  float* out;                // MODE_VECTOR: [n, out_dim] fp32 (nullable); MODE_DOT: [n]
  int32_t* extents;          // MODE_VECTOR: [n, out_dim] int32 (nullable): fused snap
  const float* snap_mean;    // [out_dim]
  const float* snap_scale;
  const float* wd;           // MODE_DOT: [256] fp32
  const float* bd;           // [1]
  int64_t n;
  int32_t T, in_dim, out_dim, mode, snap_mode, n_mma;
  int32_t K[ch::MAX_LAYERS], N[ch::MAX_LAYERS], w_off[ch::MAX_LAYERS], b_off[ch::MAX_LAYERS];
  int32_t w_bytes, a_off, bias_off, bias_count, bar_off;
};

__global__ void smem_base_probe_kernel(uint32_t* out) {
  extern __shared__ uint8_t probe_smem[];
  if (threadIdx.x == 0) *out = ptx::smem_u32(probe_smem);
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(ch::THREADS, 1)
chain_tc_kernel(const ChainParams p) {
  using namespace ch;
  extern __shared__ uint8_t smem[];                 // no static shared memory: the base is 1024-aligned (checked at pack time)
  const uint32_t sbase = ptx::smem_u32(smem);
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const int cluster = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.bar_off);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + p.bar_off + 16);
  float* s_dot = reinterpret_cast<float*>(smem + p.a_off);      // MODE_DOT: the A operand is dead during the last epilogue
  const __half* s_bias = reinterpret_cast<const __half*>(smem + p.bias_off);
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto bar = [&](int i) { return bar0 + 8u * i; };

  if (threadIdx.x == 0) {
    ptx::mbar_init(bar(B_IN), 16);      // 8 worker warps x 2 CTAs
    ptx::mbar_init(bar(B_ACC), 1);
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc_2cta(ptx::smem_u32(tmem_slot), 256);
  {
    const uint4* g = reinterpret_cast<const uint4*>(p.w) + (size_t)rank * (p.w_bytes / 16);
    uint4* s = reinterpret_cast<uint4*>(smem);
    for (int i = threadIdx.x; i < p.w_bytes / 16; i += THREADS) s[i] = __ldg(g + i);
    __half* sb = reinterpret_cast<__half*>(smem + p.bias_off);
    for (int i = threadIdx.x; i < p.bias_count; i += THREADS) sb[i] = p.bias[i];
    uint4* a = reinterpret_cast<uint4*>(smem + p.a_off);   // K padding of the first layer stays zero
    for (int i = threadIdx.x; i < ATOM / 16; i += THREADS) a[i] = make_uint4(0, 0, 0, 0);
  }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp < 8) {
    // ===== workers: thread = (row, column half) =====
    const int q = warp & 3, chh = warp >> 2;
    const uint32_t lane_base = (uint32_t)(32 * q) << 16;
    const int row = 32 * q + lane;
    const uint32_t a_row = sbase + p.a_off + (row >> 3) * 1024 + (row & 7) * 128;
    const int sw = row & 7;
    const uint32_t t_acc = tmem + lane_base;
    float xin[32];                                   // this thread's half of the next input row

    auto fetch = [&](int t) {
      const int64_t r = (int64_t)t * PAIR + rank * TILE + row;
      const bool ok = t < p.T && r < p.n;
      const float* src = p.X + (ok ? r : 0) * p.in_dim;
      if (ok && (p.in_dim & 7) == 0 && ((reinterpret_cast<uintptr_t>(p.X) & 31) == 0)) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int c = 32 * chh + 8 * j;
          if (c < p.in_dim) {
            float4 a = ptx::ldg_stream_f4(src + c), b = ptx::ldg_stream_f4(src + c + 4);
            xin[8 * j] = a.x; xin[8 * j + 1] = a.y; xin[8 * j + 2] = a.z; xin[8 * j + 3] = a.w;
            xin[8 * j + 4] = b.x; xin[8 * j + 5] = b.y; xin[8 * j + 6] = b.z; xin[8 * j + 7] = b.w;
          } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) xin[8 * j + e] = 0.f;
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int c = 32 * chh + j;
          xin[j] = (ok && c < p.in_dim) ? __ldg(src + c) : 0.f;
        }
      }
    };
    // 32 fp32 values -> 64 B of the fp16 A operand, 32-column chunk cc of the row
    auto pack32 = [&](const float (&v)[32], int cc) {
      const uint32_t base = a_row + (cc >> 1) * ATOM;
#pragma unroll
      for (int m = 0; m < 4; ++m)
        ptx::sts_v4(base + ((((cc & 1) * 4 + m) ^ sw) << 4),
                    ptx::cvt_h2_sat(v[8 * m], v[8 * m + 1]), ptx::cvt_h2_sat(v[8 * m + 2], v[8 * m + 3]),
                    ptx::cvt_h2_sat(v[8 * m + 4], v[8 * m + 5]), ptx::cvt_h2_sat(v[8 * m + 6], v[8 * m + 7]));
    };
    auto publish = [&]() {
      ptx::fence_proxy_async();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster_relaxed(bar(B_IN), 0);
    };
    auto bias32 = [&](int boff, float (&b)[32]) {     // 32 consecutive fp16 biases (64-byte aligned)
      const uint4* src = reinterpret_cast<const uint4*>(s_bias + boff);
#pragma unroll
      for (int m = 0; m < 4; ++m) {
        const uint4 u = src[m];
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[e]));
          b[8 * m + 2 * e] = f.x;
          b[8 * m + 2 * e + 1] = f.y;
        }
      }
    };

    fetch(cluster);
    uint32_t use = 0;
    for (int t = cluster; t < p.T; t += n_clusters) {
      pack32(xin, chh);                              // z -> fp16, columns [32 chh, +32) of atom 0
      publish();
      for (int l = 0; l < p.n_mma; ++l) {
        const bool last = l + 1 == p.n_mma;
        if (last) fetch(t + n_clusters);             // the next rows stream in behind the last GEMM
        ptx::mbar_wait(bar(B_ACC), use & 1);
        ++use;
        ptx::tc_fence_after();
        if (!last || p.mode == MODE_DOT) {
          // hidden layer: 128 columns of this row, bias + ReLU (+ fp16 pack | + dot)
          float dot0 = 0.f, dot1 = 0.f;
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            uint32_t v[32];
            float b[32], y[32];
            ptx::tmem_ld32(t_acc + 128 * chh + 32 * c, v);
            bias32(p.b_off[l] + 128 * chh + 32 * c, b);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) y[j] = fmaxf(__uint_as_float(v[j]) + b[j], 0.f);
            if (!last) {
              pack32(y, 4 * chh + c);
            } else {
              const float4* w4 = reinterpret_cast<const float4*>(p.wd + 128 * chh + 32 * c);
#pragma unroll
              for (int j4 = 0; j4 < 8; ++j4) {
                const float4 w = __ldg(w4 + j4);
                dot0 = fmaf(y[4 * j4], w.x, dot0);
                dot1 = fmaf(y[4 * j4 + 1], w.y, dot1);
                dot0 = fmaf(y[4 * j4 + 2], w.z, dot0);
                dot1 = fmaf(y[4 * j4 + 3], w.w, dot1);
              }
            }
          }
          if (!last) {
            publish();
          } else {
            const float dot = dot0 + dot1;
            if (chh == 1) s_dot[row] = dot;
            ptx::tc_fence_before();
            ptx::bar_sync(1, WORKERS);
            const int64_t r = (int64_t)t * PAIR + rank * TILE + row;
            if (chh == 0 && r < p.n) p.out[r] = dot + s_dot[row] + __ldg(p.bd);
            ptx::bar_sync(1, WORKERS);               // s_dot aliases the A operand the next step rewrites
          }
        } else {
          // output layer (N = 64): 32 columns of this row -> fp32 x and / or snapped extents
          uint32_t v[32];
          float b[32];
          ptx::tmem_ld32(t_acc + 32 * chh, v);
          bias32(p.b_off[l] + 32 * chh, b);
          ptx::tmem_ld_wait();
          const int64_t r = (int64_t)t * PAIR + rank * TILE + row;
          if (r < p.n) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int c = 32 * chh + j;
              if (c < p.out_dim) {
                const float x = __uint_as_float(v[j]) + b[j];
                if (p.out) p.out[r * p.out_dim + c] = x;
                if (p.extents) p.extents[r * p.out_dim + c] = snap_one(x, __ldg(p.snap_mean + c), __ldg(p.snap_scale + c), p.snap_mode);
              }
            }
          }
          ptx::tc_fence_before();
        }
      }
    }
  } else if (rank == 0) {
    // ===== MMA issuer (leader CTA) =====
    const uint64_t dA = ptx::umma_desc_sw128(sbase + p.a_off);
    uint32_t use = 0;
    for (int t = cluster; t < p.T; t += n_clusters) {
      for (int l = 0; l < p.n_mma; ++l) {
        ptx::mbar_wait(bar(B_IN), use & 1);
        ++use;
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          const uint32_t idesc = ptx::umma_idesc_f16(256, p.N[l]);
          const uint64_t dW = ptx::umma_desc_sw128(sbase + p.w_off[l]);
          const int w_atom = (p.N[l] / 2) * 128;
          for (int k = 0; k < p.K[l] / 16; ++k)
            ptx::umma_f16_2cta(tmem, dA + (uint64_t)(((k >> 2) * ATOM + (k & 3) * 32) >> 4),
                               dW + (uint64_t)(((k >> 2) * w_atom + (k & 3) * 32) >> 4), idesc, k > 0);
          ptx::umma_commit_2cta(bar(B_ACC), 3);
        }
        __syncwarp();
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  if (warp == 1) ptx::tmem_dealloc_2cta(tmem, 256);
}

// ---------------------------------------------------------------------------------------------
static __half c_clamp(float x) { return __float2half_rn(std::min(std::max(x, -65504.f), 65504.f)); }

struct ChainLayer { const std::vector<float>* w; const std::vector<float>* b; int out, in; };

// Packs `layers` (row-major [out, in] fp32) into the per-rank swizzled images + fp16 biases and fills the
// shape part of ChainParams.  The last entry of a MODE_DOT chain is the scalar head (kept fp32, not an MMA).
int chain_tc_pack(ChainPack& cp, const std::vector<ChainLayer>& layers, int mode, cudaStream_t st) {
  using namespace ch;
  cp.ready = false;
  const int n_mma = mode == MODE_DOT ? (int)layers.size() - 1 : (int)layers.size();
  if (n_mma < 1 || n_mma > MAX_LAYERS) return VAEX_OK;
  ChainParams& P = *new (cp.params_storage) ChainParams();
  static_assert(sizeof(ChainParams) <= sizeof(cp.params_storage), "ChainPack::params_storage too small");
  P.mode = mode;
  P.n_mma = n_mma;
  int woff = 0, boff = 0;
  for (int l = 0; l < n_mma; ++l) {
    const ChainLayer& L = layers[l];
    const int K = l == 0 ? K0 : HID;
    const bool out_layer = mode == MODE_VECTOR && l == n_mma - 1;
    const int N = out_layer ? 64 : HID;
    if ((l == 0 ? L.in > K0 : L.in != HID) || (out_layer ? L.out > 64 : L.out != HID)) return VAEX_OK;   // shape outside this kernel
    P.K[l] = K; P.N[l] = N; P.w_off[l] = woff; P.b_off[l] = boff;
    woff += (N / 2) * K * 2;
    boff += N;
  }
  P.in_dim = layers[0].in;
  P.out_dim = mode == MODE_VECTOR ? layers[n_mma - 1].out : 1;
  P.w_bytes = woff;
  P.a_off = woff;
  P.bias_off = P.a_off + 4 * ATOM;
  P.bias_count = boff;
  P.bar_off = (P.bias_off + boff * 2 + 15) & ~15;
  cp.smem_bytes = P.bar_off + 32;
  if (cp.smem_bytes > SMEM_MAX) return VAEX_OK;
  // the kernel relies on a 1024-byte aligned dynamic shared-memory base (no static shared memory)
  {
    uint32_t* d = nullptr;
    uint32_t h = 1;
    VAEX_CUDA(cudaMalloc(&d, 4));
    VAEX_CUDA(cudaFuncSetAttribute(smem_base_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, cp.smem_bytes));
    smem_base_probe_kernel<<<1, 32, cp.smem_bytes, st>>>(d);
    VAEX_CUDA(cudaMemcpyAsync(&h, d, 4, cudaMemcpyDeviceToHost, st));
    VAEX_CUDA(cudaStreamSynchronize(st));
    cudaFree(d);
    if (h & 1023u) return VAEX_OK;                     // stay on the SIMT path
  }
  std::vector<__half> img((size_t)woff, __float2half_rn(0.f));   // 2 ranks x (woff bytes / 2) halves
  std::vector<__half> bias((size_t)boff, __float2half_rn(0.f));
  for (int l = 0; l < n_mma; ++l) {
    const ChainLayer& L = layers[l];
    const int nh = P.N[l] / 2;
    for (int r = 0; r < 2; ++r)
      for (int u = 0; u < nh; ++u) {
        const int unit = r * nh + u;
        if (unit >= L.out) continue;
        __half* dst = img.data() + (size_t)r * (woff / 2) + P.w_off[l] / 2;
        for (int k = 0; k < L.in; ++k) dst[sw_off(nh, u, k) / 2] = c_clamp((*L.w)[(size_t)unit * L.in + k]);
      }
    for (int j = 0; j < L.out; ++j) bias[P.b_off[l] + j] = c_clamp((*L.b)[j]);
  }
  if (cp.w) cudaFree(cp.w);
  if (cp.bias) cudaFree(cp.bias);
  cp.w = nullptr; cp.bias = nullptr;
  VAEX_CUDA(cudaMalloc(&cp.w, img.size() * sizeof(__half)));
  VAEX_CUDA(cudaMalloc(&cp.bias, std::max<size_t>(bias.size(), 1) * sizeof(__half)));
  VAEX_CUDA(cudaMemcpyAsync(cp.w, img.data(), img.size() * sizeof(__half), cudaMemcpyHostToDevice, st));
  VAEX_CUDA(cudaMemcpyAsync(cp.bias, bias.data(), bias.size() * sizeof(__half), cudaMemcpyHostToDevice, st));
  VAEX_CUDA(cudaStreamSynchronize(st));
  VAEX_CUDA(cudaFuncSetAttribute(chain_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_MAX));
  cp.ready = true;
  return VAEX_OK;
}

int chain_tc_run(const ChainPack& cp, const float* X, int64_t n, float* out, int32_t* extents, const float* mean,
                 const float* scale, int snap_mode, const float* wd, const float* bd, cudaStream_t st) {
  using namespace ch;
  VAEX_CHECK(cp.ready, "tcgen05 chain not packed");
  if (n == 0) return VAEX_OK;
  ChainParams P = *reinterpret_cast<const ChainParams*>(cp.params_storage);
  P.X = X; P.w = cp.w; P.bias = cp.bias; P.out = out; P.extents = extents; P.snap_mean = mean; P.snap_scale = scale;
  P.snap_mode = snap_mode; P.wd = wd; P.bd = bd; P.n = n; P.T = (int)cdiv(n, PAIR);
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int clusters = std::max(1, std::min(sms / 2, P.T));
  chain_tc_kernel<<<2 * clusters, THREADS, cp.smem_bytes, st>>>(P);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

}  // namespace vaex
