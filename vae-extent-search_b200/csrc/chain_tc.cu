// Fused MLP chains of the extent-VAE family on tcgen05 tensor cores:
//
//   decoder   z [n, L<=64] -> relu(W0) -> relu(W1) -> relu(W2) -> W3 -> x [n, D<=64]  (+ snap to int32 extents)
//             (VAE_feature_head.decode, utils/model.py:31-39,63-64; postprocess, vae_sch_extent.ipynb:1374-1383)
//   predictor z [n, L<=64] -> relu(W0) -> relu(W1) -> w2 . h + b2 -> cost [n]
//             (VAECostPredictor.predict_cost, utils/model.py:128-141,168-170; eval mode, Dropout = identity)
//
// One persistent CTA pair (cta_group::2) per 256 rows per step; all layer weights stay resident
// in shared memory (each CTA holds half of every layer's output units: 160 KB for the decoder),
// the 64 KB A operand is rewritten in place by each layer's epilogue (bias + ReLU + fp16 pack),
// the accumulators live in TMEM: activations never touch HBM -- per row the kernel reads 4*L
// bytes and writes 4*D (or 4).
//
// Pipelining inside a tile (v2).  A 256-wide layer is issued as four MMA pieces, K-half major:
//   (n-half 0, k 0..127) (n-half 1, k 0..127) (n-half 0, k 128..255) -> commit ACC[0]
//   (n-half 1, k 128..255) -> commit ACC[1]
// The epilogue of n-half h produces exactly the A atoms of k-half h of the next layer, so the
// next layer's first two pieces run on the tensor pipe while the workers are still draining
// n-half 1 of this one (accumulators are double-buffered: 2 x 256 TMEM columns); pieces of k-half
// h wait for mbarrier IN[h].  In-place A is safe: an epilogue only overwrites atoms whose readers
// (all earlier pieces) are covered by the commit it waited for.  z of the next tile is packed into
// atom 3 (the last one the layer-0 epilogue rewrites) right after the last layer's commit, so
// layer 0 of tile t+1 runs behind the output epilogue of tile t.
// Workers: 16 warps, thread = (row, 32-column group of the n-half); warp 16 of the leader CTA issues.
#include <cstdlib>
#include <algorithm>
#include <vector>

#include "common.cuh"
#include "model.cuh"
#include "philox.cuh"
#include "snap.cuh"
#include "tc_ptx.cuh"

namespace vaex {

namespace ch {
constexpr int TILE = 128, PAIR = 256, ATOM = 128 * 128, HID = 256;   // layer 0's K is 64 or 128 (one or two operand atoms)
constexpr int WORKER_WARPS = 16, WORKERS = 32 * WORKER_WARPS, THREADS = WORKERS + 32;
constexpr int MAX_LAYERS = 4;
enum { B_IN0 = 0, B_IN1 = 1, B_ACC0 = 2, B_ACC1 = 3 };
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
  const float* bias;         // fp32 biases of the MMA layers, concatenated (zero padded to N) [+ MODE_DOT: wd[256], bd]
  float* out;                // MODE_VECTOR: [n, out_dim] fp32 (nullable); MODE_DOT: [n]
  int32_t* extents;          // MODE_VECTOR: [n, out_dim] int32 (nullable): fused snap
  const float* snap_mean;    // [out_dim]
  const float* snap_scale;
  const float* wd;           // MODE_DOT: [256] fp32
  const float* bd;           // [1]
  int64_t n;
  int32_t T, in_dim, out_dim, mode, snap_mode, n_mma;
  int32_t K[ch::MAX_LAYERS], N[ch::MAX_LAYERS], w_off[ch::MAX_LAYERS], b_off[ch::MAX_LAYERS];
  int32_t w_bytes, a_off, bias_off, bias_count, wd_off, bar_off;
  int32_t ones_col;          // >= 0: input columns ones_col, ones_col + 1 are the constant 1 that carries layer 0's bias (hi, lo) in W0
  float b_scale[ch::MAX_LAYERS];   // 1: add the stored bias; 0: the bias rides in the MMA
  // MODE_DOT with Dropout(p) after the first ReLU (mc_predict, utils/model.py:241-273): mc_T >= 1 passes per tile
  // share layer 0; pass t keeps element (row, c) iff drop_mask[t][row][c] != 0, else iff the Philox draw of
  // dropout_kernel (elementwise.cu) for seed drop_seed + t says so.  mc_T == 1: out = cost; mc_T > 1: out = mean over
  // the passes, var_out = unbiased variance.
  // encoder with both latent heads in one pass: the output layer is [fc_mu; fc_logvar] stacked (N = 2 * out_dim), columns
  // [out_dim, 2 out_dim) go to out2; W0 is then NOT resident -- its one-atom image lives in atom 2 of the operand (free
  // while layer 0 runs, overwritten by layer 0's second epilogue) and is re-stored from registers for every tile
  float* out2;
  const __half* w0_img;      // [2][16 KB] per-rank images of layer 0 when w0_in_a
  int32_t w0_in_a;
  int32_t bias_global;       // the fp32 biases (+ scalar head) are read from `bias` in global memory (L1) instead of a shared copy
  const uint8_t* drop_mask;  // [mc_T][n, 256] or null
  float* var_out;            // [n] or null
  uint64_t drop_seed;
  int64_t drop_chunk;        // rows per Philox chunk (the SIMT path re-seeds every VAE_CHUNK rows)
  float drop_p, drop_inv;    // p and 1 / (1 - p)
  int32_t mc_T;
};

__global__ void smem_base_probe_kernel(uint32_t* out) {
  extern __shared__ uint8_t probe_smem[];
  if (threadIdx.x == 0) *out = ptx::smem_u32(probe_smem);
}

template <bool BIAS_G, int KZ>     // KZ: atoms of the layer-0 operand (1: K = 64, 2: K = 128)
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(ch::THREADS, 1)
chain_tc_kernel(const ChainParams p) {
  using namespace ch;
  extern __shared__ uint8_t smem[];                 // no static shared memory: the base is 1024-aligned (checked at pack time)
  const uint32_t sbase = ptx::smem_u32(smem);
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const int cluster = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.bar_off);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + p.bar_off + 32);
  float* s_dot = reinterpret_cast<float*>(smem + p.bar_off + 64);     // MODE_DOT: [2][3][128] partial dots (double-buffered over MC passes)
  const float* s_bias = BIAS_G ? p.bias : reinterpret_cast<const float*>(smem + p.bias_off);
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto bar = [&](int i) { return bar0 + 8u * i; };

  if (threadIdx.x == 0) {
    ptx::mbar_init(bar(B_IN0), 2 * WORKER_WARPS);      // worker warps x 2 CTAs
    ptx::mbar_init(bar(B_IN1), 2 * WORKER_WARPS);
    ptx::mbar_init(bar(B_ACC0), 1);
    ptx::mbar_init(bar(B_ACC1), 1);
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc_2cta(ptx::smem_u32(tmem_slot), 512);
  {
    const uint4* g = reinterpret_cast<const uint4*>(p.w) + (size_t)rank * (p.w_bytes / 16);
    uint4* s = reinterpret_cast<uint4*>(smem);
    for (int i = threadIdx.x; i < p.w_bytes / 16; i += THREADS) s[i] = __ldg(g + i);
    if (!BIAS_G) {
      float* sb = reinterpret_cast<float*>(smem + p.bias_off);
      for (int i = threadIdx.x; i < p.bias_count; i += THREADS) sb[i] = p.bias[i];
    }
  }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp < WORKER_WARPS) {
    // ===== workers: thread = (row, column group cg of 4) =====
    const int q = warp & 3, cg = warp >> 2;
    const int row = 32 * q + lane;
    const uint32_t a_row = sbase + p.a_off + (row >> 3) * 1024 + (row & 7) * 128;
    const int sw = row & 7;
    const uint32_t t_lane = tmem + ((uint32_t)(32 * q) << 16);
    constexpr int kz = KZ;
    uint32_t xh[8 * KZ];                               // this thread's 16 kz columns of the next input row, packed fp16

    auto fetch = [&](int t) {
      const int64_t r = (int64_t)t * PAIR + rank * TILE + row;
      const bool ok = t < p.T && r < p.n;
      const float* src = p.X + (ok ? r : 0) * p.in_dim;
      const bool vec = ok && (p.in_dim & 7) == 0 && ((reinterpret_cast<uintptr_t>(p.X) & 31) == 0);
#pragma unroll
      for (int j = 0; j < 2 * kz; ++j) {
        {
          const int c = 16 * kz * cg + 8 * j;
          float f[8];
          if (vec) {                                 // full 32 B sectors per lane
            if (c < p.in_dim) {
              const ptx::F8 v = ptx::ldg_stream_f8(src + c);
#pragma unroll
              for (int e = 0; e < 8; ++e) f[e] = v.v[e];
            } else {
#pragma unroll
              for (int e = 0; e < 8; ++e) f[e] = 0.f;
            }
          } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] = (ok && c + e < p.in_dim) ? __ldg(src + c + e) : 0.f;
          }
          if (p.ones_col >= 0) {
#pragma unroll
            for (int e = 0; e < 8; ++e)
              if ((unsigned)(c + e - p.ones_col) < 2u) f[e] = 1.f;
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) xh[4 * j + e] = ptx::cvt_h2_sat(f[2 * e], f[2 * e + 1]);
        }
      }
    };
    // z: 16 kz halves -> 2 kz 16-byte chunks of the layer-0 A operand (the last kz atoms)
    auto pack_z = [&]() {
#pragma unroll
      for (int m = 0; m < 2 * kz; ++m) {
          const int gc = 2 * kz * cg + m;
          ptx::sts_v4(a_row + (4 - kz + (gc >> 3)) * ATOM + (((gc & 7) ^ sw) << 4), xh[4 * m], xh[4 * m + 1], xh[4 * m + 2], xh[4 * m + 3]);
        }
    };
    uint4 w0r[2] = {make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};      // this thread's 32 B of the W0 image (w0_in_a)
    if (p.w0_in_a) {
      const uint4* g = reinterpret_cast<const uint4*>(p.w0_img) + (size_t)rank * (ATOM / 16) + 2 * threadIdx.x;
      w0r[0] = __ldg(g);
      w0r[1] = __ldg(g + 1);
    }
    auto store_w0 = [&]() {
      if (p.w0_in_a) {
        const uint32_t dst = sbase + p.a_off + 2 * ATOM + 32u * threadIdx.x;
        ptx::sts_v4(dst, w0r[0].x, w0r[0].y, w0r[0].z, w0r[0].w);
        ptx::sts_v4(dst + 16, w0r[1].x, w0r[1].y, w0r[1].z, w0r[1].w);
      }
    };
    auto publish = [&](int b) {
      ptx::fence_proxy_async();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster_relaxed(bar(b), 0);
    };
    const bool vec_ok = (p.out_dim & 3) == 0 &&
        ((reinterpret_cast<uintptr_t>(p.out) | reinterpret_cast<uintptr_t>(p.extents) | reinterpret_cast<uintptr_t>(p.snap_mean) |
          reinterpret_cast<uintptr_t>(p.snap_scale)) & 15) == 0;
    const bool vec8_ok = (p.out_dim & 7) == 0 && ((reinterpret_cast<uintptr_t>(p.out) | reinterpret_cast<uintptr_t>(p.extents)) & 31) == 0 &&
        ((reinterpret_cast<uintptr_t>(p.snap_mean) | reinterpret_cast<uintptr_t>(p.snap_scale)) & 15) == 0;
    uint32_t acc_use0 = 0, acc_use1 = 0, lc = 0;
    fetch(cluster);
    if (cluster < p.T) {
      { pack_z(); store_w0(); }
      publish(B_IN0);
    }
    // ---- MC-dropout passes of the predictor (mc_T >= 1): keep bits of this thread's 32 columns of n-half nh ----
    auto keep_bits = [&](int pass, int64_t r, int nh) -> uint32_t {
      const int col = 128 * nh + 32 * cg;
      uint32_t bits = 0;
      if (r >= p.n) return 0u;
      if (p.drop_mask) {
        const uint8_t* mk = p.drop_mask + ((int64_t)pass * p.n + r) * HID + col;
        if ((reinterpret_cast<uintptr_t>(mk) & 15) == 0) {
#pragma unroll
          for (int h2 = 0; h2 < 2; ++h2) {
            const uint4 w = __ldg(reinterpret_cast<const uint4*>(mk) + h2);
            const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
              for (int e = 0; e < 4; ++e)
                bits |= (uint32_t)(((ww[j] >> (8 * e)) & 0xffu) != 0u) << (16 * h2 + 4 * j + e);
          }
        } else {
#pragma unroll 4
          for (int e = 0; e < 32; ++e) bits |= (uint32_t)(mk[e] != 0) << e;
        }
      } else {
        const int64_t chunk = r / p.drop_chunk;
        const PhiloxDrop gen(p.drop_seed + (uint64_t)pass + 0x9E3779B97F4A7C15ull * (uint64_t)chunk);
        const uint64_t q0 = (uint64_t)(((r - chunk * p.drop_chunk) * HID + col) >> 3);
        const uint32_t thr = dropout_threshold(p.drop_p);
#pragma unroll
        for (int j = 0; j < 4; ++j) {                      // eight 16-bit draws per call (dropout_kernel's layout)
          const uint4 u = gen(q0 + j, 0);
          const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            bits |= (uint32_t)((w[k] & 0xffffu) >= thr) << (8 * j + 2 * k);
            bits |= (uint32_t)((w[k] >> 16) >= thr) << (8 * j + 2 * k + 1);
          }
        }
      }
      return bits;
    };
    // masked copy of the packed layer-0 activations -> A operand k-half nh (register i holds columns 2i, 2i + 1)
    auto store_masked = [&](const uint32_t (&hr)[16], uint32_t bits, int nh) {
      const int cc = 4 * nh + cg;
      const uint32_t base = a_row + (cc >> 1) * ATOM;
      uint32_t o[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const uint32_t lo = 0u - ((bits >> (2 * i)) & 1u), hi = 0u - ((bits >> (2 * i + 1)) & 1u);
        o[i] = hr[i] & ((lo & 0xffffu) | (hi << 16));
      }
#pragma unroll
      for (int m = 0; m < 4; ++m)
        ptx::sts_v4(base + ((((cc & 1) * 4 + m) ^ sw) << 4), o[4 * m], o[4 * m + 1], o[4 * m + 2], o[4 * m + 3]);
    };
    for (int t = cluster; t < p.T; t += n_clusters) {
      const int64_t r = (int64_t)t * PAIR + rank * TILE + row;
      if (p.mc_T > 0) {
        // layer 0 once: relu(acc + b0) / (1 - p), packed fp16, kept in registers for every pass of this tile
        uint32_t hreg[2][16];
        uint32_t nb0 = keep_bits(0, r, 0), nb1 = keep_bits(0, r, 1);
        {
          const uint32_t t_acc = t_lane + 256u * (lc & 1);
          ++lc;
          const float bs = p.b_scale[0], inv = p.drop_inv;
#pragma unroll
          for (int nh = 0; nh < 2; ++nh) {
            const int col = 128 * nh + 32 * cg;
            const float4* b = reinterpret_cast<const float4*>(s_bias + p.b_off[0] + col);
            if (nh == 0) { ptx::mbar_wait(bar(B_ACC0), acc_use0 & 1); ++acc_use0; }
            else         { ptx::mbar_wait(bar(B_ACC1), acc_use1 & 1); ++acc_use1; }
            ptx::tc_fence_after();
            uint32_t v[32];
            ptx::tmem_ld32(t_acc + col, v);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int m = 0; m < 8; ++m) {
              const float4 bm = b[m];
              hreg[nh][2 * m] = ptx::cvt_h2_relu_sat(fmaf(bs, bm.x, __uint_as_float(v[4 * m])) * inv,
                                                     fmaf(bs, bm.y, __uint_as_float(v[4 * m + 1])) * inv);
              hreg[nh][2 * m + 1] = ptx::cvt_h2_relu_sat(fmaf(bs, bm.z, __uint_as_float(v[4 * m + 2])) * inv,
                                                         fmaf(bs, bm.w, __uint_as_float(v[4 * m + 3])) * inv);
            }
            store_masked(hreg[nh], nh == 0 ? nb0 : nb1, nh);
            publish(nh == 0 ? B_IN0 : B_IN1);
          }
        }
        float mean = 0.f, m2 = 0.f;
        for (int pass = 0; pass < p.mc_T; ++pass) {
          const bool lastp = pass + 1 == p.mc_T;
          const uint32_t t_acc = t_lane + 256u * (lc & 1);
          ++lc;
          if (lastp) {
            fetch(t + n_clusters);
            const int64_t r2 = (int64_t)(t + 2 * n_clusters) * PAIR + rank * TILE + row;
            if (32 * cg < p.in_dim && r2 < p.n) ptx::prefetch_l2(p.X + r2 * p.in_dim + 32 * cg);
          } else {                                   // the next pass's keep bits are computed behind this pass's GEMM
            nb0 = keep_bits(pass + 1, r, 0);
            nb1 = keep_bits(pass + 1, r, 1);
          }
          float dot0 = 0.f, dot1 = 0.f;
#pragma unroll
          for (int nh = 0; nh < 2; ++nh) {
            const int col = 128 * nh + 32 * cg;
            const float4* b = reinterpret_cast<const float4*>(s_bias + p.b_off[1] + col);
            const float4* w4 = reinterpret_cast<const float4*>(s_bias + p.wd_off + col);
            if (nh == 0) { ptx::mbar_wait(bar(B_ACC0), acc_use0 & 1); ++acc_use0; }
            else         { ptx::mbar_wait(bar(B_ACC1), acc_use1 & 1); ++acc_use1; }
            ptx::tc_fence_after();
            uint32_t v[32];
            ptx::tmem_ld32(t_acc + col, v);
            if (!lastp) {                            // k-half nh is free: every piece that reads it is behind this commit
              store_masked(hreg[nh], nh == 0 ? nb0 : nb1, nh);
              ptx::tmem_ld_wait();
              publish(nh == 0 ? B_IN0 : B_IN1);
            } else {
              ptx::tmem_ld_wait();
            }
#pragma unroll
            for (int m = 0; m < 8; ++m) {
              const float4 w = w4[m], bm = b[m];
              dot0 = fmaf(fmaxf(__uint_as_float(v[4 * m]) + bm.x, 0.f), w.x, dot0);
              dot1 = fmaf(fmaxf(__uint_as_float(v[4 * m + 1]) + bm.y, 0.f), w.y, dot1);
              dot0 = fmaf(fmaxf(__uint_as_float(v[4 * m + 2]) + bm.z, 0.f), w.z, dot0);
              dot1 = fmaf(fmaxf(__uint_as_float(v[4 * m + 3]) + bm.w, 0.f), w.w, dot1);
            }
          }
          if (lastp) {
            ptx::tc_fence_before();
            if (t + n_clusters < p.T) {
              { pack_z(); store_w0(); }
              publish(B_IN0);
            }
          }
          const float dot = dot0 + dot1;
          float* sd = s_dot + (pass & 1) * 3 * TILE;
          if (cg) sd[(cg - 1) * TILE + row] = dot;
          ptx::bar_sync(1, WORKERS);
          if (cg == 0) {                             // Welford over the passes (unbiased variance, torch.var)
            const float c = ((dot + sd[row]) + (sd[TILE + row] + sd[2 * TILE + row])) + s_bias[p.wd_off + HID];
            const float d = c - mean;
            mean += d / (float)(pass + 1);
            m2 = fmaf(d, c - mean, m2);
          }
        }
        if (cg == 0 && r < p.n) {
          p.out[r] = mean;
          if (p.var_out) p.var_out[r] = m2 / (float)(p.mc_T - 1);
        }
        continue;
      }
      for (int l = 0; l < p.n_mma; ++l) {
        const bool last = l + 1 == p.n_mma;
        const uint32_t t_acc = t_lane + 256u * (lc & 1);
        ++lc;
        if (last) {
          fetch(t + n_clusters);                     // the next rows stream in behind the last GEMM
          const int64_t r2 = (int64_t)(t + 2 * n_clusters) * PAIR + rank * TILE + row;
          if (32 * cg < p.in_dim && r2 < p.n) ptx::prefetch_l2(p.X + r2 * p.in_dim + 32 * cg);   // and the rows after those reach L2
        }
        if (!last || p.mode == MODE_DOT) {
          // hidden layer: n-half nh -> 32 columns of this row: bias + ReLU (+ fp16 pack | + dot)
          float dot0 = 0.f, dot1 = 0.f;
#pragma unroll 1
          for (int nh = 0; nh < 2; ++nh) {
            const int col = 128 * nh + 32 * cg;
            const float4* b = reinterpret_cast<const float4*>(s_bias + p.b_off[l] + col);
            const float bs = p.b_scale[l];             // fmaf(1, b, acc) == acc + b bit for bit; 0 skips a folded bias
            if (nh == 0) { ptx::mbar_wait(bar(B_ACC0), acc_use0 & 1); ++acc_use0; }
            else         { ptx::mbar_wait(bar(B_ACC1), acc_use1 & 1); ++acc_use1; }
            ptx::tc_fence_after();
            uint32_t v[32];
            ptx::tmem_ld32(t_acc + col, v);
            ptx::tmem_ld_wait();
            if (!last) {
              const int cc = 4 * nh + cg;
              const uint32_t base = a_row + (cc >> 1) * ATOM;
#pragma unroll
              for (int m = 0; m < 4; ++m) {
                const float4 b0 = b[2 * m], b1 = b[2 * m + 1];
                ptx::sts_v4(base + ((((cc & 1) * 4 + m) ^ sw) << 4),
                            ptx::cvt_h2_relu_sat(fmaf(bs, b0.x, __uint_as_float(v[8 * m])), fmaf(bs, b0.y, __uint_as_float(v[8 * m + 1]))),
                            ptx::cvt_h2_relu_sat(fmaf(bs, b0.z, __uint_as_float(v[8 * m + 2])), fmaf(bs, b0.w, __uint_as_float(v[8 * m + 3]))),
                            ptx::cvt_h2_relu_sat(fmaf(bs, b1.x, __uint_as_float(v[8 * m + 4])), fmaf(bs, b1.y, __uint_as_float(v[8 * m + 5]))),
                            ptx::cvt_h2_relu_sat(fmaf(bs, b1.z, __uint_as_float(v[8 * m + 6])), fmaf(bs, b1.w, __uint_as_float(v[8 * m + 7]))));
              }
              publish(nh == 0 ? B_IN0 : B_IN1);
            } else {
              const float4* w4 = reinterpret_cast<const float4*>(s_bias + p.wd_off + col);
#pragma unroll
              for (int m = 0; m < 8; ++m) {
                const float4 w = w4[m], bm = b[m];
                dot0 = fmaf(fmaxf(__uint_as_float(v[4 * m]) + bm.x, 0.f), w.x, dot0);
                dot1 = fmaf(fmaxf(__uint_as_float(v[4 * m + 1]) + bm.y, 0.f), w.y, dot1);
                dot0 = fmaf(fmaxf(__uint_as_float(v[4 * m + 2]) + bm.z, 0.f), w.z, dot0);
                dot1 = fmaf(fmaxf(__uint_as_float(v[4 * m + 3]) + bm.w, 0.f), w.w, dot1);
              }
            }
          }
          if (last) {
            // every MMA of this tile has completed: z of the next tile goes in first, then the scalar head
            ptx::tc_fence_before();
            if (t + n_clusters < p.T) {
              { pack_z(); store_w0(); }
              publish(B_IN0);
            }
            const float dot = dot0 + dot1;
            if (cg) s_dot[(cg - 1) * TILE + row] = dot;
            ptx::bar_sync(1, WORKERS);
            if (cg == 0 && r < p.n) p.out[r] = ((dot + s_dot[row]) + (s_dot[TILE + row] + s_dot[2 * TILE + row])) + s_bias[p.wd_off + HID];
          }
        } else {
          // output layer (N <= 64): 16 columns of this row -> fp32 x and / or snapped extents; two stacked heads (N = 128,
          // out2): 32 columns, the upper 64 go to out2
          const bool two = p.out2 != nullptr;
          ptx::mbar_wait(bar(B_ACC0), acc_use0 & 1);
          ++acc_use0;
          ptx::tc_fence_after();
          uint32_t v[16], v2[16];
          ptx::tmem_ld16(t_acc + (two ? 32 * cg : 16 * cg), v);
          if (two) ptx::tmem_ld16(t_acc + 32 * cg + 16, v2);
          if (t + n_clusters < p.T) {
            { pack_z(); store_w0(); }
            publish(B_IN0);
          }
          ptx::tmem_ld_wait();
          if (two) {
            if (r < p.n) {
              const int cb = 32 * cg;                                   // column in the stacked layer
              float* dst = (cb < p.out_dim ? p.out : p.out2) + r * p.out_dim + (cb & (p.out_dim - 1));
              const float* bb = s_bias + p.b_off[l] + cb;
#pragma unroll
              for (int j = 0; j < 2; ++j) {
                ptx::stg_f8(dst + 8 * j, __uint_as_float(v[8 * j]) + bb[8 * j], __uint_as_float(v[8 * j + 1]) + bb[8 * j + 1],
                            __uint_as_float(v[8 * j + 2]) + bb[8 * j + 2], __uint_as_float(v[8 * j + 3]) + bb[8 * j + 3],
                            __uint_as_float(v[8 * j + 4]) + bb[8 * j + 4], __uint_as_float(v[8 * j + 5]) + bb[8 * j + 5],
                            __uint_as_float(v[8 * j + 6]) + bb[8 * j + 6], __uint_as_float(v[8 * j + 7]) + bb[8 * j + 7]);
                ptx::stg_f8(dst + 16 + 8 * j, __uint_as_float(v2[8 * j]) + bb[16 + 8 * j], __uint_as_float(v2[8 * j + 1]) + bb[17 + 8 * j],
                            __uint_as_float(v2[8 * j + 2]) + bb[18 + 8 * j], __uint_as_float(v2[8 * j + 3]) + bb[19 + 8 * j],
                            __uint_as_float(v2[8 * j + 4]) + bb[20 + 8 * j], __uint_as_float(v2[8 * j + 5]) + bb[21 + 8 * j],
                            __uint_as_float(v2[8 * j + 6]) + bb[22 + 8 * j], __uint_as_float(v2[8 * j + 7]) + bb[23 + 8 * j]);
              }
            }
            ptx::tc_fence_before();
            continue;
          }
          const int c0 = 16 * cg;
          const float4* b = reinterpret_cast<const float4*>(s_bias + p.b_off[l] + c0);
          if (r < p.n && c0 < p.N[l] && vec8_ok) {
            // two 256-bit stores per output: full sectors for the lane-per-row pattern
            float y[16];
#pragma unroll
            for (int m = 0; m < 4; ++m) {
              const float4 bm = b[m];
              y[4 * m] = __uint_as_float(v[4 * m]) + bm.x; y[4 * m + 1] = __uint_as_float(v[4 * m + 1]) + bm.y;
              y[4 * m + 2] = __uint_as_float(v[4 * m + 2]) + bm.z; y[4 * m + 3] = __uint_as_float(v[4 * m + 3]) + bm.w;
            }
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const int c = c0 + 8 * j;
              if (c < p.out_dim) {
                if (p.out) ptx::stg_f8(p.out + r * p.out_dim + c, y[8 * j], y[8 * j + 1], y[8 * j + 2], y[8 * j + 3], y[8 * j + 4],
                                       y[8 * j + 5], y[8 * j + 6], y[8 * j + 7]);
                if (p.extents) {
                  int q[8];
                  const float y8[8] = {y[8 * j], y[8 * j + 1], y[8 * j + 2], y[8 * j + 3], y[8 * j + 4], y[8 * j + 5], y[8 * j + 6], y[8 * j + 7]};
                  snap_vec8(y8, p.snap_mean + c, p.snap_scale + c, p.snap_mode, q);
                  ptx::stg_i8(p.extents + r * p.out_dim + c, q[0], q[1], q[2], q[3], q[4], q[5], q[6], q[7]);
                }
              }
            }
          } else if (r < p.n && c0 < p.N[l]) {
#pragma unroll
            for (int m = 0; m < 4; ++m) {
              const int c = c0 + 4 * m;
              const float4 bm = b[m];
              const float y0 = __uint_as_float(v[4 * m]) + bm.x, y1 = __uint_as_float(v[4 * m + 1]) + bm.y;
              const float y2 = __uint_as_float(v[4 * m + 2]) + bm.z, y3 = __uint_as_float(v[4 * m + 3]) + bm.w;
              if (vec_ok) {
                if (c < p.out_dim) {
                  if (p.out) *reinterpret_cast<float4*>(p.out + r * p.out_dim + c) = make_float4(y0, y1, y2, y3);
                  if (p.extents) {
                    const float4 mu = __ldg(reinterpret_cast<const float4*>(p.snap_mean + c));
                    const float4 sc = __ldg(reinterpret_cast<const float4*>(p.snap_scale + c));
                    *reinterpret_cast<int4*>(p.extents + r * p.out_dim + c) =
                        make_int4(snap_one(y0, mu.x, sc.x, p.snap_mode), snap_one(y1, mu.y, sc.y, p.snap_mode),
                                  snap_one(y2, mu.z, sc.z, p.snap_mode), snap_one(y3, mu.w, sc.w, p.snap_mode));
                  }
                }
              } else {
                const float y[4] = {y0, y1, y2, y3};
#pragma unroll
                for (int e = 0; e < 4; ++e)
                  if (c + e < p.out_dim) {
                    if (p.out) p.out[r * p.out_dim + c + e] = y[e];
                    if (p.extents)
                      p.extents[r * p.out_dim + c + e] = snap_one(y[e], __ldg(p.snap_mean + c + e), __ldg(p.snap_scale + c + e), p.snap_mode);
                  }
              }
            }
          }
          ptx::tc_fence_before();
        }
      }
    }
  } else if (rank == 0) {
    // ===== MMA issuer (leader CTA) =====
    uint32_t in_use0 = 0, in_use1 = 0, lc = 0;
    for (int t = cluster; t < p.T; t += n_clusters) {
      const int n_issue = p.mc_T > 0 ? 1 + p.mc_T : p.n_mma;
      for (int li = 0; li < n_issue; ++li) {
        const int l = p.mc_T > 0 ? min(li, 1) : li;
        const int K = p.K[l], N = p.N[l];
        const int n_kh = K == HID ? 2 : 1, n_nh = N == HID ? 2 : 1;
        const int n_piece = N == HID ? 128 : N;                 // columns per MMA
        const int w_atom = (N / 2) * 128;                       // bytes per 64-column atom of the W image
        const uint32_t idesc = ptx::umma_idesc_f16(256, n_piece);
        const uint32_t acc = tmem + 256u * (lc & 1);
        ++lc;
        const uint64_t dA = ptx::umma_desc_sw128(sbase + p.a_off + (K == HID ? 0 : (4 - (K >> 6)) * ATOM));
        for (int kh = 0; kh < n_kh; ++kh) {
          if (kh == 0) { ptx::mbar_wait(bar(B_IN0), in_use0 & 1); ++in_use0; }
          else         { ptx::mbar_wait(bar(B_IN1), in_use1 & 1); ++in_use1; }
          ptx::tc_fence_after();
          if (ptx::elect_one()) {
            const int steps = K == HID ? 8 : K >> 4;
            for (int nh = 0; nh < n_nh; ++nh) {
              const uint64_t dW = ptx::umma_desc_sw128(sbase + p.w_off[l] + nh * 64 * 128);
              for (int i = 0; i < steps; ++i) {
                const int k = steps * kh + i;
                ptx::umma_f16_2cta(acc + 128u * nh, dA + (uint64_t)(((k >> 2) * ATOM + (k & 3) * 32) >> 4),
                                   dW + (uint64_t)(((k >> 2) * w_atom + (k & 3) * 32) >> 4), idesc, k > 0);
              }
              if (kh == n_kh - 1) ptx::umma_commit_2cta(bar(nh == 0 ? B_ACC0 : B_ACC1), 3);
            }
          }
          __syncwarp();
        }
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  if (warp == 1) ptx::tmem_dealloc_2cta(tmem, 512);
}

// ---------------------------------------------------------------------------------------------
static __half c_clamp(float x) { return __float2half_rn(std::min(std::max(x, -65504.f), 65504.f)); }


// Packs `layers` (row-major [out, in] fp32) into the per-rank swizzled images + fp16 biases and fills the
// shape part of ChainParams.  The last entry of a MODE_DOT chain is the scalar head (kept fp32, not an MMA).
static int chain2_pack(ChainPack& cp, const std::vector<ChainLayer>& layers, int mode, cudaStream_t st, int two_heads) {
  using namespace ch;
  cp.v2 = false;
  const int n_mma = mode == MODE_DOT ? (int)layers.size() - 1 : (int)layers.size();
  if (n_mma < 1 || n_mma > MAX_LAYERS) return VAEX_OK;
  ChainParams& P = *new (cp.params_storage) ChainParams();
  static_assert(sizeof(ChainParams) <= sizeof(cp.params_storage), "ChainPack::params_storage too small");
  P.mode = mode;
  P.n_mma = n_mma;
  const int out_max = two_heads ? 128 : 64;              // two stacked 64-wide heads: one N = 128 output layer
  // When the fp32 biases do not fit beside the weights and the operand (the encoder: 160 KB + 64 KB + 3.3 KB), layer 0's
  // bias rides in its GEMM instead: two spare input columns hold the constant 1 and W0 gets fp16 hi / lo parts of b0.
  // When even that does not fit (both latent heads: 176 KB of weights), W0's one-atom image moves into the operand.
  if (layers[0].in > 128) return VAEX_OK;
  const int K0 = layers[0].in <= 64 ? 64 : 128;          // one or two atoms of layer-0 operand
  const int dot_bytes = mode == MODE_DOT ? (n_mma == 2 ? 6 : 3) * TILE * 4 : 0;   // double-buffered over MC passes (two-layer predictor only)
  auto smem_need = [&](bool fold0, bool w0a, bool bias_g = false) {
    int w = 0, b = 0;
    for (int l = 0; l < n_mma; ++l) {
      const bool out_layer = mode == MODE_VECTOR && l == n_mma - 1;
      const int N = out_layer ? std::max(16, (layers[l].out + 15) / 16 * 16) : HID;
      if (!(w0a && l == 0)) w += (N / 2) * (l == 0 ? K0 : HID) * 2;
      if (!(fold0 && l == 0)) b += N;
    }
    if (mode == MODE_DOT) b += HID + 4;
    return w + 4 * ATOM + (bias_g ? 0 : b * 4) + 64 + dot_bytes;
  };
  const bool can_fold = n_mma > 1 && layers[0].in + 2 <= K0 && layers[0].out == HID;
  const bool fold0 = smem_need(false, false) > SMEM_MAX && can_fold;
  const bool w0a = fold0 && K0 == 64 && smem_need(true, false) > SMEM_MAX && smem_need(true, true) <= SMEM_MAX;
  // last resort (the segment cost predictor, 128 -> 256 -> 256 -> 256 -> 1: 160 KB of weights and no spare input column
  // to fold a bias into): the 4 KB of fp32 biases are read through L1 instead of from a shared copy
  const bool bias_g = K0 == 128 && smem_need(fold0, w0a) > SMEM_MAX;
  if (smem_need(fold0, w0a, bias_g) > SMEM_MAX) return VAEX_OK;
  P.bias_global = bias_g ? 1 : 0;
  P.ones_col = fold0 ? layers[0].in : -1;
  P.w0_in_a = w0a ? 1 : 0;
  int woff = 0, boff = 0;
  for (int l = 0; l < n_mma; ++l) {
    const ChainLayer& L = layers[l];
    const int K = l == 0 ? K0 : HID;
    const bool out_layer = mode == MODE_VECTOR && l == n_mma - 1;
    const int N = out_layer ? std::max(16, (L.out + 15) / 16 * 16) : HID;
    if ((l == 0 ? L.in > K0 : L.in != HID) || (out_layer ? L.out > out_max : L.out != HID)) return VAEX_OK;   // shape outside this kernel
    P.K[l] = K; P.N[l] = N; P.w_off[l] = woff; P.b_off[l] = boff;
    P.b_scale[l] = (fold0 && l == 0) ? 0.f : 1.f;
    if (!(w0a && l == 0)) woff += (N / 2) * K * 2;
    if (!(fold0 && l == 0)) boff += N;
  }
  P.in_dim = layers[0].in;
  P.out_dim = mode == MODE_VECTOR ? (two_heads ? layers[n_mma - 1].out / 2 : layers[n_mma - 1].out) : 1;
  if (two_heads && (mode != MODE_VECTOR || layers[n_mma - 1].out != 128)) return VAEX_OK;
  P.w_bytes = woff;
  P.a_off = woff;
  if (w0a) P.w_off[0] = P.a_off + 2 * ATOM;              // atom 2 of the operand
  P.wd_off = boff;
  if (mode == MODE_DOT) {
    if (layers[n_mma].in != HID || layers[n_mma].out != 1) return VAEX_OK;
    boff += HID + 4;
  }
  P.bias_off = P.a_off + 4 * ATOM;
  P.bias_count = boff;
  P.bar_off = P.bias_off + (bias_g ? 0 : boff * 4);
  cp.smem_bytes = P.bar_off + 64 + dot_bytes;
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
  std::vector<__half> img0(w0a ? (size_t)ATOM : 0, __float2half_rn(0.f));   // 2 ranks x one atom (16 KB) when W0 lives in the operand
  std::vector<float> bias((size_t)boff, 0.f);
  for (int l = 0; l < n_mma; ++l) {
    const ChainLayer& L = layers[l];
    const int nh = P.N[l] / 2;
    for (int r = 0; r < 2; ++r)
      for (int u = 0; u < nh; ++u) {
        // a 256-wide layer is issued as two 128-column MMAs: image rows [64 h, +64) of rank r = units 128 h + 64 r + [0, 64)
        const int unit = P.N[l] == HID ? 128 * (u >> 6) + 64 * r + (u & 63) : r * nh + u;
        if (unit >= L.out) continue;
        __half* dst = (w0a && l == 0) ? img0.data() + (size_t)r * (ATOM / 2) : img.data() + (size_t)r * (woff / 2) + P.w_off[l] / 2;
        for (int k = 0; k < L.in; ++k) dst[sw_off(nh, u, k) / 2] = c_clamp((*L.w)[(size_t)unit * L.in + k]);
        if (fold0 && l == 0) {
          const __half hi = c_clamp((*L.b)[unit]);
          dst[sw_off(nh, u, L.in) / 2] = hi;
          dst[sw_off(nh, u, L.in + 1) / 2] = c_clamp((*L.b)[unit] - __half2float(hi));
        }
      }
    if (!(fold0 && l == 0))
      for (int j = 0; j < L.out; ++j) bias[P.b_off[l] + j] = (*L.b)[j];
  }
  if (mode == MODE_DOT) {
    for (int k = 0; k < HID; ++k) bias[P.wd_off + k] = (*layers[n_mma].w)[k];
    bias[P.wd_off + HID] = (*layers[n_mma].b)[0];
  }
  if (cp.w) cudaFree(cp.w);
  if (cp.bias) cudaFree(cp.bias);
  cp.w = nullptr; cp.bias = nullptr;
  VAEX_CUDA(cudaMalloc(&cp.w, img.size() * sizeof(__half)));
  VAEX_CUDA(cudaMalloc(&cp.bias, std::max<size_t>(bias.size(), 1) * sizeof(float)));
  VAEX_CUDA(cudaMemcpyAsync(cp.w, img.data(), img.size() * sizeof(__half), cudaMemcpyHostToDevice, st));
  VAEX_CUDA(cudaMemcpyAsync(cp.bias, bias.data(), bias.size() * sizeof(float), cudaMemcpyHostToDevice, st));
  if (w0a) {
    if (!cp.w0) VAEX_CUDA(cudaMalloc(&cp.w0, 2 * ATOM));
    VAEX_CUDA(cudaMemcpyAsync(cp.w0, img0.data(), 2 * ATOM, cudaMemcpyHostToDevice, st));
  }
  VAEX_CUDA(cudaStreamSynchronize(st));
  VAEX_CUDA(cudaFuncSetAttribute(chain_tc_kernel<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_MAX));
  VAEX_CUDA(cudaFuncSetAttribute(chain_tc_kernel<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_MAX));
  VAEX_CUDA(cudaFuncSetAttribute(chain_tc_kernel<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_MAX));
  cp.v2 = true;
  return VAEX_OK;
}

// Packs both kernels' images: chain_tc3.cu (two tiles in flight, streamed weights) runs every plain chain it has a shape
// for, chain_mc3.cu the MC-dropout passes; this file is the fallback for both (VAEX_CHAIN_V2=1 or a shape they do not take).
int chain_tc_pack(ChainPack& cp, const std::vector<ChainLayer>& layers, int mode, cudaStream_t st, int two_heads) {
  cp.ready = false;
  TRY(chain2_pack(cp, layers, mode, st, two_heads));
  TRY(chain3_pack(cp, layers, mode, st, two_heads));
  if (mode == ch::MODE_DOT) TRY(chain_mc3_pack(cp, layers, st));
  cp.ready = cp.v2 || cp.v3 != nullptr;
  return VAEX_OK;
}

void chain_tc_release(ChainPack& cp) {
  if (cp.w) cudaFree(cp.w);
  if (cp.w0) cudaFree(cp.w0);
  if (cp.bias) cudaFree(cp.bias);
  cp.w = nullptr; cp.w0 = nullptr; cp.bias = nullptr;
  chain3_release(cp);
  chain_mc3_release(cp);
  cp.ready = cp.v2 = false;
}

// VAEX_CHAIN_V2=1 keeps every chain on this file's kernel (A/B runs)
static bool chain_force_v2() {
  static const bool f = [] { const char* e = getenv("VAEX_CHAIN_V2"); return e && e[0] == '1'; }();
  return f;
}

int chain_tc_run(const ChainPack& cp, const float* X, int64_t n, float* out, int32_t* extents, const float* mean,
                 const float* scale, int snap_mode, const float* wd, const float* bd, cudaStream_t st, const ChainDropout* drop,
                 float* out2) {
  using namespace ch;
  VAEX_CHECK(cp.ready, "tcgen05 chain not packed");
  if (n == 0) return VAEX_OK;
  if (cp.v3 && !drop && !(cp.v2 && chain_force_v2())) return chain3_run(cp, X, n, out, extents, mean, scale, snap_mode, st, out2);
  if (drop && cp.mc3 && !(cp.v2 && chain_force_v2())) return chain_mc3_run(cp, X, n, out, *drop, st);
  VAEX_CHECK(cp.v2, "this chain has no resident-weight image (MC-dropout passes need the two-layer 64-wide predictor)");
  ChainParams P = *reinterpret_cast<const ChainParams*>(cp.params_storage);
  P.X = X; P.w = cp.w; P.bias = cp.bias; P.out = out; P.extents = extents; P.snap_mean = mean; P.snap_scale = scale;
  P.snap_mode = snap_mode; P.wd = wd; P.bd = bd; P.n = n; P.T = (int)cdiv(n, PAIR);
  P.out2 = out2; P.w0_img = cp.w0;
  VAEX_CHECK(!out2 || (P.mode == MODE_VECTOR && P.N[P.n_mma - 1] == 128 && P.out_dim == 64 && out), "two-head output needs the stacked 128-wide layer");
  VAEX_CHECK(out2 || P.mode != MODE_VECTOR || P.N[P.n_mma - 1] <= 64, "the stacked two-head chain needs both outputs");
  P.mc_T = 0; P.drop_mask = nullptr; P.var_out = nullptr; P.drop_seed = 0; P.drop_chunk = 1; P.drop_p = 0.f; P.drop_inv = 1.f;
  if (drop) {
    VAEX_CHECK(P.mode == MODE_DOT && P.n_mma == 2 && drop->passes >= 1, "dropout passes need the two-layer predictor chain");
    P.mc_T = drop->passes; P.drop_mask = drop->mask; P.var_out = drop->var_out; P.drop_seed = drop->seed;
    P.drop_chunk = std::max<int64_t>(drop->chunk_rows, 1); P.drop_p = drop->p; P.drop_inv = 1.f / (1.f - drop->p);
  }
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int clusters = std::max(1, std::min(sms / 2, P.T));
  if (P.K[0] == 64) chain_tc_kernel<false, 1><<<2 * clusters, THREADS, cp.smem_bytes, st>>>(P);
  else if (P.bias_global) chain_tc_kernel<true, 2><<<2 * clusters, THREADS, cp.smem_bytes, st>>>(P);
  else chain_tc_kernel<false, 2><<<2 * clusters, THREADS, cp.smem_bytes, st>>>(P);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

}  // namespace vaex
