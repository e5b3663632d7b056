// Multi-step latent ascent through the NORTH-STAR graph  score(z) = head(decoder_trunk(z))  (BASELINE.json north_star (3):
// "a fused backward pass through the decoder and cost model"; SURVEY 8d config 3: 1.115 MFLOP per restart and step):
//
//   a1 = relu(D0 z  + c0)      decoder.0   [64 -> 256]   (VAE_feature_head.decoder, utils/model.py:31-39)
//   a2 = relu(D1 a1 + c1)      decoder.2   [256 -> 256]
//   a3 = relu(D2 a2 + c2)      decoder.4   [256 -> 256]
//   y1 = relu(L0 a3 + e0) + a3 l0          (SegmentSumMLPModule's residual head on the 256-d vector, mlp_model.py:210-221)
//   y2 = relu(L1 y1 + e1) + y1 l1
//   score = wd . y2 + bd       decoder
//   z += lr * d score / d z,   `steps` times; the single-step precedent is utils/select.py:128-135.
//
// Same scheme as ascent_cluster.cu, one size up: ALL steps of a restart stay on chip.  A cluster of EIGHT CTAs owns
// sixteen restarts; each CTA keeps the fp32 weights of its 32 units of every layer in shared memory (140 KB, k-major with
// an odd row stride: the forward W x and the backward W^T g walks are both bank-conflict free).  Forward: a layer's 32 x 16
// outputs are pushed into every CTA's activation buffer through distributed shared memory, one cluster barrier per layer.
// Backward: every CTA forms the partial sums of W^T g over ITS units for all inputs, one cluster barrier, then pulls the
// eight partials of its own 32 inputs in rank order (fixed summation order: reruns are bit-identical and the eight copies
// of z never diverge).  Nine cluster barriers per step, no global memory traffic between the prologue and the final store.
#include <cooperative_groups.h>

#include "common.cuh"
#include "model.cuh"

namespace cg = cooperative_groups;

namespace vaex {
namespace ascx {

constexpr int L = 64, H = 256, CL = 8, Q = H / CL, ROWS = 16, T = 256, WS = Q + 1, NL = 5;
// shared memory layout (floats); activations are k-major with the 16 rows innermost
constexpr int O_W0 = 0;                           // [L][WS]   Wt0[k][u] = D0[rank*Q + u][k]
constexpr int O_W1 = O_W0 + L * WS;               // [4][H][WS]
constexpr int O_B = O_W1 + 4 * H * WS;            // [NL][Q] biases of my units
constexpr int O_WD = O_B + NL * Q;                // [Q]
constexpr int O_Z = ((O_WD + Q + 3) / 4) * 4;     // [L][ROWS]  full z (identical in the eight CTAs)
constexpr int O_ACTA = O_Z + L * ROWS;            // [H][ROWS]  full activation (written by all CTAs) / backward partials
constexpr int O_ACTB = O_ACTA + H * ROWS;
constexpr int O_R = O_ACTB + H * ROWS;            // [NL][Q][ROWS] relu outputs of my units (masks of the backward pass)
constexpr int O_G = O_R + NL * Q * ROWS;          // [Q][ROWS]  gradient at my units' pre-activations
constexpr int O_GRES = O_G + Q * ROWS;            // [Q][ROWS]  gradient at my units' outputs (the residual branch)
constexpr int O_GZP = O_GRES + Q * ROWS;          // [L][ROWS]  my partial of d z, read by the others
constexpr int O_RED = O_GZP + L * ROWS;           // [4][Q][ROWS] split-K scratch
constexpr int O_SC = O_RED + 4 * Q * ROWS;        // [ROWS]
constexpr int SMEM_FLOATS = O_SC + ROWS;
constexpr size_t SMEM_BYTES = SMEM_FLOATS * sizeof(float);
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");

// partial sums over a quarter of K of out[u][r] for my Q units: thread = (up: units up, up + 16; rg: rows 4rg..4rg+3; kq)
template <int K>
__device__ __forceinline__ void dense_quarter(const float* __restrict__ Wt, const float* __restrict__ actT, float* __restrict__ red) {
  const int t = threadIdx.x, up = t & 15, rg = (t >> 4) & 3, kq = t >> 6;
  float a0[4] = {0.f, 0.f, 0.f, 0.f}, a1[4] = {0.f, 0.f, 0.f, 0.f};
  const int k0 = kq * (K / 4);
#pragma unroll 4
  for (int k = k0; k < k0 + K / 4; ++k) {
    const float w0 = Wt[k * WS + up], w1 = Wt[k * WS + up + 16];
    const float4 x = *reinterpret_cast<const float4*>(actT + k * ROWS + 4 * rg);
    a0[0] = fmaf(x.x, w0, a0[0]); a0[1] = fmaf(x.y, w0, a0[1]); a0[2] = fmaf(x.z, w0, a0[2]); a0[3] = fmaf(x.w, w0, a0[3]);
    a1[0] = fmaf(x.x, w1, a1[0]); a1[1] = fmaf(x.y, w1, a1[1]); a1[2] = fmaf(x.z, w1, a1[2]); a1[3] = fmaf(x.w, w1, a1[3]);
  }
  float* r = red + kq * (Q * ROWS);
  *reinterpret_cast<float4*>(r + up * ROWS + 4 * rg) = make_float4(a0[0], a0[1], a0[2], a0[3]);
  *reinterpret_cast<float4*>(r + (up + 16) * ROWS + 4 * rg) = make_float4(a1[0], a1[1], a1[2], a1[3]);
}
__device__ __forceinline__ float red4(const float* red, int o) {
  return ((red[o] + red[Q * ROWS + o]) + red[2 * Q * ROWS + o]) + red[3 * Q * ROWS + o];
}
// part[k][r] = sum_{u mine} g[u][r] Wt[k][u] for all H inputs k: thread = k, sixteen accumulators
__device__ __forceinline__ void backward_partial(const float* __restrict__ Wt, const float* __restrict__ g, float* __restrict__ part) {
  const int k = threadIdx.x;
  float a[ROWS];
#pragma unroll
  for (int r = 0; r < ROWS; ++r) a[r] = 0.f;
#pragma unroll 2
  for (int u = 0; u < Q; ++u) {
    const float w = Wt[k * WS + u];
#pragma unroll
    for (int q4 = 0; q4 < ROWS / 4; ++q4) {
      const float4 gv = *reinterpret_cast<const float4*>(g + u * ROWS + 4 * q4);
      a[4 * q4] = fmaf(gv.x, w, a[4 * q4]); a[4 * q4 + 1] = fmaf(gv.y, w, a[4 * q4 + 1]);
      a[4 * q4 + 2] = fmaf(gv.z, w, a[4 * q4 + 2]); a[4 * q4 + 3] = fmaf(gv.w, w, a[4 * q4 + 3]);
    }
  }
#pragma unroll
  for (int q4 = 0; q4 < ROWS / 4; ++q4)
    *reinterpret_cast<float4*>(part + k * ROWS + 4 * q4) = make_float4(a[4 * q4], a[4 * q4 + 1], a[4 * q4 + 2], a[4 * q4 + 3]);
}

struct Params {
  const float* W[NL];      // D0 [H, L], D1, D2, L0, L1 [H, H]
  const float* b[NL];
  const float *wd, *bd;
  float* z;                // [n, L] in / out
  float* score;            // [n]
  int n, steps;
  float lr;
};

__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(T, 1) ascent_compose_kernel(const Params p) {
  extern __shared__ __align__(16) float sm[];
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int t = threadIdx.x;
  const int row0 = (blockIdx.x / CL) * ROWS;
  float* Wt[NL] = {sm + O_W0, sm + O_W1, sm + O_W1 + H * WS, sm + O_W1 + 2 * H * WS, sm + O_W1 + 3 * H * WS};
  float *bias = sm + O_B, *wd = sm + O_WD, *zT = sm + O_Z, *red = sm + O_RED, *g = sm + O_G, *gres = sm + O_GRES, *gzp = sm + O_GZP;
  float* R = sm + O_R;
  float* rem[CL];
#pragma unroll
  for (int c = 0; c < CL; ++c) rem[c] = cluster.map_shared_rank(sm, c);

  // ---- prologue: my 32 units of every layer, transposed into shared memory ----
  for (int i = t; i < Q * L; i += T) { const int u = i / L, k = i - u * L; Wt[0][k * WS + u] = p.W[0][(size_t)(rank * Q + u) * L + k]; }
  for (int l = 1; l < NL; ++l)
    for (int i = t; i < Q * H; i += T) { const int u = i / H, k = i - u * H; Wt[l][k * WS + u] = p.W[l][(size_t)(rank * Q + u) * H + k]; }
  if (t < Q) {
    for (int l = 0; l < NL; ++l) bias[l * Q + t] = p.b[l][rank * Q + t];
    wd[t] = p.wd[rank * Q + t];
  }
  for (int i = t; i < L * ROWS; i += T) { const int k = i / ROWS, r = i % ROWS; zT[i] = row0 + r < p.n ? p.z[(size_t)(row0 + r) * L + k] : 0.f; }
  cluster.sync();

  for (int s = 0; s <= p.steps; ++s) {           // the last pass is the forward half only: the score of the final z
    // ================= forward =================
    // layers 0..2: r_l = relu(W_l in + b_l), pushed into every CTA's next activation buffer
    // layer 3: r_3 = relu(L0 a3 + e0), y1 = r_3 + a3 pushed;  layer 4: r_4 = relu(L1 y1 + e1), y2 = r_4 + y1 (local)
    float4 sc4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
    for (int l = 0; l < NL; ++l) {
      const float* in = l == 0 ? zT : ((l & 1) ? sm + O_ACTA : sm + O_ACTB);       // l = 1, 3 read A; l = 2, 4 read B
      const int out_off = (l & 1) ? O_ACTB : O_ACTA;                                // l = 0, 2 write A; l = 1, 3 write B
      if (l == 0) dense_quarter<L>(Wt[0], in, red); else dense_quarter<H>(Wt[l], in, red);
      __syncthreads();
      {
        // epilogue: four consecutive rows of one unit per thread item (128 items); the two thread halves compute the same
        // values and push them to four CTAs each as 16-byte distributed-shared-memory stores (4-byte pushes made the
        // remote-store count, 4096 per layer and CTA, the cost of a step)
        const int item = t & 127, half = t >> 7, o = 4 * item, u = o / ROWS;
        const float bu = bias[l * Q + u];
        float4 v = make_float4(fmaxf(red4(red, o) + bu, 0.f), fmaxf(red4(red, o + 1) + bu, 0.f),
                               fmaxf(red4(red, o + 2) + bu, 0.f), fmaxf(red4(red, o + 3) + bu, 0.f));
        if (half == 0) *reinterpret_cast<float4*>(R + l * Q * ROWS + o) = v;
        if (l >= 3) {                                                               // residual: my slice of the layer's input
          const float4 x = *reinterpret_cast<const float4*>(in + rank * (Q * ROWS) + o);
          v.x += x.x; v.y += x.y; v.z += x.z; v.w += x.w;
        }
        if (l < NL - 1) {
#pragma unroll
          for (int c = 0; c < CL / 2; ++c)
            *reinterpret_cast<float4*>(rem[4 * half + c] + out_off + rank * (Q * ROWS) + o) = v;
        } else if (half == 0) {
          const float w = wd[u];
          sc4 = make_float4(w * v.x, w * v.y, w * v.z, w * v.w);
          *reinterpret_cast<float4*>(gres + o) = make_float4(w, w, w, w);           // d score / d y2 at my units
        }
      }
      if (l < NL - 1) cluster.sync(); else __syncthreads();
    }
    if (s == p.steps) {
      // score partial of my units per row: reduce over u within the CTA (o = u * ROWS + r)
      if (t < 128) *reinterpret_cast<float4*>(red + 4 * t) = sc4;
      __syncthreads();
      if (t < ROWS) {
        float c = 0.f;
        for (int u = 0; u < Q; ++u) c += red[u * ROWS + t];
        sm[O_SC + t] = c;
      }
      break;
    }
    // ================= backward =================
    // entering layer l: gres = d score / d (output of layer l at my units).  g = gres * relu'(r_l); partial W_l^T g for all
    // inputs -> my partial buffer; barrier; pull my 32 inputs' sums (+ gres for the residual layers) -> gres of layer l - 1.
#pragma unroll 1
    for (int l = NL - 1; l >= 1; --l) {
      for (int o = t; o < Q * ROWS; o += T) g[o] = R[l * Q * ROWS + o] > 0.f ? gres[o] : 0.f;
      __syncthreads();
      const int part_off = (l & 1) ? O_ACTB : O_ACTA;      // alternate: the others may still be pulling the previous layer's
      backward_partial(Wt[l], g, sm + part_off);
      cluster.sync();
      if (t < 128) {                                         // 16-byte pulls, rank order
        const int o = 4 * t, oo = part_off + rank * (Q * ROWS) + o;
        float4 a = *reinterpret_cast<const float4*>(rem[0] + oo);
#pragma unroll
        for (int c = 1; c < CL; ++c) {
          const float4 b = *reinterpret_cast<const float4*>(rem[c] + oo);
          a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        }
        if (l >= 3) {                                        // residual: the identity branch carries the output gradient
          const float4 b = *reinterpret_cast<const float4*>(gres + o);
          a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        }
        *reinterpret_cast<float4*>(gres + o) = a;
      }
      __syncthreads();
    }
    // layer 0: d z partial over my units: thread = (j, rg)
    for (int o = t; o < Q * ROWS; o += T) g[o] = R[o] > 0.f ? gres[o] : 0.f;
    __syncthreads();
    {
      const int j = t & 63, rg = t >> 6;
      float a[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
      for (int u = 0; u < Q; ++u) {
        const float w = Wt[0][j * WS + u];
        const float4 gv = *reinterpret_cast<const float4*>(g + u * ROWS + 4 * rg);
        a[0] = fmaf(gv.x, w, a[0]); a[1] = fmaf(gv.y, w, a[1]); a[2] = fmaf(gv.z, w, a[2]); a[3] = fmaf(gv.w, w, a[3]);
      }
      *reinterpret_cast<float4*>(gzp + j * ROWS + 4 * rg) = make_float4(a[0], a[1], a[2], a[3]);
    }
    cluster.sync();
    {
      const int o = 4 * t;                                   // L * ROWS = 1024 = 256 threads x 4
      float4 a = *reinterpret_cast<const float4*>(rem[0] + O_GZP + o);
#pragma unroll
      for (int c = 1; c < CL; ++c) {
        const float4 b = *reinterpret_cast<const float4*>(rem[c] + O_GZP + o);
        a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
      }
      float4 zc = *reinterpret_cast<const float4*>(zT + o);
      zc.x = fmaf(p.lr, a.x, zc.x); zc.y = fmaf(p.lr, a.y, zc.y); zc.z = fmaf(p.lr, a.z, zc.z); zc.w = fmaf(p.lr, a.w, zc.w);
      *reinterpret_cast<float4*>(zT + o) = zc;
    }
    __syncthreads();
  }

  cluster.sync();
  if (rank == 0) {
    if (t < ROWS && row0 + t < p.n) {
      float c = rem[0][O_SC + t];
#pragma unroll
      for (int k = 1; k < CL; ++k) c += rem[k][O_SC + t];
      p.score[row0 + t] = c + p.bd[0];
    }
    for (int i = t; i < L * ROWS; i += T) { const int k = i / ROWS, r = i % ROWS; if (row0 + r < p.n) p.z[(size_t)(row0 + r) * L + k] = zT[i]; }
  }
  cluster.sync();                                // nobody leaves while its shared memory may still be read
}

}  // namespace ascx

// z [n, 64] in place, score [n]
int ascent_compose_run(vaex_model* vae, vaex_model* mlp, float* z, int64_t n, int steps, float lr, float* score, cudaStream_t st) {
  using namespace ascx;
  VAEX_CHECK(vae->L == L && vae->H == H && mlp->H == H && !mlp->use_norm && !mlp->add_sigmoid,
             "the composed ascent needs L = 64, hidden 256 on both models, no BatchNorm / sigmoid on the head");
  Params p;
  const char* wn[NL] = {"decoder.0", "decoder.2", "decoder.4", "l0.0", "l1.0"};
  for (int l = 0; l < NL; ++l) {
    vaex_model* m = l < 3 ? vae : mlp;
    p.W[l] = m->w(std::string(wn[l]) + ".weight");
    p.b[l] = m->w(std::string(wn[l]) + ".bias");
    VAEX_CHECK(p.W[l] && p.b[l], "weights missing for the composed ascent");
  }
  p.wd = mlp->w("decoder.weight"); p.bd = mlp->w("decoder.bias");
  VAEX_CHECK(p.wd && p.bd, "head weights missing");
  p.z = z; p.score = score; p.n = (int)n; p.steps = steps; p.lr = lr;
  VAEX_CUDA(cudaFuncSetAttribute(ascent_compose_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
  const unsigned clusters = (unsigned)cdiv(n, ROWS);
  ascent_compose_kernel<<<clusters * CL, T, SMEM_BYTES, st>>>(p);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

}  // namespace vaex
