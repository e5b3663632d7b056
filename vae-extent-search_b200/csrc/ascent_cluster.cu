// Multi-step latent ascent z += lr * d cost / d z through the extent-family cost predictor (BASELINE config 3; the
// single-step precedent is select_topk_z_grad, scripts/pre_experiments/model_myself/utils/select.py:111-148) as ONE
// kernel: all `steps` iterations of a restart stay on chip.
//
// Restarts are independent rows, so nothing needs a grid-wide exchange: a cluster of four CTAs owns eight restarts and
// runs every step for them.  Each CTA keeps the weights of its quarter of the hidden units in shared memory in fp32
// (W0 quarter 16 KB + W1 quarter 65 KB -- the 256 KB of W1 do not fit one CTA) and the four exchange activations through
// distributed shared memory: per step three cluster barriers (h1 gathered; the partial sums of d h1 over each CTA's
// rows of W1; the partial sums of d z), no global memory traffic at all.  The launch-per-layer version replayed from a
// CUDA graph took 63 us per step for 256 restarts (8 dependent launches); the arithmetic is 84 MFLOP.
//
// Per CTA and step (256 threads, fp32 FMA, fixed summation order -> deterministic, the four CTAs hold identical z):
//   A  h1[r, u]  = relu(b0[u] + sum_k z[r, k] W0[u, k])            my 64 u, 8 rows, split-K over 4 thread groups, pushed to all
//   B  h2[r, u]  = relu(b1[u] + sum_k h1[r, k] W1[u, k]);  g2 = wd[u] * (h2 > 0)
//   C  g1p[r, k] = sum_{u mine} g2[r, u] W1[u, k]                  partial over my rows of W1, all 256 k
//   D  g1[r, u]  = (h1[r, u] > 0) * sum_{c} g1p_c[r, u]            my 64 u, pulled from the four CTAs in rank order
//   E  gzp[r, j] = sum_{u mine} g1[r, u] W0[u, j]                  partial over my u
//   F  z[r, j]  += lr * sum_c gzp_c[r, j]
#include <cooperative_groups.h>

#include "common.cuh"
#include "model.cuh"

namespace cg = cooperative_groups;

namespace vaex {
namespace asc {

constexpr int L = 64, PH = 256, CL = 4, Q = PH / CL, ROWS = 8, T = 256, WS = Q + 1;   // WS: odd row stride (bank spread)
// shared memory layout (floats); activations are stored k-major with the 8 rows innermost ("T" = transposed)
constexpr int O_W0T = 0;                      // [L][WS]   W0t[k][u] = W0[rank*Q + u][k]
constexpr int O_W1T = O_W0T + L * WS;         // [PH][WS]  W1t[k][u] = W1[rank*Q + u][k]
constexpr int O_B0 = O_W1T + PH * WS;         // [Q]
constexpr int O_B1 = O_B0 + Q;
constexpr int O_WD = O_B1 + Q;
constexpr int O_ZT = ((O_WD + Q + 3) / 4) * 4;   // [L][8]      (16-byte aligned from here on)
constexpr int O_H1T = O_ZT + L * ROWS;        // [PH][8]   full h1, written by all four CTAs
constexpr int O_H2T = O_H1T + PH * ROWS;      // [Q][8]
constexpr int O_G2T = O_H2T + Q * ROWS;       // [Q][8]
constexpr int O_G1P = O_G2T + Q * ROWS;       // [PH][8]   my partial of d h1, read by the others
constexpr int O_G1T = O_G1P + PH * ROWS;      // [Q][8]
constexpr int O_GZP = O_G1T + Q * ROWS;       // [L][8]    my partial of d z, read by the others
constexpr int O_RED = O_GZP + L * ROWS;       // [4096]    split-K scratch
constexpr int O_COST = O_RED + 4096;          // [8]
constexpr int SMEM_FLOATS = O_COST + ROWS;
constexpr size_t SMEM_BYTES = SMEM_FLOATS * sizeof(float);

// out[u][r] partial sums for my Q units: thread = (up: units up and up + 32, rg: rows 4rg..4rg+3, kq: quarter of K)
template <int K>
__device__ __forceinline__ void dense_quarter(const float* __restrict__ Wt, const float* __restrict__ actT, float* __restrict__ red) {
  const int t = threadIdx.x, up = t & 31, rg = (t >> 5) & 1, kq = t >> 6;
  float a0[4] = {0.f, 0.f, 0.f, 0.f}, a1[4] = {0.f, 0.f, 0.f, 0.f};
  const int k0 = kq * (K / 4);
#pragma unroll 4
  for (int k = k0; k < k0 + K / 4; ++k) {
    const float w0 = Wt[k * WS + up], w1 = Wt[k * WS + up + 32];
    const float4 x = *reinterpret_cast<const float4*>(actT + k * ROWS + 4 * rg);
    a0[0] = fmaf(x.x, w0, a0[0]); a0[1] = fmaf(x.y, w0, a0[1]); a0[2] = fmaf(x.z, w0, a0[2]); a0[3] = fmaf(x.w, w0, a0[3]);
    a1[0] = fmaf(x.x, w1, a1[0]); a1[1] = fmaf(x.y, w1, a1[1]); a1[2] = fmaf(x.z, w1, a1[2]); a1[3] = fmaf(x.w, w1, a1[3]);
  }
  float* r = red + kq * (Q * ROWS);
  *reinterpret_cast<float4*>(r + up * ROWS + 4 * rg) = make_float4(a0[0], a0[1], a0[2], a0[3]);
  *reinterpret_cast<float4*>(r + (up + 32) * ROWS + 4 * rg) = make_float4(a1[0], a1[1], a1[2], a1[3]);
}
__device__ __forceinline__ float red4(const float* red, int o) {
  return ((red[o] + red[Q * ROWS + o]) + red[2 * Q * ROWS + o]) + red[3 * Q * ROWS + o];
}

struct Params {
  const float *W0, *b0, *W1, *b1, *wd, *bd;   // cost_predictor.{0,3,6}.{weight,bias}
  float* z;                                   // [n, L] in / out
  float* cost;                                // [n]
  int n, steps;
  float lr;
};

__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(T, 1) ascent_cluster_kernel(const Params p) {
  extern __shared__ __align__(16) float sm[];
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int t = threadIdx.x;
  const int row0 = (blockIdx.x / CL) * ROWS;   // first restart of this cluster
  float *W0t = sm + O_W0T, *W1t = sm + O_W1T, *b0 = sm + O_B0, *b1 = sm + O_B1, *wd = sm + O_WD, *zT = sm + O_ZT;
  float *h1T = sm + O_H1T, *h2T = sm + O_H2T, *g2T = sm + O_G2T, *g1p = sm + O_G1P, *g1T = sm + O_G1T, *gzp = sm + O_GZP;
  float *red = sm + O_RED, *costp = sm + O_COST;
  float* rem[CL];                              // the other CTAs' shared memory (rank order, mine included)
#pragma unroll
  for (int c = 0; c < CL; ++c) rem[c] = cluster.map_shared_rank(sm, c);

  // ---- prologue: my quarter of the weights (coalesced global reads, transposed into shared memory), biases, z ----
  for (int i = t; i < Q * L; i += T) { const int u = i / L, k = i - u * L; W0t[k * WS + u] = p.W0[(size_t)(rank * Q + u) * L + k]; }
  for (int i = t; i < Q * PH; i += T) { const int u = i / PH, k = i - u * PH; W1t[k * WS + u] = p.W1[(size_t)(rank * Q + u) * PH + k]; }
  if (t < Q) { b0[t] = p.b0[rank * Q + t]; b1[t] = p.b1[rank * Q + t]; wd[t] = p.wd[rank * Q + t]; }
  for (int i = t; i < L * ROWS; i += T) { const int k = i >> 3, r = i & 7; zT[i] = row0 + r < p.n ? p.z[(size_t)(row0 + r) * L + k] : 0.f; }
  cluster.sync();                              // everybody's shared memory is live before the first remote store

  for (int s = 0; s <= p.steps; ++s) {         // the last pass is the forward half only: the cost of the final z
    // ---- A: h1 quarter, pushed into every CTA's h1T ----
    dense_quarter<L>(W0t, zT, red);
    __syncthreads();
#pragma unroll
    for (int o = t; o < Q * ROWS; o += T) {
      const float v = fmaxf(red4(red, o) + b0[o >> 3], 0.f);
#pragma unroll
      for (int c = 0; c < CL; ++c) rem[c][O_H1T + rank * (Q * ROWS) + o] = v;
    }
    cluster.sync();                            // (1) h1 complete everywhere
    // ---- B: h2 quarter, g2 = wd * relu'(h2) ----
    dense_quarter<PH>(W1t, h1T, red);
    __syncthreads();
#pragma unroll
    for (int o = t; o < Q * ROWS; o += T) {
      const float v = fmaxf(red4(red, o) + b1[o >> 3], 0.f);
      h2T[o] = v;
      g2T[o] = v > 0.f ? wd[o >> 3] : 0.f;
    }
    __syncthreads();
    if (s == p.steps) break;
    // ---- C: g1p[k][r] = sum_{u mine} g2[u][r] W1t[k][u]; thread = (kp: k = kp and kp + 128, us: half of my u) ----
    {
      const int kp = t & 127, us = t >> 7;
      float a0[ROWS], a1[ROWS];
#pragma unroll
      for (int r = 0; r < ROWS; ++r) { a0[r] = 0.f; a1[r] = 0.f; }
#pragma unroll 4
      for (int u = us * 32; u < us * 32 + 32; ++u) {
        const float w0 = W1t[kp * WS + u], w1 = W1t[(kp + 128) * WS + u];
        const float4 ga = *reinterpret_cast<const float4*>(g2T + u * ROWS), gb = *reinterpret_cast<const float4*>(g2T + u * ROWS + 4);
        const float g[ROWS] = {ga.x, ga.y, ga.z, ga.w, gb.x, gb.y, gb.z, gb.w};
#pragma unroll
        for (int r = 0; r < ROWS; ++r) { a0[r] = fmaf(g[r], w0, a0[r]); a1[r] = fmaf(g[r], w1, a1[r]); }
      }
      float* rr = red + us * (PH * ROWS);
      *reinterpret_cast<float4*>(rr + kp * ROWS) = make_float4(a0[0], a0[1], a0[2], a0[3]);
      *reinterpret_cast<float4*>(rr + kp * ROWS + 4) = make_float4(a0[4], a0[5], a0[6], a0[7]);
      *reinterpret_cast<float4*>(rr + (kp + 128) * ROWS) = make_float4(a1[0], a1[1], a1[2], a1[3]);
      *reinterpret_cast<float4*>(rr + (kp + 128) * ROWS + 4) = make_float4(a1[4], a1[5], a1[6], a1[7]);
    }
    __syncthreads();
    for (int o = t; o < PH * ROWS; o += T) g1p[o] = red[o] + red[PH * ROWS + o];
    cluster.sync();                            // (2) every CTA's partial of d h1 is complete
    // ---- D: my quarter of d h1: sum of the four partials (rank order), ReLU mask of h1 ----
    for (int o = t; o < Q * ROWS; o += T) {
      const int oo = rank * (Q * ROWS) + o;
      const float s4 = ((rem[0][O_G1P + oo] + rem[1][O_G1P + oo]) + rem[2][O_G1P + oo]) + rem[3][O_G1P + oo];
      g1T[o] = h1T[oo] > 0.f ? s4 : 0.f;
    }
    __syncthreads();
    // ---- E: gzp[j][r] = sum_{u mine} g1[u][r] W0t[j][u]; thread = (j, rg: rows 4rg.., uh: half of my u) ----
    {
      const int j = t & 63, rg = (t >> 6) & 1, uh = t >> 7;
      float a[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
      for (int u = uh * 32; u < uh * 32 + 32; ++u) {
        const float w = W0t[j * WS + u];
        const float4 g = *reinterpret_cast<const float4*>(g1T + u * ROWS + 4 * rg);
        a[0] = fmaf(g.x, w, a[0]); a[1] = fmaf(g.y, w, a[1]); a[2] = fmaf(g.z, w, a[2]); a[3] = fmaf(g.w, w, a[3]);
      }
      *reinterpret_cast<float4*>(red + uh * (L * ROWS) + j * ROWS + 4 * rg) = make_float4(a[0], a[1], a[2], a[3]);
    }
    __syncthreads();
    for (int o = t; o < L * ROWS; o += T) gzp[o] = red[o] + red[L * ROWS + o];
    cluster.sync();                            // (3) every CTA's partial of d z is complete
    // ---- F: the step (all four CTAs apply the same sums in the same order: their copies of z stay identical) ----
    for (int o = t; o < L * ROWS; o += T)
      zT[o] = fmaf(p.lr, ((rem[0][O_GZP + o] + rem[1][O_GZP + o]) + rem[2][O_GZP + o]) + rem[3][O_GZP + o], zT[o]);
    __syncthreads();
  }

  // ---- cost of the final z: cost[r] = bd + sum_u wd[u] h2[r, u], reduced over the four CTAs by rank 0 ----
  if (t < ROWS) {
    float c = 0.f;
    for (int u = 0; u < Q; ++u) c = fmaf(wd[u], h2T[u * ROWS + t], c);
    costp[t] = c;
  }
  cluster.sync();
  if (rank == 0) {
    if (t < ROWS && row0 + t < p.n)
      p.cost[row0 + t] = (((rem[0][O_COST + t] + rem[1][O_COST + t]) + rem[2][O_COST + t]) + rem[3][O_COST + t]) + p.bd[0];
    for (int i = t; i < L * ROWS; i += T) { const int k = i >> 3, r = i & 7; if (row0 + r < p.n) p.z[(size_t)(row0 + r) * L + k] = zT[i]; }
  }
  cluster.sync();                              // nobody leaves while its shared memory may still be read
}

}  // namespace asc

// z [n, 64] in place, cost [n]; returns VAEX_E_UNSUPPORTED for other shapes (the caller then takes the launch-per-layer path)
int ascent_cluster_run(vaex_model* m, float* z, int64_t n, int steps, float lr, float* cost, cudaStream_t st) {
  using namespace asc;
  if (m->L != L || m->PH != PH) return VAEX_E_UNSUPPORTED;
  Params p;
  p.W0 = m->w("cost_predictor.0.weight"); p.b0 = m->w("cost_predictor.0.bias");
  p.W1 = m->w("cost_predictor.3.weight"); p.b1 = m->w("cost_predictor.3.bias");
  p.wd = m->w("cost_predictor.6.weight"); p.bd = m->w("cost_predictor.6.bias");
  VAEX_CHECK(p.W0 && p.b0 && p.W1 && p.b1 && p.wd && p.bd, "cost predictor weights missing");
  p.z = z; p.cost = cost; p.n = (int)n; p.steps = steps; p.lr = lr;
  // per device and cheap: no process-wide "done" flag
  VAEX_CUDA(cudaFuncSetAttribute(ascent_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
  const unsigned clusters = (unsigned)cdiv(n, ROWS);
  ascent_cluster_kernel<<<clusters * CL, T, SMEM_BYTES, st>>>(p);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

}  // namespace vaex
