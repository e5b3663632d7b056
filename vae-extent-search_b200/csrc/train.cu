// Training step of the SegmentSumMLP cost model on the GPU (fp32):
//   forward with saved activations -> ranking / regression loss and dL/dpred -> backward ->
//   global-norm clip + Adam (or SGD) on one flat parameter buffer.
//   Reference: MLPModelInternal._fit_a_model / _fine_tune_a_model, mlp_model.py:512-600,609-651;
//   losses RankNetLoss :869-876, LambdaRankLoss :879-925, ListNetLoss :928-944, MSELoss.
// A training batch is 512 programs (~8.4 K rows): every GEMM is ~1 GFLOP, so the step is launch-
// bound; the kernels are fp32 SIMT (gradients match the fp32 reference to ~1e-5) and every
// reduction is two-pass deterministic (no float atomics).  The flat parameter / gradient / Adam
// buffers belong to the caller (torch tensors), so data-parallel training is one NCCL all-reduce
// of the gradient buffer between vaex_trainer_forward_backward and vaex_adam_step.
#include <algorithm>
#include <cmath>
#include <string>
#include <vector>

#include "../../include/vaex_b200.h"
#include "common.cuh"
#include "model.cuh"
#include "simt_kernels.cuh"

namespace vaex {

// ---------------------------------------------------------------------------------------------
// dW[m, k] = sum_n dY[n, m] * X[n, k]   (TN GEMM, reduction over the long dimension n)
// 64 x 64 output tile, 16 rows of n per stage, blockIdx.z = split of n; partials [splits][M*K].
// ---------------------------------------------------------------------------------------------
constexpr int WG_T = 64, WG_N = 16;

__global__ void __launch_bounds__(256)
wgrad_partial_kernel(const float* __restrict__ dY, int64_t ldy, const float* __restrict__ X, int64_t ldx, int64_t N,
                     int M, int K, int64_t rows_per_split, float* __restrict__ part) {
  __shared__ float sY[WG_N][WG_T + 4];
  __shared__ float sX[WG_N][WG_T + 4];
  const int tid = threadIdx.x;
  const int k0 = blockIdx.x * WG_T, m0 = blockIdx.y * WG_T;
  const int64_t nb = (int64_t)blockIdx.z * rows_per_split, ne = min(N, nb + rows_per_split);
  const int lr = tid >> 4;            // 0..15: row of the stage this thread loads
  const int lc = (tid & 15) * 4;      // 0..60: first of 4 columns
  const int tm = (tid >> 4) * 4, tk = (tid & 15) * 4;
  float acc[4][4] = {};
  for (int64_t n0 = nb; n0 < ne; n0 += WG_N) {
    const int64_t n = n0 + lr;
    float yv[4] = {0.f, 0.f, 0.f, 0.f}, xv[4] = {0.f, 0.f, 0.f, 0.f};
    if (n < ne) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (m0 + lc + j < M) yv[j] = __ldg(dY + n * ldy + m0 + lc + j);
        if (k0 + lc + j < K) xv[j] = __ldg(X + n * ldx + k0 + lc + j);
      }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) { sY[lr][lc + j] = yv[j]; sX[lr][lc + j] = xv[j]; }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < WG_N; ++r) {
      const float4 a = *reinterpret_cast<const float4*>(&sY[r][tm]);
      const float4 b = *reinterpret_cast<const float4*>(&sX[r][tk]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
  }
  float* out = part + (size_t)blockIdx.z * M * K;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (m0 + tm + i < M && k0 + tk + j < K) out[(size_t)(m0 + tm + i) * K + k0 + tk + j] = acc[i][j];
}

// column sums of dY [N, M] (bias gradients): partials [splits][M]
__global__ void colsum_partial_kernel(const float* __restrict__ dY, int64_t ldy, int64_t N, int M, int64_t rows_per_split,
                                      float* __restrict__ part) {
  __shared__ float red[8][33];
  const int m = blockIdx.x * 32 + threadIdx.x;
  const int64_t nb = (int64_t)blockIdx.y * rows_per_split, ne = min(N, nb + rows_per_split);
  float s = 0.f;
  if (m < M)
    for (int64_t n = nb + threadIdx.y; n < ne; n += 8) s += __ldg(dY + n * ldy + m);
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && m < M) {
    float t = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) t += red[r][threadIdx.x];
    part[(size_t)blockIdx.y * M + m] = t;
  }
}

__global__ void reduce_partials_kernel(const float* __restrict__ part, int splits, int64_t count, float* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  float s = 0.f;
  for (int z = 0; z < splits; ++z) s += part[(size_t)z * count + i];
  out[i] = s;
}

__global__ void transpose_kernel(const float* __restrict__ W, int rows, int cols, float* __restrict__ Wt) {
  __shared__ float tile[32][33];
  int c = blockIdx.x * 32 + threadIdx.x, r = blockIdx.y * 32 + threadIdx.y;
  for (int j = 0; j < 32; j += 8)
    if (r + j < rows && c < cols) tile[threadIdx.y + j][threadIdx.x] = W[(size_t)(r + j) * cols + c];
  __syncthreads();
  c = blockIdx.y * 32 + threadIdx.x;   // column of Wt = row of W
  r = blockIdx.x * 32 + threadIdx.y;   // row of Wt = column of W
  for (int j = 0; j < 32; j += 8)
    if (r + j < cols && c < rows) Wt[(size_t)(r + j) * rows + c] = tile[threadIdx.x][threadIdx.y + j];
}

// dZ = dY * ((A - B) > 0): ReLU backward where relu(z) = A - B was stored as a residual sum (B may be null)
__global__ void relu_bwd_kernel(const float* __restrict__ dY, const float* __restrict__ A, const float* __restrict__ B,
                                int64_t count, float* __restrict__ dZ) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const float r = B ? A[i] - B[i] : A[i];
  dZ[i] = r > 0.f ? dY[i] : 0.f;
}

// segment-sum backward fused with the row encoder's last ReLU: dZ[row, :] = dS[prog(row), :] * (B[row, :] > 0)
__global__ void segsum_bwd_kernel(const float* __restrict__ dS, const int32_t* __restrict__ off, int64_t P, int H,
                                  const float* __restrict__ B, float* __restrict__ dZ) {
  const int64_t p = blockIdx.x;
  if (p >= P) return;
  const int r0 = off[p], r1 = off[p + 1];
  for (int j = threadIdx.x; j < H; j += blockDim.x) {
    const float g = dS[p * H + j];
    for (int r = r0; r < r1; ++r) dZ[(size_t)r * H + j] = B[(size_t)r * H + j] > 0.f ? g : 0.f;
  }
}

// dU[p, j] = dout[p] * w[j]
__global__ void outer_kernel(const float* __restrict__ dout, const float* __restrict__ w, int64_t P, int H, float* __restrict__ dU) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P * H) return;
  dU[i] = dout[i / H] * w[i % H];
}

// shuffled mini-batch assembly on the device (SegmentDataLoader._fetch_indices, mlp_model.py:129-141)
__global__ void gather_sizes_kernel(const int32_t* __restrict__ off, const int64_t* __restrict__ idx, int64_t n,
                                    int32_t* __restrict__ sizes) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) sizes[i] = off[idx[i] + 1] - off[idx[i]];
}
__global__ void gather_rows_kernel(const float* __restrict__ feats, int F, const int32_t* __restrict__ off,
                                   const int64_t* __restrict__ idx, const int32_t* __restrict__ out_off,
                                   float* __restrict__ out, const float* __restrict__ labels, float* __restrict__ out_labels) {
  const int64_t i = blockIdx.x;
  const int64_t src = (int64_t)off[idx[i]] * F, dst = (int64_t)out_off[i] * F;
  const int64_t cnt = (int64_t)(out_off[i + 1] - out_off[i]) * F;
  for (int64_t j = threadIdx.x; j < cnt; j += blockDim.x) out[dst + j] = feats[src + j];
  if (labels && threadIdx.x == 0) out_labels[i] = labels[idx[i]];
}

struct GradWs {
  float* part = nullptr;   // partial sums [splits][<= 256*256]
  size_t floats = 0;
};

static int splits_for(int64_t N) { return (int)std::max<int64_t>(1, std::min<int64_t>(32, cdiv(N, 256))); }

static int wgrad(const float* dY, int64_t ldy, const float* X, int64_t ldx, int64_t N, int M, int K, float* part, float* dW,
                 cudaStream_t st) {
  const int splits = splits_for(N);
  const int64_t rps = cdiv(cdiv(N, splits), WG_N) * WG_N;
  dim3 grid((unsigned)cdiv(K, WG_T), (unsigned)cdiv(M, WG_T), (unsigned)splits);
  wgrad_partial_kernel<<<grid, 256, 0, st>>>(dY, ldy, X, ldx, N, M, K, rps, part);
  VAEX_LAUNCH_CHECK();
  reduce_partials_kernel<<<(unsigned)cdiv((int64_t)M * K, 256), 256, 0, st>>>(part, splits, (int64_t)M * K, dW);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

static int colsum(const float* dY, int64_t ldy, int64_t N, int M, float* part, float* db, cudaStream_t st) {
  const int splits = splits_for(N);
  const int64_t rps = cdiv(N, splits);
  colsum_partial_kernel<<<dim3((unsigned)cdiv(M, 32), (unsigned)splits), dim3(32, 8), 0, st>>>(dY, ldy, N, M, rps, part);
  VAEX_LAUNCH_CHECK();
  reduce_partials_kernel<<<(unsigned)cdiv(M, 256), 256, 0, st>>>(part, splits, M, db);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

static int transpose_to(const float* W, int rows, int cols, float* Wt, cudaStream_t st) {
  transpose_kernel<<<dim3((unsigned)cdiv(cols, 32), (unsigned)cdiv(rows, 32)), dim3(32, 8), 0, st>>>(W, rows, cols, Wt);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

}  // namespace vaex

#include "train_loss.inc"
#include "train_step.inc"
#include "train_epoch.inc"
#include "train_vae.inc"
#include "train_segvae.inc"
#include "train_segreg.inc"
