// fp32 SIMT kernels: the exact-precision path for the small extent-VAE family and the
// on-GPU check path (VAEX_F_FORCE_SIMT) for the tcgen05 SegmentSumMLP kernel.
#include "common.cuh"
#include "simt_kernels.cuh"

namespace vaex {

// ---------------------------------------------------------------------------------------
// Tiled SGEMM with fused epilogue.  64x64 output tile, BK=16, 256 threads, 4x4 per thread.
// ---------------------------------------------------------------------------------------
constexpr int BM = 64, BN = 64, BK = 16;

template <int EPI>
__global__ void __launch_bounds__(256)
linear_kernel(const float* __restrict__ X, int64_t ldx, const float* __restrict__ W,
              const float* __restrict__ bias, float* __restrict__ Y, int64_t ldy, int64_t N, int M,
              int K, const float* __restrict__ aux, int64_t ldaux) {
  __shared__ float sX[BK][BM + 4];
  __shared__ float sW[BK][BN + 4];
  const int tid = threadIdx.x;
  const int64_t n0 = (int64_t)blockIdx.x * BM;
  const int m0 = blockIdx.y * BN;
  const int lr = tid >> 2;         // 0..63 : row inside the tile this thread loads
  const int lk = (tid & 3) * 4;    // 0,4,8,12
  const int tn = (tid >> 4) * 4;   // output rows  tn..tn+3
  const int tm = (tid & 15) * 4;   // output cols  tm..tm+3
  float acc[4][4] = {};
  const bool vec = ((K & 3) == 0) && ((ldx & 3) == 0);
  for (int k0 = 0; k0 < K; k0 += BK) {
    float xv[4] = {0.f, 0.f, 0.f, 0.f}, wv[4] = {0.f, 0.f, 0.f, 0.f};
    const int64_t n = n0 + lr;
    const int m = m0 + lr;
    const int k = k0 + lk;
    if (n < N) {
      if (vec && k + 3 < K) {
        float4 t = __ldg(reinterpret_cast<const float4*>(X + n * ldx + k));
        xv[0] = t.x; xv[1] = t.y; xv[2] = t.z; xv[3] = t.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) if (k + j < K) xv[j] = __ldg(X + n * ldx + k + j);
      }
    }
    if (m < M) {
      if (((K & 3) == 0) && k + 3 < K) {
        float4 t = __ldg(reinterpret_cast<const float4*>(W + (int64_t)m * K + k));
        wv[0] = t.x; wv[1] = t.y; wv[2] = t.z; wv[3] = t.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) if (k + j < K) wv[j] = __ldg(W + (int64_t)m * K + k + j);
      }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) { sX[lk + j][lr] = xv[j]; sW[lk + j][lr] = wv[j]; }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float4 a = *reinterpret_cast<const float4*>(&sX[kk][tn]);
      float4 b = *reinterpret_cast<const float4*>(&sW[kk][tm]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t n = n0 + tn + i;
    if (n >= N) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int m = m0 + tm + j;
      if (m >= M) continue;
      float v = acc[i][j] + (bias ? bias[m] : 0.f);
      if (EPI == EPI_RELU) v = fmaxf(v, 0.f);
      if (EPI == EPI_RELU_RESIDUAL) v = fmaxf(v, 0.f) + aux[n * ldaux + m];
      if (EPI == EPI_MUL_MASK) v = aux[n * ldaux + m] > 0.f ? v : 0.f;
      if (EPI == EPI_ADD_AUX) v += aux[n * ldaux + m];
      Y[n * ldy + m] = v;
    }
  }
}

int linear_simt(const float* X, int64_t ldx, const float* W, const float* b, float* Y, int64_t ldy,
                int64_t N, int M, int K, int epi, const float* aux, int64_t ldaux,
                cudaStream_t st) {
  if (N == 0) return VAEX_OK;
  dim3 grid((unsigned)cdiv(N, BM), (unsigned)cdiv(M, BN));
  switch (epi) {
    case EPI_NONE: linear_kernel<EPI_NONE><<<grid, 256, 0, st>>>(X, ldx, W, b, Y, ldy, N, M, K, aux, ldaux); break;
    case EPI_RELU: linear_kernel<EPI_RELU><<<grid, 256, 0, st>>>(X, ldx, W, b, Y, ldy, N, M, K, aux, ldaux); break;
    case EPI_RELU_RESIDUAL: linear_kernel<EPI_RELU_RESIDUAL><<<grid, 256, 0, st>>>(X, ldx, W, b, Y, ldy, N, M, K, aux, ldaux); break;
    case EPI_MUL_MASK: linear_kernel<EPI_MUL_MASK><<<grid, 256, 0, st>>>(X, ldx, W, b, Y, ldy, N, M, K, aux, ldaux); break;
    case EPI_ADD_AUX: linear_kernel<EPI_ADD_AUX><<<grid, 256, 0, st>>>(X, ldx, W, b, Y, ldy, N, M, K, aux, ldaux); break;
    default: set_error("bad epilogue"); return VAEX_E_INVALID;
  }
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

// ---------------------------------------------------------------------------------------
// y[n] = act(dot(X[n, :K], w) + b)   -- the out_dim=1 decoder / last predictor layer.
// one warp per row.
// ---------------------------------------------------------------------------------------
__global__ void rowdot_kernel(const float* __restrict__ X, int64_t ldx, const float* __restrict__ w,
                              const float* __restrict__ b, float* __restrict__ y, int64_t N, int K,
                              int sigmoid, const uint8_t* __restrict__ valid) {
  const int64_t n = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (n >= N) return;
  const int lane = threadIdx.x & 31;
  float s = 0.f;
  for (int k = lane; k < K; k += 32) s = fmaf(X[n * ldx + k], w[k], s);
#pragma unroll
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) {
    s += b[0];
    if (sigmoid) s = 1.f / (1.f + expf(-s));
    if (valid && !valid[n]) s = -INFINITY;
    y[n] = s;
  }
}

int rowdot(const float* X, int64_t ldx, const float* w, const float* b, float* y, int64_t N, int K,
           int sigmoid, const uint8_t* valid, cudaStream_t st) {
  if (N == 0) return VAEX_OK;
  rowdot_kernel<<<(unsigned)cdiv(N, 8), 256, 0, st>>>(X, ldx, w, b, y, N, K, sigmoid, valid);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

// ---------------------------------------------------------------------------------------
// stand-alone segment sum: out[p, :] = sum_{r in [off[p], off[p+1])} rows[r - row_base, :]
// one CTA of H threads per program (coalesced over the hidden dim), deterministic order.
// ---------------------------------------------------------------------------------------
__global__ void segsum_kernel(const float* __restrict__ rows, int64_t row_base,
                              const int32_t* __restrict__ off, int64_t p0, int64_t P, int H,
                              float* __restrict__ out) {
  const int64_t p = p0 + blockIdx.x;
  if (p >= p0 + P) return;
  const int64_t r0 = off[p] - row_base, r1 = off[p + 1] - row_base;
  for (int c = threadIdx.x; c < H; c += blockDim.x) {
    float s = 0.f;
    for (int64_t r = r0; r < r1; ++r) s += rows[r * H + c];
    out[(p - p0) * H + c] = s;
  }
}

int segsum_simt(const float* rows, int64_t row_base, const int32_t* off, int64_t p0, int64_t P,
                int H, float* out, cudaStream_t st) {
  if (P == 0) return VAEX_OK;
  segsum_kernel<<<(unsigned)P, H < 256 ? H : 256, 0, st>>>(rows, row_base, off, p0, P, H, out);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

// valid[p] = any(feats[rows of p] != 0)  -- MLPModel.predict's invalid rule (mlp_model.py:842-844)
__global__ void program_valid_kernel(const float* __restrict__ feats, const int32_t* __restrict__ off,
                                     int64_t P, int F, uint8_t* __restrict__ valid) {
  const int64_t p = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (p >= P) return;
  const int lane = threadIdx.x & 31;
  const int64_t e0 = (int64_t)off[p] * F, e1 = (int64_t)off[p + 1] * F;
  int any = 0;
  for (int64_t e = e0 + lane; e < e1 && !any; e += 32) any |= (feats[e] != 0.f);
  any = __any_sync(0xffffffffu, any);
  if (lane == 0) valid[p] = (uint8_t)any;
}

int program_valid(const float* feats, const int32_t* off, int64_t P, int F, uint8_t* valid,
                  cudaStream_t st) {
  if (P == 0) return VAEX_OK;
  program_valid_kernel<<<(unsigned)cdiv(P, 8), 256, 0, st>>>(feats, off, P, F, valid);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

}  // namespace vaex
