// k-center greedy (farthest-first) diversity selection on L2-normalised latents
// (select_topk_latent_diversity, scripts/pre_experiments/model_myself/utils/select.py:185-270).
//
// The reference runs, per selected point, an argmax + .item() host sync + a chunked torch.cdist; here the
// whole selection is three launches with no host round trip:
//   1. normalise the gathered candidate / centre rows        z / (||z|| + eps)            (:210-211)
//   2. min_d[j] = min_u ||zc_j - zu_u||  (one CTA per candidate; +inf when there are no centres, :224-244)
//   3. ONE CTA loops k times: argmax(min_d) (first index on ties, like torch.argmax) -> record ->
//      min_d[j*] = -inf -> min_d = min(min_d, ||zc - zc_j*||)                              (:248-264)
// Distances are the direct sqrt(sum (a - b)^2) in fp32 (torch.cdist's small-case formula).
#include <cfloat>

#include "../../include/vaex_b200.h"
#include "common.cuh"

namespace vaex {

// out[i, :] = z[idx[i], :] / (||z[idx[i], :]|| + eps), or the raw row when `raw`; idx null = the identity; one warp per row
__global__ void kc_normalise_kernel(const float* __restrict__ z, int64_t n_total, int L, const int64_t* __restrict__ idx,
                                    int64_t count, float eps, float* __restrict__ out, int* __restrict__ bad, int raw = 0) {
  const int64_t i = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (i >= count) return;
  const int lane = threadIdx.x & 31;
  const int64_t r = idx ? idx[i] : i;
  if (r < 0 || r >= n_total) {   // reported to the host by the caller's status read; the row is zeroed
    if (lane == 0) atomicExch(bad, 1);
    for (int c = lane; c < L; c += 32) out[i * L + c] = 0.f;
    return;
  }
  float ss = 0.f;
  for (int c = lane; c < L; c += 32) { const float v = z[r * L + c]; ss = fmaf(v, v, ss); }
#pragma unroll
  for (int o = 16; o; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
  const float inv = raw ? 1.f : 1.f / (sqrtf(ss) + eps);
  for (int c = lane; c < L; c += 32) out[i * L + c] = z[r * L + c] * inv;
}

// farthest-point sampling with a given first pick: record it and take it out of the running
__global__ void kc_first_kernel(const int64_t* __restrict__ cand, int64_t first, float* __restrict__ min_d, int64_t* __restrict__ out_sel) {
  out_sel[0] = cand ? cand[first] : first;
  min_d[first] = -INFINITY;
}

__device__ __forceinline__ float kc_dist(const float* __restrict__ a, const float* __restrict__ b, int L) {
  float s = 0.f;
  for (int c = 0; c < L; ++c) { const float d = a[c] - b[c]; s = fmaf(d, d, s); }
  return sqrtf(s);
}

// min_d[j] = min over centres; grid = M, block = 256
__global__ void kc_init_kernel(const float* __restrict__ zc, const float* __restrict__ zu, int64_t U, int L,
                               float* __restrict__ min_d) {
  extern __shared__ float sh[];       // [L] candidate row + [8] warp minima
  float* row = sh;
  float* wmin = sh + L;
  const int64_t j = blockIdx.x;
  for (int c = threadIdx.x; c < L; c += blockDim.x) row[c] = zc[j * L + c];
  __syncthreads();
  float best = INFINITY;
  for (int64_t u = threadIdx.x; u < U; u += blockDim.x) best = fminf(best, kc_dist(row, zu + u * L, L));
#pragma unroll
  for (int o = 16; o; o >>= 1) best = fminf(best, __shfl_xor_sync(0xffffffffu, best, o));
  if ((threadIdx.x & 31) == 0) wmin[threadIdx.x >> 5] = best;
  __syncthreads();
  if (threadIdx.x == 0) {
    float b = wmin[0];
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) b = fminf(b, wmin[w]);
    min_d[j] = b;
  }
}

// one CTA, 1024 threads: the k greedy steps
__global__ void __launch_bounds__(1024)
kc_greedy_kernel(const float* __restrict__ zc, int64_t M, int L, int k, float* __restrict__ min_d,
                 const int64_t* __restrict__ cand, int64_t* __restrict__ out_sel) {
  extern __shared__ float sh[];       // [L] new centre
  __shared__ float s_val[32];
  __shared__ int s_idx[32];
  __shared__ int s_best;
  float* centre = sh;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int t = 0; t < k; ++t) {
    // argmax with the smallest index on ties (NaN never wins: the reference's inputs have none)
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    for (int64_t j = tid; j < M; j += blockDim.x) {
      const float v = min_d[j];
      if (v > bv || (v == bv && (int)j < bi)) { bv = v; bi = (int)j; }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    if (lane == 0) { s_val[warp] = bv; s_idx[warp] = bi; }
    __syncthreads();
    if (warp == 0) {
      bv = lane < (int)(blockDim.x >> 5) ? s_val[lane] : -INFINITY;
      bi = lane < (int)(blockDim.x >> 5) ? s_idx[lane] : 0x7fffffff;
#pragma unroll
      for (int o = 16; o; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
      }
      if (lane == 0) {
        if (bi == 0x7fffffff) bi = 0;          // every entry -inf cannot happen for k <= M; keep the index valid
        s_best = bi;
        out_sel[t] = cand ? cand[bi] : (int64_t)bi;
        min_d[bi] = -INFINITY;
      }
    }
    __syncthreads();
    const int jb = s_best;
    for (int c = tid; c < L; c += blockDim.x) centre[c] = zc[(int64_t)jb * L + c];
    __syncthreads();
    for (int64_t j = tid; j < M; j += blockDim.x) min_d[j] = fminf(min_d[j], kc_dist(zc + j * L, centre, L));
    __syncthreads();
  }
}

}  // namespace vaex

using namespace vaex;

extern "C" int vaex_kcenter_select(const float* z, int64_t n, int32_t L, const int64_t* cand, int64_t M, const int64_t* used,
                                   int64_t U, int32_t k, float eps, int64_t* out_sel, void* stream) {
  VAEX_CHECK(n >= 0 && L > 0 && M >= 0 && U >= 0 && k >= 0, "bad argument");
  VAEX_CHECK(k <= M, "cannot select more points than there are candidates");
  if (k == 0) return VAEX_OK;
  VAEX_CHECK(z && cand && out_sel && (used || U == 0), "null argument");
  VAEX_CHECK(M < (1ll << 31) && L <= 8192, "candidate set / latent width too large");
  cudaStream_t st = (cudaStream_t)stream;
  float* ws = nullptr;
  const size_t floats = (size_t)(M + U) * L + M + 1;
  VAEX_CUDA(scratch_alloc((void**)&ws, floats * sizeof(float), st));
  float* zc = ws;
  float* zu = zc + (size_t)M * L;
  float* min_d = zu + (size_t)U * L;
  int* bad = reinterpret_cast<int*>(min_d + M);
  int rc = VAEX_OK;
  do {
    if (cudaMemsetAsync(bad, 0, sizeof(int), st) != cudaSuccess) { rc = VAEX_E_CUDA; break; }
    kc_normalise_kernel<<<(unsigned)cdiv(M, 8), 256, 0, st>>>(z, n, L, cand, M, eps, zc, bad);
    if (U) kc_normalise_kernel<<<(unsigned)cdiv(U, 8), 256, 0, st>>>(z, n, L, used, U, eps, zu, bad);
    kc_init_kernel<<<(unsigned)M, 256, (L + 8) * sizeof(float), st>>>(zc, zu, U, L, min_d);
    kc_greedy_kernel<<<1, 1024, L * sizeof(float), st>>>(zc, M, L, k, min_d, cand, out_sel);
    if (cudaGetLastError() != cudaSuccess) { set_error("k-center launch failed"); rc = VAEX_E_CUDA; break; }
    for (int i = 0; i < (U ? 4 : 3); ++i) count_launch();
    int bad_h = 0;
    if (cudaMemcpyAsync(&bad_h, bad, sizeof(int), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
        cudaStreamSynchronize(st) != cudaSuccess) { set_error("k-center status read failed"); rc = VAEX_E_CUDA; break; }
    if (bad_h) { set_error("k-center: candidate / used index out of range"); rc = VAEX_E_INVALID; }
  } while (0);
  cudaFreeAsync(ws, st);
  return rc;
}

// Farthest-point sampling on RAW latent rows (no row normalisation) from a given first pick.
extern "C" int vaex_fps_select(const float* z, int64_t n, int32_t L, const int64_t* cand, int64_t M, int64_t first, int32_t k,
                               int64_t* out_sel, void* stream) {
  VAEX_CHECK(n >= 0 && L > 0 && M >= 0 && k >= 0, "bad argument");
  VAEX_CHECK(k <= M, "cannot select more points than there are candidates");
  if (k == 0) return VAEX_OK;
  VAEX_CHECK(z && out_sel && first >= 0 && first < M, "null argument / first pick outside the candidate list");
  VAEX_CHECK(cand || M <= n, "without a candidate list the candidates are rows [0, M) of z");
  VAEX_CHECK(M < (1ll << 31) && L <= 8192, "candidate set / latent width too large");
  cudaStream_t st = (cudaStream_t)stream;
  float* ws = nullptr;
  VAEX_CUDA(scratch_alloc((void**)&ws, ((size_t)M * L + M + 1) * sizeof(float), st));
  float* zc = ws;
  float* min_d = zc + (size_t)M * L;
  int* bad = reinterpret_cast<int*>(min_d + M);
  int rc = VAEX_OK;
  do {
    if (cudaMemsetAsync(bad, 0, sizeof(int), st) != cudaSuccess) { rc = VAEX_E_CUDA; break; }
    kc_normalise_kernel<<<(unsigned)cdiv(M, 8), 256, 0, st>>>(z, n, L, cand, M, 0.f, zc, bad, 1);
    kc_init_kernel<<<(unsigned)M, 256, (L + 8) * sizeof(float), st>>>(zc, zc + (size_t)first * L, 1, L, min_d);
    kc_first_kernel<<<1, 1, 0, st>>>(cand, first, min_d, out_sel);
    if (k > 1) kc_greedy_kernel<<<1, 1024, L * sizeof(float), st>>>(zc, M, L, k - 1, min_d, cand, out_sel + 1);
    if (cudaGetLastError() != cudaSuccess) { set_error("farthest-point launch failed"); rc = VAEX_E_CUDA; break; }
    for (int i = 0; i < (k > 1 ? 4 : 3); ++i) count_launch();
    int bad_h = 0;
    if (cudaMemcpyAsync(&bad_h, bad, sizeof(int), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
        cudaStreamSynchronize(st) != cudaSuccess) { set_error("farthest-point status read failed"); rc = VAEX_E_CUDA; break; }
    if (bad_h) { set_error("farthest-point sampling: candidate index out of range"); rc = VAEX_E_INVALID; }
  } while (0);
  cudaFreeAsync(ws, st);
  return rc;
}

// out_d[j] = min_u ||zc_j - zu_u||_2 (raw rows; +inf when U = 0)
extern "C" int vaex_latent_min_dist(const float* zc, int64_t M, const float* zu, int64_t U, int32_t L, float* out_d, void* stream) {
  VAEX_CHECK(M >= 0 && U >= 0 && L > 0 && L <= 8192 && M < (1ll << 31), "bad argument");
  if (M == 0) return VAEX_OK;
  VAEX_CHECK(zc && out_d && (zu || U == 0), "null argument");
  kc_init_kernel<<<(unsigned)M, 256, (L + 8) * sizeof(float), (cudaStream_t)stream>>>(zc, zu, U, L, out_d);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}
