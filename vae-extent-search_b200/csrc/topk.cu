// Deterministic top-k by MSD radix select over a unique 64-bit composite key
//   key = (order-preserving u32 of the fp32 score) << 32 | ~index
// so the result is exactly "k largest, ties -> smaller index", independent of scheduling.
// Replaces np.argsort(pred[remaining])[-k:] (utils/select.py:54-81).  HBM-bound: each pass
// that is needed reads 4 B/score (+1 B/mask); later passes early-exit once the k-th key is
// isolated (normally after 1-3 passes).
#include <algorithm>

#include "common.cuh"
#include "simt_kernels.cuh"

namespace vaex {

constexpr int TOPK_MAX = 1024;

struct TopkState {
  unsigned int hist[256];
  unsigned long long prefix;   // threshold key bits decided so far
  unsigned int k_rem;          // how many still to take from the current prefix bucket
  unsigned int done;
  unsigned int count;          // collected so far
  unsigned int n_out;
  unsigned long long cand[TOPK_MAX];
};

size_t topk_workspace_bytes() { return sizeof(TopkState); }

__device__ __forceinline__ unsigned long long make_key(float s, int64_t idx) {
  unsigned int u = __float_as_uint(s);
  if (s != s) u = 0u;  // NaN sorts last
  else u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  return ((unsigned long long)u << 32) | (unsigned int)(~(unsigned int)idx);
}
__device__ __forceinline__ float key_score(unsigned long long key) {
  unsigned int u = (unsigned int)(key >> 32);
  u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
  return __uint_as_float(u);
}

__global__ void topk_init(TopkState* st, int k) {
  if (threadIdx.x < 256) st->hist[threadIdx.x] = 0;
  if (threadIdx.x == 0) { st->prefix = 0; st->k_rem = k; st->done = 0; st->count = 0; st->n_out = 0; }
}

__global__ void __launch_bounds__(256)
topk_hist(const float* __restrict__ s, const uint8_t* __restrict__ mask, int64_t n, TopkState* st, int pass) {
  if (st->done) return;
  __shared__ unsigned int h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  const unsigned long long prefix = st->prefix;
  const int hs = 64 - 8 * pass;  // bits above hs are decided
  const int ds = 56 - 8 * pass;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    if (mask && !mask[i]) continue;
    const unsigned long long key = make_key(s[i], i);
    if (pass > 0 && (key >> hs) != (prefix >> hs)) continue;
    atomicAdd(&h[(unsigned int)(key >> ds) & 255u], 1u);
  }
  __syncthreads();
  if (h[threadIdx.x]) atomicAdd(&st->hist[threadIdx.x], h[threadIdx.x]);
}

__global__ void topk_decide(TopkState* st, int pass) {
  __shared__ unsigned int was_done;
  if (threadIdx.x == 0) was_done = st->done;
  __syncthreads();
  if (was_done) return;
  if (threadIdx.x == 0) {
    const int ds = 56 - 8 * pass;
    unsigned int cum = 0, k = st->k_rem;
    int b = 255;
    for (; b >= 0; --b) {
      const unsigned int c = st->hist[b];
      if (cum + c >= k) break;
      cum += c;
    }
    if (b < 0) {            // fewer eligible than k: take everything
      st->prefix = 0; st->done = 1; st->n_out = cum;  // (only reachable in pass 0)
    } else {
      st->prefix |= (unsigned long long)b << ds;
      st->k_rem = k - cum;
      if (st->hist[b] == k - cum || pass == 7) st->done = 1;
    }
  }
  __syncthreads();
  if (threadIdx.x < 256) st->hist[threadIdx.x] = 0;
}

__global__ void __launch_bounds__(256)
topk_collect(const float* __restrict__ s, const uint8_t* __restrict__ mask, int64_t n, TopkState* st) {
  const unsigned long long thr = st->prefix;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    if (mask && !mask[i]) continue;
    const unsigned long long key = make_key(s[i], i);
    if (key >= thr) {
      const unsigned int slot = atomicAdd(&st->count, 1u);
      if (slot < TOPK_MAX) st->cand[slot] = key;
    }
  }
}

__global__ void __launch_bounds__(TOPK_MAX)
topk_sort(TopkState* st, int k, float* __restrict__ out_vals, int64_t* __restrict__ out_idx,
          int32_t* __restrict__ n_out) {
  __shared__ unsigned long long a[TOPK_MAX];
  const unsigned int cnt = min(st->count, (unsigned int)TOPK_MAX);
  const int t = threadIdx.x;
  a[t] = (unsigned int)t < cnt ? st->cand[t] : 0ull;
  __syncthreads();
  for (int size = 2; size <= TOPK_MAX; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      const int j = t ^ stride;
      if (j > t) {
        const bool desc = (t & size) == 0;
        const unsigned long long x = a[t], y = a[j];
        if (desc ? (x < y) : (x > y)) { a[t] = y; a[j] = x; }
      }
      __syncthreads();
    }
  }
  const int m = min((int)cnt, k);
  if (t < m) {
    out_vals[t] = key_score(a[t]);
    out_idx[t] = (int64_t)(unsigned int)(~(unsigned int)(a[t] & 0xffffffffull));
  }
  if (t == 0 && n_out) *n_out = m;
}

int topk(const float* scores, int64_t n, int k, const uint8_t* mask, float* out_vals,
         int64_t* out_idx, int32_t* n_out, void* workspace, cudaStream_t stream) {
  VAEX_CHECK(k >= 0 && k <= TOPK_MAX, "k must be in [0, 1024]");
  VAEX_CHECK(n >= 0 && n < (1ll << 32), "n must fit in 32 bits");
  TopkState* st = reinterpret_cast<TopkState*>(workspace);
  topk_init<<<1, 256, 0, stream>>>(st, k);
  VAEX_LAUNCH_CHECK();
  if (k > 0 && n > 0) {
    const int grid = (int)std::min<int64_t>(cdiv(n, 256), 148 * 8);
    for (int pass = 0; pass < 8; ++pass) {
      topk_hist<<<grid, 256, 0, stream>>>(scores, mask, n, st, pass);
      VAEX_LAUNCH_CHECK();
      topk_decide<<<1, 256, 0, stream>>>(st, pass);
      VAEX_LAUNCH_CHECK();
    }
    topk_collect<<<grid, 256, 0, stream>>>(scores, mask, n, st);
    VAEX_LAUNCH_CHECK();
  }
  topk_sort<<<1, TOPK_MAX, 0, stream>>>(st, k, out_vals, out_idx, n_out);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

}  // namespace vaex
