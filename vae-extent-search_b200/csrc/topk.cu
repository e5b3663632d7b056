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

constexpr int TOPK_BUCKETS = 4096;
// digits of the 64-bit key, most significant first: 12 + 12 + 8 bits cover the score (three reads of the scores decide
// any threshold that is not a tie), three more digits walk the index bits when the k-th score is tied
__constant__ int TOPK_SHIFT[6] = {52, 40, 32, 20, 8, 0};
__constant__ int TOPK_BITS[6] = {12, 12, 8, 12, 12, 8};
constexpr int TOPK_PASSES = 6;

struct TopkState {
  unsigned int hist[TOPK_BUCKETS];
  unsigned long long prefix;   // threshold key bits decided so far
  unsigned int k_rem;          // how many still to take from the current prefix bucket
  unsigned int done;
  unsigned int count;          // collected so far
  unsigned int n_out;
  unsigned int ticket;         // CTAs of the current pass that have merged their histogram
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
  for (int b = threadIdx.x; b < TOPK_BUCKETS; b += blockDim.x) st->hist[b] = 0;
  if (threadIdx.x == 0) { st->prefix = 0; st->k_rem = k; st->done = 0; st->count = 0; st->n_out = 0; st->ticket = 0; }
}

// Fixes one digit of the threshold from a complete histogram; called by all 256 threads of a CTA.
template <bool GLOBAL>
__device__ __forceinline__ void decide_digit(const unsigned int* hist, TopkState* st, unsigned long long prefix, int pass,
                                             int shift, int nb, unsigned int* part) {
  // ---- decide, in parallel (a serial walk by one thread costs ~30 clk per shared and ~1 us per global load):
  // thread t owns the group of `per` buckets that is t-th from the TOP; an inclusive scan over t gives the number of
  // keys in its group and above, and exactly one thread finds the k-th key inside its own group ----
  const int per = nb / 256, t = threadIdx.x, g0 = (255 - t) * per;
  unsigned int cnt[16];
  unsigned int mine = 0;
#pragma unroll
  for (int b = 0; b < 16; ++b) {
    cnt[b] = b < per ? (GLOBAL ? __ldcg(&hist[g0 + b]) : hist[g0 + b]) : 0u;
    mine += cnt[b];
  }
  unsigned int incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned int v = __shfl_up_sync(0xffffffffu, incl, o);
    if ((t & 31) >= o) incl += v;
  }
  if ((t & 31) == 31) part[t >> 5] = incl;
  __syncthreads();
  unsigned int above = 0;                       // keys in the groups above this thread's
  for (int w = 0; w < (t >> 5); ++w) above += part[w];
  incl += above;
  above = incl - mine;
  const unsigned int k = st->k_rem;
  __syncthreads();
  if (t == 255 && incl < k) {                   // fewer eligible than k: take everything (only reachable in pass 0)
    st->prefix = 0; st->done = 1; st->n_out = incl;
  }
  if (above < k && incl >= k) {                 // the k-th key is in this thread's group
    unsigned int cum = above, c = 0;
    int b = 0;
    bool found = false;
#pragma unroll
    for (int q = 15; q >= 0; --q) {
      if (q < per && !found) {
        if (cum + cnt[q] >= k) { found = true; b = q; c = cnt[q]; } else cum += cnt[q];
      }
    }
    st->prefix = prefix | ((unsigned long long)(g0 + b) << shift);
    st->k_rem = k - cum;
    if (c == k - cum || pass == TOPK_PASSES - 1) st->done = 1;
  }
}

// One launch per digit: every CTA histograms its share of the still-eligible keys in shared memory and merges it into
// the global histogram; the CTA that merges last (ticket) walks the buckets from the top, fixes the digit of the
// threshold and clears the histogram for the next pass.  Launches after the threshold is isolated return at once.
__global__ void __launch_bounds__(256)
topk_pass(const float* __restrict__ s, const uint8_t* __restrict__ mask, int64_t n, TopkState* st, int pass) {
  if (st->done) return;
  __shared__ unsigned int h[TOPK_BUCKETS];
  __shared__ unsigned int part[8];
  __shared__ int is_last;
  const int shift = TOPK_SHIFT[pass], nb = 1 << TOPK_BITS[pass], hs = shift + TOPK_BITS[pass];
  for (int b = threadIdx.x; b < nb; b += 256) h[b] = 0;
  __syncthreads();
  const unsigned long long prefix = st->prefix;
#pragma unroll 4
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    if (mask && !mask[i]) continue;
    const unsigned long long key = make_key(s[i], i);
    if (pass > 0 && (key >> hs) != (prefix >> hs)) continue;
    const unsigned int bucket = (unsigned int)(key >> shift) & (nb - 1);
    if (pass == 0) {
      // every key takes part and scores of one search cluster in a few exponents: one atomic per distinct bucket of the
      // warp instead of up to 32 serialised ones on the same address
      const unsigned int peers = __match_any_sync(__activemask(), bucket);
      if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&h[bucket], __popc(peers));
    } else {
      atomicAdd(&h[bucket], 1u);
    }
  }
  __syncthreads();
  for (int b = threadIdx.x; b < nb; b += 256)
    if (h[b]) atomicAdd(&st->hist[b], h[b]);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(&st->ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  decide_digit<true>(st->hist, st, prefix, pass, shift, nb, part);
  if (threadIdx.x == 0) st->ticket = 0;
  __syncthreads();
  for (int b = threadIdx.x; b < nb; b += 256) st->hist[b] = 0;
}

// Few scores (config 1's 4 096 latents, the 5k pools of select_programs): the whole select in ONE CTA -- histogram in shared
// memory, the same digit decision, the collect -- so that the call is 2 launches (this + the sort) instead of 9; at this
// size the launches, not the bytes, are the cost.
constexpr int TOPK_SMALL_N = 32768;
__global__ void __launch_bounds__(256)
topk_small(const float* __restrict__ s, const uint8_t* __restrict__ mask, int n, TopkState* st, int k) {
  __shared__ unsigned int h[TOPK_BUCKETS];
  __shared__ unsigned int part[8];
  if (threadIdx.x == 0) { st->prefix = 0; st->k_rem = k; st->done = 0; st->count = 0; st->n_out = 0; st->ticket = 0; }
  __syncthreads();
  for (int pass = 0; pass < TOPK_PASSES; ++pass) {
    if (st->done) break;
    const int shift = TOPK_SHIFT[pass], nb = 1 << TOPK_BITS[pass], hs = shift + TOPK_BITS[pass];
    for (int b = threadIdx.x; b < nb; b += 256) h[b] = 0;
    const unsigned long long prefix = st->prefix;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += 256) {
      if (mask && !mask[i]) continue;
      const unsigned long long key = make_key(s[i], i);
      if (pass > 0 && (key >> hs) != (prefix >> hs)) continue;
      atomicAdd(&h[(unsigned int)(key >> shift) & (nb - 1)], 1u);
    }
    __syncthreads();
    decide_digit<false>(h, st, prefix, pass, shift, nb, part);
    __syncthreads();
  }
  const unsigned long long thr = st->prefix;
  for (int i = threadIdx.x; i < n; i += 256) {
    if (mask && !mask[i]) continue;
    const unsigned long long key = make_key(s[i], i);
    if (key >= thr) {
      const unsigned int slot = atomicAdd(&st->count, 1u);
      if (slot < TOPK_MAX) st->cand[slot] = key;
    }
  }
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
  int S = 2;                                    // the network only has to cover the candidates there are
  while (S < (int)cnt) S <<= 1;
  for (int size = 2; size <= S; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      const int j = t ^ stride;
      if (j > t && j < S) {
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
  if (k > 0 && n > 0 && n <= TOPK_SMALL_N) {
    topk_small<<<1, 256, 0, stream>>>(scores, mask, (int)n, st, k);
    VAEX_LAUNCH_CHECK();
    topk_sort<<<1, TOPK_MAX, 0, stream>>>(st, k, out_vals, out_idx, n_out);
    VAEX_LAUNCH_CHECK();
    return VAEX_OK;
  }
  topk_init<<<1, 256, 0, stream>>>(st, k);
  VAEX_LAUNCH_CHECK();
  if (k > 0 && n > 0) {
    // few, fat CTAs: every CTA pays for clearing and merging a 4096-bucket histogram and for one ticket (a returning
    // atomic on one address, serialised in L2): 1184 CTAs made a pass over 2 M scores 24 us, 8 MB of reads notwithstanding
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(cdiv(n, 256 * 16), 148 * 2));
    for (int pass = 0; pass < TOPK_PASSES; ++pass) {
      topk_pass<<<grid, 256, 0, stream>>>(scores, mask, n, st, pass);
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
