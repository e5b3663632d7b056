// Bandwidth-bound helpers: scan, reparameterisation, integer snap / legalise, dropout,
// ReLU-mask products for the fused backward, z update.
#include <algorithm>
#include <mutex>
#include <string>

#include "common.cuh"
#include "simt_kernels.cuh"
#include "snap.cuh"
#include "philox.cuh"

namespace vaex {

// ------------------------------------------------------------------ exclusive scan (int32)
constexpr int SCAN_T = 256, SCAN_I = 8, SCAN_B = SCAN_T * SCAN_I;

__device__ __forceinline__ int block_excl_scan(int v, int* total) {
  __shared__ int wsum[SCAN_T / 32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) wsum[w] = inc;
  __syncthreads();
  if (w == 0) {
    int s = lane < SCAN_T / 32 ? wsum[lane] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= o) s += t;
    }
    if (lane < SCAN_T / 32) wsum[lane] = s;
  }
  __syncthreads();
  const int base = w ? wsum[w - 1] : 0;
  *total = wsum[SCAN_T / 32 - 1];
  __syncthreads();
  return base + inc - v;
}

__global__ void scan_block_sums(const int32_t* __restrict__ in, int64_t n, int32_t* __restrict__ bsum) {
  const int64_t base = (int64_t)blockIdx.x * SCAN_B + threadIdx.x * SCAN_I;
  int s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_I; ++i) if (base + i < n) s += in[base + i];
  int total;
  block_excl_scan(s, &total);
  if (threadIdx.x == 0) bsum[blockIdx.x] = total;
}

__global__ void scan_of_sums(int32_t* bsum, int nb) {  // single block, in place exclusive
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int b0 = 0; b0 < nb; b0 += SCAN_T) {
    const int i = b0 + threadIdx.x;
    const int v = i < nb ? bsum[i] : 0;
    int total;
    const int ex = block_excl_scan(v, &total);
    const int c = carry;
    if (i < nb) bsum[i] = c + ex;
    __syncthreads();
    if (threadIdx.x == 0) carry = c + total;
    __syncthreads();
  }
}

__global__ void scan_final(const int32_t* __restrict__ in, int64_t n, const int32_t* __restrict__ bsum,
                           int32_t* __restrict__ out) {
  const int64_t base = (int64_t)blockIdx.x * SCAN_B + threadIdx.x * SCAN_I;
  int v[SCAN_I], s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_I; ++i) { v[i] = base + i < n ? in[base + i] : 0; s += v[i]; }
  int total;
  int run = block_excl_scan(s, &total) + bsum[blockIdx.x];
#pragma unroll
  for (int i = 0; i < SCAN_I; ++i) {
    if (base + i < n) out[base + i] = run;
    run += v[i];
    if (base + i == n - 1) out[n] = run;
  }
}

int exclusive_scan_i32(const int32_t* sizes, int64_t n, int32_t* offsets, cudaStream_t st) {
  if (n == 0) {
    VAEX_CUDA(cudaMemsetAsync(offsets, 0, sizeof(int32_t), st));
    return VAEX_OK;
  }
  const int nb = (int)cdiv(n, SCAN_B);
  int32_t* bsum = nullptr;
  VAEX_CUDA(scratch_alloc((void**)&bsum, sizeof(int32_t) * nb, st));
  scan_block_sums<<<nb, SCAN_T, 0, st>>>(sizes, n, bsum);
  VAEX_LAUNCH_CHECK();
  scan_of_sums<<<1, SCAN_T, 0, st>>>(bsum, nb);
  VAEX_LAUNCH_CHECK();
  scan_final<<<nb, SCAN_T, 0, st>>>(sizes, n, bsum, offsets);
  VAEX_LAUNCH_CHECK();
  VAEX_CUDA(cudaFreeAsync(bsum, st));
  return VAEX_OK;
}

// z = mu + eps * exp(0.5 logvar)   (utils/model.py:58-61); 4 elements per thread
__global__ void reparam_kernel(const float* __restrict__ mu, const float* __restrict__ logvar,
                               const float* __restrict__ eps, uint64_t seed, uint64_t offset,
                               float* __restrict__ z, int64_t count) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t e0 = q * 4;
  if (e0 >= count) return;
  float n4[4];
  if (eps == nullptr) {
    const uint4 r = Philox(seed)((uint64_t)q, offset);
    const float r0 = sqrtf(-2.f * logf(u01(r.x))), r1 = sqrtf(-2.f * logf(u01(r.z)));
    float s0, c0, s1, c1;
    sincospif(2.f * u01(r.y), &s0, &c0);
    sincospif(2.f * u01(r.w), &s1, &c1);
    n4[0] = r0 * c0; n4[1] = r0 * s0; n4[2] = r1 * c1; n4[3] = r1 * s1;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t e = e0 + i;
    if (e < count) {
      const float n = eps ? eps[e] : n4[i];
      z[e] = mu[e] + n * expf(0.5f * logvar[e]);
    }
  }
}

int reparam(const float* mu, const float* logvar, const float* eps, uint64_t seed, uint64_t offset,
            float* z, int64_t count, cudaStream_t st) {
  if (count == 0) return VAEX_OK;
  reparam_kernel<<<(unsigned)cdiv(cdiv(count, 4), 256), 256, 0, st>>>(mu, logvar, eps, seed, offset, z, count);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

// ------------------------------------------------------------------ snap + legalise
// c | rem, for rem < 2^21: k = rint(rem / c) with the 2-ulp approximate division is the exact quotient whenever c
// divides rem (|error| < 2^-22 * k < 0.5), and k * c == rem can only hold for a divisor.  Six instructions instead of
// the ~25 of a 32-bit integer modulo.
__device__ __forceinline__ bool divides_small(int rem, float remf, int c, float cf) {
  return __float2int_rn(__fdividef(remf, cf)) * c == rem;
}

__device__ __forceinline__ int nearest_divisor(int rem, int v, int cap) {
  int bound = rem;
  if (cap > 0 && cap < bound) bound = cap;
  if (v > bound) v = bound;
  if (v < 1) v = 1;
  if (rem < (1 << 21)) {
    const float remf = (float)rem;
    float lof = (float)v, hif = lof;
    for (int lo = v, hi = v;; --lo, ++hi, lof -= 1.f, hif += 1.f) {
      if (lo >= 1 && divides_small(rem, remf, lo, lof)) return lo;   // ties -> smaller
      if (hi <= bound && divides_small(rem, remf, hi, hif)) return hi;
    }
  }
  for (int delta = 0;; ++delta) {
    const int lo = v - delta, hi = v + delta;
    if (lo >= 1 && rem % lo == 0) return lo;   // ties -> smaller
    if (hi <= bound && rem % hi == 0) return hi;
  }
}

__global__ void snap_kernel(const float* __restrict__ x, int64_t n, int d, const float* __restrict__ mean,
                            const float* __restrict__ scale, int mode, int32_t* __restrict__ out) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n * d) return;
  const int j = (int)(e % d);
  out[e] = snap_one(x[e], mean[j], scale[j], mode);
}

__global__ void legalise_kernel(int32_t* __restrict__ ext, int64_t n, int d,
                                const int32_t* __restrict__ slot_extent,
                                const uint8_t* __restrict__ slot_last, int max_innermost,
                                const int32_t* __restrict__ order, int n_order) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int rem = 0;
  // walk order: the slots as printed (order == nullptr), or the caller's list in which the lengths of one
  // SplitStep are consecutive (a printed loop nest interleaves the splits of different axes)
  const int steps = order ? n_order : d;
  for (int jj = 0; jj < steps; ++jj) {
    const int j = order ? order[jj] : jj;
    const int se = slot_extent[j];
    if (se <= 0) { rem = 0; continue; }
    if (rem == 0) rem = se;
    const int last = slot_last[j];
    const int q = nearest_divisor(rem, ext[i * d + j], last ? max_innermost : 0);
    ext[i * d + j] = q;
    rem /= q;
    if (last) rem = 0;
  }
}

// Same walk with the rows staged through shared memory: a thread per row reads d consecutive ints (192 B apart from its
// neighbour's for d = 48), which costs a full sector per element; here a CTA copies its 128 rows in and out with
// coalesced 4-byte accesses and the walk runs on a padded (d + 1) tile without bank conflicts.  Used for d <= 64.
//
// Answer tables for small loop extents.  Every remaining extent of a split divides the loop extent E of its slots, so for
// E <= 255 the whole function (remaining extent r | E, target v) -> nearest divisor fits a table of ndiv(E) * (E + 1)
// 16-bit entries (twice: with and without the innermost cap): low byte = the divisor q, high byte = the index of r / q
// among E's divisors.  The walk therefore carries the INDEX of the remaining extent; a slot costs one 16-byte step
// record, the row element and one table load, instead of a data-dependent search and an integer division --
// GetFactors memoises per extent on the host for the same reason (utils.cc:446-463).  Each persistent CTA builds the
// tables of its distinct extents once in shared memory, one thread per table row: a forward pass leaves the largest
// divisor <= v, a backward pass compares it with the smallest one >= v (ties -> smaller, like the +-1 search), targets
// above the bound answer like the bound.  Larger extents, whatever does not fit LEG_LUT bytes, and slot tables in which
// a split changes its extent half way keep the +-1 search.  History and measurements: profiles/r01_ncu_legalise_notes.md
// (tried and dropped: sorted divisor lists + bisection, the next tile held in registers, a software-pipelined walk).
#ifndef VAEX_LEG_ROWS
#define VAEX_LEG_ROWS 128
#define VAEX_LEG_CTAS 5
#endif
constexpr int LEG_ROWS = VAEX_LEG_ROWS, LEG_CTAS = VAEX_LEG_CTAS;   // rows (= threads) per CTA, CTAs per SM
constexpr int LEG_LUT = 12 * 1024;             // bytes of answer tables per CTA
constexpr int LEG_LUT_MAX_E = 255;

__global__ void __launch_bounds__(LEG_ROWS, LEG_CTAS) legalise_tiled_kernel(int32_t* __restrict__ ext, int64_t n, int d,
                                                                  const int32_t* __restrict__ slot_extent,
                                                                  const uint8_t* __restrict__ slot_last, int max_innermost,
                                                                  const int32_t* __restrict__ order, int n_order, int use_lut) {
  extern __shared__ __align__(16) int32_t tile[];   // [LEG_ROWS][d + 1], se[d], sl[d], so[steps], nd[d], t0[d], ridx[d], rows[d], lut bytes, step records
  const int ds = d + 1;
  const int steps = order ? n_order : d;
  int32_t* se = tile + LEG_ROWS * ds;
  int32_t* sl = se + d;
  int32_t* so = sl + d;
  int32_t* nd = so + steps;                    // owner slots: number of divisors (sharers: minus that); no table: 0
  int32_t* t0 = nd + d;                        // first task (table row) of an owner slot
  int32_t* ridx = t0 + d;                      // byte offset of the extent's divisor-index map [E + 1], -1: no table
  int32_t* rows = ridx + d;                    // byte offset of its answer rows: [2][ndiv][E + 1]
  uint8_t* lut = reinterpret_cast<uint8_t*>(tile + ((LEG_ROWS * ds + 6 * d + steps + 3) & ~3));   // 16-byte aligned, still a shared-space pointer
  __shared__ int n_tasks;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int j = tid; j < d; j += LEG_ROWS) { se[j] = slot_extent[j]; sl[j] = slot_last[j]; nd[j] = 0; ridx[j] = -1; }
  for (int j = tid; j < steps; j += LEG_ROWS) so[j] = order ? order[j] : j;
  if (tid == 0) n_tasks = 0;
  __syncthreads();
  if (use_lut) {
    // slot j's owner = the first slot with the same (small) extent; t0 holds it until the tables are laid out.  Serial
    // scans of shared memory by one thread are what this prologue must avoid: ~30 clk per dependent load.
    for (int j = tid; j < d; j += LEG_ROWS) {
      const int E = se[j];
      int o = -1;
      if (E > 0 && E <= LEG_LUT_MAX_E) { o = 0; while (se[o] != E) ++o; }
      t0[j] = o;
    }
    __syncthreads();
    for (int j = warp; j < d; j += LEG_ROWS / 32) {      // a warp counts the divisors of an owner's extent
      if (t0[j] != j) continue;
      const int E = se[j];
      int cnt = 0;
      for (int c0 = 1; c0 <= E; c0 += 32) cnt += __popc(__ballot_sync(0xffffffffu, c0 + lane <= E && E % (c0 + lane) == 0));
      if (lane == 0) nd[j] = cnt;
    }
    __syncthreads();
    int own = tid < d ? t0[tid] : -1;                    // (d <= 64 < LEG_ROWS)
    __syncthreads();
    if (tid == 0) {
      int bytes = 0, tasks = 0;
      for (int j = 0; j < d; ++j) {
        const int cnt = nd[j];
        if (cnt == 0) continue;
        const int E = se[j], head = (E + 1 + 3) & ~3, need = head + 2 * cnt * (E + 1) * 2;
        if (bytes + need > LEG_LUT) { nd[j] = 0; continue; }
        ridx[j] = bytes;
        rows[j] = bytes + head;
        t0[j] = tasks;
        bytes = (bytes + need + 3) & ~3;
        tasks += 2 * cnt;
      }
      n_tasks = tasks;
    }
    __syncthreads();
    if (own >= 0 && own != tid && nd[own] > 0) {         // the other slots of an extent share its owner's tables
      ridx[tid] = ridx[own];
      rows[tid] = rows[own];
      nd[tid] = -nd[own];
    }
    __syncthreads();
    // one table row per task: owner slot j, variant (capped or not), i-th divisor r of E.  First the map divisor -> index,
    // then the rows: entry (i, v) = nearest divisor q of r to v | index of r / q << 8, for every v in [1, E] (targets
    // above the bound answer like the bound), so the walk carries the INDEX of the remaining extent and a slot is one
    // 16-bit load whose address does not depend on anything but the previous slot's entry.
    auto task_of = [&](int task, int& j, int& k, int& i, int& r) {
      j = 0;
      while (!(nd[j] > 0 && task >= t0[j] && task < t0[j] + 2 * nd[j])) ++j;
      const int E = se[j];
      k = task - t0[j];
      i = k >= nd[j] ? k - nd[j] : k;
      r = 0;
      for (int c = 1, seen = 0; c <= E; ++c)
        if (divides_small(E, (float)E, c, (float)c) && seen++ == i) { r = c; break; }
    };
    for (int task = tid; task < n_tasks; task += LEG_ROWS) {
      int j, k, i, r;
      task_of(task, j, k, i, r);
      if (k < nd[j]) lut[ridx[j] + r] = (uint8_t)i;
    }
    __syncthreads();
    for (int task = tid; task < n_tasks; task += LEG_ROWS) {
      int j, k, i, r;
      task_of(task, j, k, i, r);
      const int E = se[j], capped = k >= nd[j];
      uint16_t* trow = reinterpret_cast<uint16_t*>(lut + rows[j]) + k * (E + 1);
      const uint8_t* index_of = lut + ridx[j];
      const int bound = (capped && max_innermost > 0 && max_innermost < r) ? max_innermost : r;
      const float rf = (float)r;
      int down = 1;
      for (int v = 1; v <= bound; ++v) {       // largest divisor <= v
        if (divides_small(r, rf, v, (float)v)) down = v;
        trow[v] = (uint16_t)down;
      }
      int up = 0;
      for (int v = bound; v >= 1; --v) {       // smallest divisor in [v, bound]; ties -> smaller
        if (divides_small(r, rf, v, (float)v)) up = v;
        const int dn = trow[v];
        const int q = (up != 0 && up - v < v - dn) ? up : dn;
        trow[v] = (uint16_t)(q | (index_of[__float2int_rn(__fdividef(rf, (float)q))] << 8));
      }
      for (int v = bound + 1; v <= E; ++v) trow[v] = trow[bound];
    }
    __syncthreads();
  }
  // one 16-byte record per walk step: {slot | last << 8, extent, uint16 offset of the answer rows of this slot's variant
  // (-1: search), index of E among its divisors}
  int4* stp = reinterpret_cast<int4*>(lut + LEG_LUT);
  const uint16_t* tab16 = reinterpret_cast<const uint16_t*>(lut);
  for (int jj = tid; jj < steps; jj += LEG_ROWS) {
    const int j = so[jj], E = se[j], last = sl[j] ? 1 : 0, ndiv = nd[j] < 0 ? -nd[j] : nd[j];
    stp[jj] = make_int4(j | (last << 8), E, ridx[j] >= 0 ? rows[j] / 2 + (last ? ndiv * (E + 1) : 0) : -1, ndiv - 1);
  }
  __syncthreads();
  // slot tables in which a split changes its extent half way (no innermost mark in between) carry the remaining extent
  // over; only the search reproduces that, so such tables switch the answer tables off
  if (tid == 0 && n_tasks > 0) {
    bool open = false, odd = false;
    int prev = 0;
    for (int jj = 0; jj < steps; ++jj) {
      const int E = stp[jj].y;
      if (E <= 0) { open = false; continue; }
      if (open && E != prev) odd = true;
      open = (stp[jj].x >> 8) == 0;
      prev = E;
    }
    if (odd) n_tasks = -1;
  }
  __syncthreads();
  if (n_tasks < 0) {
    for (int jj = tid; jj < steps; jj += LEG_ROWS) stp[jj].z = -1;
    __syncthreads();
  }
  const int64_t n_tiles = (n + LEG_ROWS - 1) / LEG_ROWS;
  const bool vec4 = d % 4 == 0 && (reinterpret_cast<uintptr_t>(ext) & 15) == 0;   // a tile starts 128 * d * 4 bytes further: aligned too
  const int dq = vec4 ? d / 4 : d;             // units per row (int4 or int)
  const int r_first = tid / dq, c_first = tid % dq, r_step = LEG_ROWS / dq, c_step = LEG_ROWS % dq;
  // int4 path: the 128 x d/4 units of a tile are d/4 (<= 16) per thread, moved eight at a time (eight loads in flight per
  // thread); one thread asks L2 for the next tile before this one is walked.  (Holding the next tile in 64 registers
  // instead made ptxas recompute the walk's loop invariants on every step: same time, more instructions.)
  for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    const int64_t r0 = t * LEG_ROWS;
    const int nrows = (int)min((int64_t)LEG_ROWS, n - r0);
    int32_t* g = ext + r0 * d;
    // (row, column) of element tid + 128 k kept incrementally: a division by the run-time d per element costs more than the walk
    if (vec4) {
      const int4* g4 = reinterpret_cast<const int4*>(g);
      const int units = nrows * dq;
      if (tid == 0 && t + gridDim.x < n_tiles) {
        const int64_t tn = t + gridDim.x;
        const uint32_t bytes = (uint32_t)min((int64_t)LEG_ROWS, n - tn * LEG_ROWS) * (uint32_t)d * 4u;
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(ext + tn * LEG_ROWS * d), "r"(bytes) : "memory");
      }
      int r = r_first, c = c_first;
      for (int e0 = tid; e0 < units; e0 += 8 * LEG_ROWS) {
        int4 x[8];
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (e0 + LEG_ROWS * k < units) x[k] = __ldcs(g4 + e0 + LEG_ROWS * k);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          if (e0 + LEG_ROWS * k < units) {
            int32_t* w = tile + r * ds + 4 * c;
            w[0] = x[k].x; w[1] = x[k].y; w[2] = x[k].z; w[3] = x[k].w;
          }
          r += r_step; c += c_step;
          if (c >= dq) { c -= dq; ++r; }
        }
      }
    } else {
      for (int e = tid, r = r_first, c = c_first; e < nrows * d; e += LEG_ROWS) {
        tile[r * ds + c] = g[e];
        r += r_step; c += c_step;
        if (c >= dq) { c -= dq; ++r; }
      }
    }
    __syncthreads();
    if (tid < nrows) {
      int32_t* row = tile + tid * ds;
      int rem = 0, ri = -1;                    // search groups carry the remaining extent, table groups its divisor index
      for (int jj = 0; jj < steps; ++jj) {
        const int4 sp = stp[jj];
        const int j = sp.x & 255, last = sp.x >> 8, sext = sp.y;
        if (sext <= 0) { rem = 0; ri = -1; continue; }
        int q;
        if (sp.z >= 0) {
          if (ri < 0) ri = sp.w;
          const int v = min(max(row[j], 1), sext);
          const int e = tab16[sp.z + ri * (sext + 1) + v];
          q = e & 255;
          ri = last ? -1 : e >> 8;
        } else {
          if (rem == 0) rem = sext;
          q = nearest_divisor(rem, row[j], last ? max_innermost : 0);
          rem = last ? 0 : rem / q;
        }
        row[j] = q;
      }
    }
    __syncthreads();
    if (vec4) {
      int4* g4 = reinterpret_cast<int4*>(g);
      for (int e = tid, r = r_first, c = c_first; e < nrows * dq; e += LEG_ROWS) {
        const int32_t* w = tile + r * ds + 4 * c;
        __stcs(g4 + e, make_int4(w[0], w[1], w[2], w[3]));
        r += r_step; c += c_step;
        if (c >= dq) { c -= dq; ++r; }
      }
    } else {
      for (int e = tid, r = r_first, c = c_first; e < nrows * d; e += LEG_ROWS) {
        g[e] = tile[r * ds + c];
        r += r_step; c += c_step;
        if (c >= dq) { c -= dq; ++r; }
      }
    }
    __syncthreads();
  }
}

int legalise_extents(int32_t* ext, int64_t n, int d, const int32_t* slot_extent, const uint8_t* slot_last, int max_innermost,
                     cudaStream_t st, const int32_t* order, int n_order) {
  if (n == 0) return VAEX_OK;
  if (d <= 64) {
    const int steps = order ? n_order : d;
    const size_t sh = sizeof(int32_t) * ((size_t)LEG_ROWS * (d + 1) + 6 * (size_t)d + (size_t)steps + 4) + LEG_LUT + 16 * (size_t)steps;
    static int sms = 0;
    if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
    const int64_t tiles = cdiv(n, LEG_ROWS);
    const unsigned grid = (unsigned)std::min<int64_t>(tiles, (int64_t)std::max(sms, 1) * LEG_CTAS);
    if (sh > 48 * 1024) cudaFuncSetAttribute(legalise_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh);
    // the tables cost a CTA about as much as walking one tile: worth it from a few tiles per CTA on
    legalise_tiled_kernel<<<grid, LEG_ROWS, sh, st>>>(ext, n, d, slot_extent, slot_last, max_innermost, order, n_order,
                                                      tiles >= 3 * (int64_t)grid ? 1 : 0);
  } else {
    legalise_kernel<<<(unsigned)cdiv(n, 128), 128, 0, st>>>(ext, n, d, slot_extent, slot_last, max_innermost, order, n_order);
  }
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

int snap_extents(const float* x, int64_t n, int d, const float* mean, const float* scale, int mode,
                 const int32_t* slot_extent, const uint8_t* slot_last, int max_innermost,
                 int32_t* out, cudaStream_t st) {
  if (n == 0) return VAEX_OK;
  snap_kernel<<<(unsigned)cdiv(n * d, 256), 256, 0, st>>>(x, n, d, mean, scale, mode, out);
  VAEX_LAUNCH_CHECK();
  if (mode & VAEX_SNAP_DIVISOR) {
    VAEX_CHECK(slot_extent && slot_last, "divisor mode needs slot tables");
    TRY(legalise_extents(out, n, d, slot_extent, slot_last, max_innermost, st, nullptr, 0));
  }
  return VAEX_OK;
}

// ------------------------------------------------------------------ dropout / backward helpers
// Dropout(p) in place: keep element e iff mask[e] != 0, or -- without a mask -- iff the 16-bit draw of e is >= round(p * 65536):
// Philox call q = e >> 3 yields eight 16-bit draws (word (e >> 1) & 3, half e & 1).  Sixteen bits per element instead of
// thirty-two halve the generator work, which is what bounds the MC-dropout passes of the tcgen05 predictor chain
// (chain_tc.cu draws the same bits); the drop probability is quantised to 1 / 65536 (0.100006 for p = 0.1).
__global__ void dropout_kernel(float* __restrict__ h, int64_t count, const uint8_t* __restrict__ mask,
                               uint64_t seed, float p) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t e0 = q * 8;
  if (e0 >= count) return;
  const float inv = 1.f / (1.f - p);
  const uint32_t thr = dropout_threshold(p);
  uint4 r = make_uint4(0, 0, 0, 0);
  if (!mask) r = PhiloxDrop(seed)((uint64_t)q, 0);
  const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t e = e0 + i;
    if (e < count) {
      const bool keep = mask ? mask[e] != 0 : ((rr[i >> 1] >> (16 * (i & 1))) & 0xffffu) >= thr;
      h[e] = keep ? h[e] * inv : 0.f;
    }
  }
}

int dropout_apply(float* h, int64_t n, int cols, const uint8_t* mask, uint64_t seed, float p,
                  cudaStream_t st) {
  const int64_t count = n * cols;
  if (count == 0) return VAEX_OK;
  dropout_kernel<<<(unsigned)cdiv(cdiv(count, 8), 256), 256, 0, st>>>(h, count, mask, seed, p);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

// out[n, c] = (h[n, c] > 0) ? w[c] : 0    (d cost / d pre-activation of the last hidden layer)
__global__ void relu_mask_rowvec_kernel(const float* __restrict__ h, const float* __restrict__ w,
                                        float* __restrict__ out, int64_t count, int cols) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e < count) out[e] = h[e] > 0.f ? w[e % cols] : 0.f;
}
int relu_mask_rowvec(const float* h, const float* w, float* out, int64_t n, int cols, cudaStream_t st) {
  const int64_t count = n * cols;
  if (count == 0) return VAEX_OK;
  relu_mask_rowvec_kernel<<<(unsigned)cdiv(count, 256), 256, 0, st>>>(h, w, out, count, cols);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

__global__ void axpy_kernel(float* __restrict__ z, const float* __restrict__ g, float lr, int64_t count) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e < count) z[e] = z[e] + lr * g[e];
}
int axpy(float* z, const float* g, float lr, int64_t count, cudaStream_t st) {
  if (count == 0) return VAEX_OK;
  axpy_kernel<<<(unsigned)cdiv(count, 256), 256, 0, st>>>(z, g, lr, count);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

__global__ void row_l2norm_kernel(const float* __restrict__ g, int64_t n, int cols, float* __restrict__ out) {
  const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= n) return;
  const int lane = threadIdx.x & 31;
  float s = 0.f;
  for (int c = lane; c < cols; c += 32) { const float v = g[r * cols + c]; s = fmaf(v, v, s); }
#pragma unroll
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) out[r] = sqrtf(s);
}
int row_l2norm(const float* g, int64_t n, int cols, float* out, cudaStream_t st) {
  if (n == 0) return VAEX_OK;
  row_l2norm_kernel<<<(unsigned)cdiv(n, 8), 256, 0, st>>>(g, n, cols, out);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

}  // namespace vaex

// ------------------------------------------------------------------ allocation-free exclusive scan
// vaex_segment_offsets_noalloc: the same three scan kernels as exclusive_scan_i32, but the per-block sums live in a
// file-scope device array instead of a stream-ordered allocation, for callers that may not allocate (CUDA graph
// capture, tight serving loops).  Every GPU that loads the library holds the 64 KiB until exit.  In host code the
// array's name is only a host-side stand-in: the kernels get the device address that cudaGetSymbolAddress returns,
// resolved once per device with that device current.
namespace vaex {
constexpr int64_t NOALLOC_MAX_BLOCKS = 16384;                       // 64 KiB of scratch
constexpr int64_t NOALLOC_MAX_N = NOALLOC_MAX_BLOCKS * SCAN_B;      // = VAEX_SEGMENT_OFFSETS_NOALLOC_MAX
static_assert(NOALLOC_MAX_N == VAEX_SEGMENT_OFFSETS_NOALLOC_MAX, "header and kernel disagree on the limit");
static __device__ __align__(256) int32_t g_scan_block_sums[NOALLOC_MAX_BLOCKS];

constexpr int NOALLOC_MAX_DEVICES = 64;
static int32_t* g_scan_block_sums_addr[NOALLOC_MAX_DEVICES];        // device address per device ordinal, 0 = not resolved
static std::mutex g_scan_addr_mutex;

static int scan_scratch_for_current_device(int32_t** out) {
  int dev = -1;
  VAEX_CUDA(cudaGetDevice(&dev));
  VAEX_CHECK(dev >= 0 && dev < NOALLOC_MAX_DEVICES, "device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_scan_addr_mutex);
  if (!g_scan_block_sums_addr[dev]) {
    int32_t* p = nullptr;
    VAEX_CUDA(cudaGetSymbolAddress((void**)&p, g_scan_block_sums));
    VAEX_CHECK(p != nullptr, "cudaGetSymbolAddress returned a null address");
    g_scan_block_sums_addr[dev] = p;
  }
  *out = g_scan_block_sums_addr[dev];
  return VAEX_OK;
}
}  // namespace vaex

extern "C" int vaex_segment_offsets_noalloc(const int32_t* sizes_dev, int64_t n, int32_t* offsets_dev, void* stream) {
  using namespace vaex;
  VAEX_CHECK(n >= 0 && offsets_dev && (sizes_dev || n == 0), "bad argument");
  if (n > NOALLOC_MAX_N) {
    set_error("vaex_segment_offsets_noalloc: n = " + std::to_string(n) + " exceeds the static scratch limit of " +
              std::to_string(NOALLOC_MAX_N) + " sizes (VAEX_SEGMENT_OFFSETS_NOALLOC_MAX); use vaex_segment_offsets");
    return VAEX_E_INVALID;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {
    VAEX_CUDA(cudaMemsetAsync(offsets_dev, 0, sizeof(int32_t), st));
    return VAEX_OK;
  }
  int32_t* bsum = nullptr;
  TRY(scan_scratch_for_current_device(&bsum));
  const int nb = (int)cdiv(n, SCAN_B);
  // nothing has to start at zero: scan_block_sums overwrites bsum[0, nb) before anything reads it
  scan_block_sums<<<nb, SCAN_T, 0, st>>>(sizes_dev, n, bsum);
  VAEX_LAUNCH_CHECK();
  scan_of_sums<<<1, SCAN_T, 0, st>>>(bsum, nb);
  VAEX_LAUNCH_CHECK();
  scan_final<<<nb, SCAN_T, 0, st>>>(sizes_dev, n, bsum, offsets_dev);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}
