// Fused SegmentSumMLP row encoder + per-program segment sum on tcgen05 tensor cores.
//
//   S[p, :] = sum_{r in program p} relu(W1 relu(W0 x_r + b0) + b1)
//   (SegmentSumMLPModule.forward, python/tvm/auto_scheduler/cost_model/mlp_model.py:189-208;
//    SegmentVAE.segment_sum, scripts/vae_experiments/models/vae.py:68-85)
//
// One persistent CTA pair (cluster of 2, tcgen05 cta_group::2) per two SMs; a pair tile is 256
// feature rows (128 per CTA).  fp16 operands (11-bit significand), fp32 accumulation in TMEM.
//
//   layer 0:  D0[256 rows x 256 hid] = [X | 1][256 x K0] * [W0 | b0]^T   (two N=128 halves)
//             TMEM cols 0..255           A = X tile (this CTA's 128 rows), B = this CTA's 128 hidden units
//   epi 0  :  TMEM -> ReLU + fp16 pack (one F2FP per pair) -> A1 smem (K-major, 128B swizzle); lane = row
//   layer 1:  D1[256 hid x 256 rows] = W1[256 x 256] * A1^T              (two K halves, as epi 0 delivers them)
//             TMEM cols 256..511         A = W1 (this CTA's 128 hidden units), B = A1 (this CTA's 128 rows)
//   epi 1  :  TMEM -> +b1, ReLU, sum over *columns* (= rows of the tile) between program
//             boundaries -> S.  lane = hidden unit: no shuffles, no shared memory, boundaries are
//             uniform across the CTA; each segment is fetched with power-of-two TMEM loads.
//
// Layer 1 is computed transposed on purpose: it puts the segment dimension on TMEM columns.
// Weights stay resident in shared memory (each CTA holds half of W0 and of W1: 112 KB); X is
// streamed with 128-bit loads behind an L2 bulk prefetch, cast to fp16 and written straight into
// the swizzled operand layout (the cast rules out a TMA tensor load of X).  b0 rides in a spare
// K column of layer 0 (X gets a constant-1 column), so epi 0 has no bias add.
//
// Warp roles (21 warps): 0-7 loaders (16 rows each; the first half of the next tile is already
// in registers when the tensor pipe releases A0), 8-11 epi 0 (one warp per TMEM lane quarter;
// hidden half 0 while the MMAs of half 1 run), 12-19 epi 1 in two groups of four (group g drains
// layer-1 block g while the tensor pipe fills the other block and runs the next tile's layer 0),
// 20 MMA issuer (leader CTA).  All hand-offs are mbarriers; tcgen05.commit signals MMA completion.
//
// epi 1 walks the 64 accumulator columns of a sub-block with a *static* register index and a
// warp-uniform 64-bit "last row of a piece" mask built by a pre-pass: no per-segment TMEM
// round trip.  When F <= 174, b1 rides in the spare k-step (columns 176/177) of the resident
// W0 / A0 images as a 17th K step of layer 1 (hi + lo fp16 halves, so the bias keeps ~22 bits).
//
// Programs cut by a 64-row sub-block boundary receive partial sums through atomicAdd on zeroed
// rows (two addends when S_i <= 64: order-independent, so results are deterministic).
#include <cuda.h>

#include <algorithm>
#include <cstring>
#include <type_traits>
#include <vector>

#include "common.cuh"
#include "model.cuh"
#include "tc_ptx.cuh"

namespace vaex {

namespace tc {
constexpr int HID = 256;
constexpr int TILE = 128;            // rows per CTA per tile (= rows per epi-1 block)
constexpr int PAIR = 256;            // rows per CTA pair per tile
constexpr int K0P = 192;             // layer-0 K padded to 3 swizzle atoms of 64 fp16
constexpr int ATOM = 128 * 128;      // bytes of one [128 rows x 64 fp16] swizzle atom
constexpr int W0_OFF = 0, W0_BYTES = 3 * ATOM;
constexpr int W1_OFF = W0_OFF + W0_BYTES, W1_BYTES = 4 * ATOM;
constexpr int A0_OFF = W1_OFF + W1_BYTES, A0_BYTES = 3 * ATOM;
constexpr int A1_OFF = A0_OFF + A0_BYTES, A1_BYTES = 4 * ATOM;
constexpr int SMEM_TILES = A1_OFF + A1_BYTES;          // 229376
constexpr int SMEM_DYN = SMEM_TILES + 1024;            // + slack to align the base to 1024 B
constexpr int SUB = 64;              // rows per segment-descriptor sub-block (= rows one CTA feeds a layer-1 block)
constexpr int LOAD_WARPS = 8, EPI1_WARP0 = 12, MMA_WARP = 20;   // 0-7 loaders, 8-11 epi 0, 12-19 epi 1, 20 MMA issuer
constexpr int THREADS = 32 * (MMA_WARP + 1);   // 672
// L1_N256: layer 1 as ONE N = 256 MMA per k-step instead of two N = 128 row blocks.  tcgen05.mma M256 N128 K16 issues
// every 93.6 clk whatever N <= 128 is, M256 N256 K16 every 128.1 clk (tools/mma_rate.cu, profiles/r02_mma_rate.txt: 1 494
// vs 2 210 TFLOP/s chip-wide), so the 34 layer-1 MMAs of a tile (3 182 clk) would become 17 (2 178 clk).  Measured: SLOWER
// (3.06 vs 2.76 ms on 16.5 M rows, results identical).  With one 256-column accumulator both row blocks complete together
// and all sixteen epi-1 warps drain at once (~4 K clk: the SM's issue slots, not the tensor pipe, bound the drain) before
// the next tile's layer 1 may start, where the two N = 128 blocks let block 0 drain under block 1's MMAs; TMEM has no
// room for a second 256-column accumulator beside acc0.  Kept as a switch for the record.
constexpr bool L1_N256 = false;
// L0_N256: layer 0 as ONE N = 256 MMA per k-step (11 x 128.1 = 1 409 clk instead of 22 x 93.6 = 2 059; accumulator
// column c then holds hidden unit c, so W1's K order is the identity).  Measured: the image path gets SLOWER (3.06 vs
// 2.76-2.90 ms), the fp32-row path faster (3.69 vs 3.96-4.23 ms): with both hidden halves completing together, epi 0's
// four warps start on half 0 no earlier than before and layer 1 waits for them with an idle pipe, where the two N = 128
// halves keep the pipe busy under epi 0.  Both switches on: 2.95 ms (the 3.6 K clk of MMAs per tile would allow 1.5 x, but
// a block's drain takes 2.5 K clk -- profiles/r02_chain_cycles.md -- and nothing hides it).  Off; kept for the record.
constexpr bool L0_N256 = false;
constexpr int B1_COL = 176;          // k columns of the b1 hi/lo halves (spare k-step 11 of the W0 / A0 images)
enum { B_A0_FULL = 0, B_A0_EMPTY, B_ACC0_FULL0, B_ACC0_FULL1, B_A1_FULL0, B_A1_FULL1, B_ACC1_FULL0, B_ACC1_FULL1,
       B_ACC1_EMPTY0, B_ACC1_EMPTY1, B_COUNT };

__host__ __device__ inline uint32_t sw128_off(int row, int k) {
  return (uint32_t)((k >> 6) * ATOM + (row >> 3) * 1024 + (row & 7) * 128 + ((((k & 63) >> 3) ^ (row & 7)) << 4) +
                    (k & 7) * 2);
}
// hidden unit held by layer-0 accumulator column c (two N=128 halves, 64 units per CTA each)
__host__ __device__ inline int hidden_of_col(int c) {
  if (L0_N256) return c;
  const int h = c >> 7, j = c & 127;
  return j < 64 ? 64 * h + j : 128 + 64 * h + (j - 64);
}
}  // namespace tc

struct TcParams {
  const float* feats;          // [n_rows, F] fp32, or fp16 (already saturated) for the HALF_IN instantiation
  const int32_t* off;          // [P + 1]
  const uint4* meta;           // [4T][2] per 64-row sub-block: {endmask lo, hi, partialmask lo, hi}, {first program, flags, -, -}
  const uint2* desc;           // [4T][64] piece list {program, col | len<<8 | partial<<16}; only written for irregular sub-blocks
  const __half* w0;            // [2][W0_BYTES/2] swizzled images per CTA rank ([W0 | b0])
  const __half* w1;            // [2][W1_BYTES/2] (K permuted to the layer-0 column order)
  const float* b1;             // [256]
  float* out;                  // [P, 256]
  uint32_t* rowmask;           // [T * 8] bit r: feature row has a non-zero
  int64_t n_rows;
  int32_t P, T, F, k0_steps;
};

// ---------------------------------------------------------------------------------------------
// loader: the CTA's 128 x F fp32 half tile is one contiguous range of global memory.  Each of
// the 8 loader warps owns 16 consecutive rows = a contiguous run of 16*U vector units (U = F/VEC)
// and sweeps it with fully coalesced 16/8/4-byte loads, SLOTS units per lane, issued in two
// batches so that ~45 KB per SM are in flight.  A unit is cast to fp16 and stored at its
// swizzled (row, k) position; a ballot per slot yields the per-row "has a non-zero" bits.
template <int VEC>
struct Vec;
template <> struct Vec<4> { float4 v; __device__ void ld(const float* p) { v = ptx::ldg_stream_f4(p); }
  __device__ bool nz() const { return ((__float_as_uint(v.x) | __float_as_uint(v.y) | __float_as_uint(v.z) | __float_as_uint(v.w)) << 1) != 0; }
  __device__ void zero() { v = make_float4(0.f, 0.f, 0.f, 0.f); }
  __device__ void st(uint32_t a) const { ptx::sts_v2(a, ptx::cvt_h2_sat(v.x, v.y), ptx::cvt_h2_sat(v.z, v.w)); } };
template <> struct Vec<2> { float2 v; __device__ void ld(const float* p) { v = ptx::ldg_stream_f2(p); }
  __device__ bool nz() const { return ((__float_as_uint(v.x) | __float_as_uint(v.y)) << 1) != 0; }
  __device__ void zero() { v = make_float2(0.f, 0.f); }
  __device__ void st(uint32_t a) const { ptx::sts_b32(a, ptx::cvt_h2_sat(v.x, v.y)); } };
template <> struct Vec<1> { float v; __device__ void ld(const float* p) { v = __ldg(p); }
  __device__ bool nz() const { return (__float_as_uint(v) << 1) != 0; }
  __device__ void zero() { v = 0.f; }
  __device__ void st(uint32_t a) const { ptx::sts_b16(a, ptx::cvt_h2_sat(v, 0.f)); } };

template <int VEC>
__device__ __forceinline__ uint32_t load_convert_rows(const TcParams& p, uint32_t A0, int64_t row0, int wrow0, int lane) {
  using namespace tc;
  constexpr int ROWS = TILE / LOAD_WARPS;             // 16 rows per warp
  constexpr int MAXU = (K0P - 1) / VEC;               // F <= K0P - 1
  constexpr int SLOTS = (ROWS * MAXU + 31) / 32;
  constexpr int BATCH = VEC == 4 ? 12 : 16;           // slots in flight per lane
  const int U = p.F / VEC;
  const int total = ROWS * U;                         // units owned by this warp
  const int64_t grow0 = row0 + wrow0;                 // first global row of this warp
  const int64_t valid_rows = min((int64_t)ROWS, p.n_rows - grow0);
  const int valid_units = valid_rows > 0 ? (int)valid_rows * U : 0;
  const float* src = p.feats + grow0 * p.F;
  uint32_t nzmask = 0;
  int row = lane / U, col = lane - row * U;           // (row, unit) of this lane's unit in slot 0
  int urow = 0, ucol = 0;                             // (row, unit) of lane 0's unit (warp-uniform)
  for (int s0 = 0; s0 * 32 < total; s0 += BATCH) {
    Vec<VEC> x[BATCH];
#pragma unroll
    for (int j = 0; j < BATCH; ++j) {
      const int idx = (s0 + j) * 32 + lane;
      if (j < SLOTS && idx < valid_units) x[j].ld(src + (size_t)idx * VEC); else x[j].zero();
    }
#pragma unroll
    for (int j = 0; j < BATCH; ++j) {
      const int base = (s0 + j) * 32;
      if (j < SLOTS && base < total) {
        if (base + lane < total) x[j].st(A0 + sw128_off(wrow0 + row, col * VEC));
        const uint32_t bal = __ballot_sync(0xffffffffu, x[j].nz());
        // rows touched by units [base, base+32): uniform scalar walk (1-2 rows when U >= 32)
        int r = urow, lo = 0, rem = U - ucol;
        while (lo < 32 && r < ROWS) {
          const int hi = min(32, lo + rem);
          const uint32_t m = (hi >= 32 ? 0xffffffffu : ((1u << hi) - 1u)) & ~((1u << lo) - 1u);
          if (bal & m) nzmask |= 1u << r;
          lo = hi;
          rem = U;
          ++r;
        }
        col += 32;
        while (col >= U) { col -= U; ++row; }
        ucol += 32;
        while (ucol >= U) { ucol -= U; ++urow; }
      }
    }
  }
  return nzmask;
}

// Fast path (F % 4 == 0, 16-byte aligned rows): the swizzled A0 offset of every (slot, lane) unit
// is the same for all tiles and, up to +2048*warp, for all warps -> a 1.5 KB table in shared
// memory built once per CTA.  Per slot and lane: LDG.128, LDS.U16, 2 F2FP, STS.64 and the row's
// non-zero bit (row = bits 7..10 of the offset); one REDUX.OR per tile merges the lanes' bits.
// The 24 slots of a lane are split in two batches: issue() puts 12 LDG.128 in flight, commit()
// converts and stores them, so batch 0 of the next tile can be fetched before A0 is released.
constexpr int FAST_SLOTS = 24;                        // 16 rows * 47 units / 32 lanes
constexpr int FAST_BATCH = 12;
struct FastLoader {
  static constexpr int BATCH = FAST_BATCH;
  using Unit = float4;
  const float4* src;
  int valid_units, total;
  __device__ __forceinline__ void open(const TcParams& p, int64_t grow0, int lane) {
    constexpr int ROWS = tc::TILE / tc::LOAD_WARPS;
    total = ROWS * (p.F >> 2);
    const int64_t vr = min((int64_t)ROWS, p.n_rows - grow0);
    valid_units = vr > 0 ? (int)vr * (p.F >> 2) : 0;
    src = reinterpret_cast<const float4*>(p.feats + (vr > 0 ? grow0 : 0) * p.F) + lane;
  }
  __device__ __forceinline__ void issue(int s0, int lane, float4 (&x)[FAST_BATCH]) const {
#pragma unroll
    for (int j = 0; j < FAST_BATCH; ++j) {
      const int idx = (s0 + j) * 32 + lane;
      if (idx < valid_units) x[j] = ptx::ldg_stream_f4(reinterpret_cast<const float*>(src + (s0 + j) * 32));
      else x[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  __device__ __forceinline__ uint32_t commit(int s0, int lane, const float4 (&x)[FAST_BATCH], uint32_t A0w,
                                             const uint16_t* tab) const {
    uint32_t mask = 0;
#pragma unroll
    for (int j = 0; j < FAST_BATCH; ++j) {
      const int idx = (s0 + j) * 32 + lane;
      if (idx < total) {
        const uint32_t o = tab[idx];
        ptx::sts_v2(A0w + o, ptx::cvt_h2_sat(x[j].x, x[j].y), ptx::cvt_h2_sat(x[j].z, x[j].w));
        const uint32_t any = (__float_as_uint(x[j].x) | __float_as_uint(x[j].y) | __float_as_uint(x[j].z) | __float_as_uint(x[j].w)) << 1;
        if (any) mask |= 1u << ((o >> 7) & 15);
      }
    }
    return mask;
  }
};

// Same walk for features that are already fp16 in HBM (VAEX_F_FEATS_F16: the host path narrows them before the
// PCIe copy).  A lane moves 16 bytes = 8 halves per slot with one LDG.128 - two consecutive 4-feature units of the
// table above, so a slot that straddles two rows (F % 8 == 4) simply stores its halves at two table offsets.  A
// warp's 16-row run is a whole number of 16-byte units (16 * F * 2 bytes); only the ragged last tile can end on a
// half unit, which is fetched with one LDG.64.  11 slots per lane (F = 164) instead of 21, no F2FP, and as many
// bytes in flight per load as the fp32 walk (LDG.64 per 4-feature unit measured SLOWER than fp32 rows: 4.28 vs
// 4.05 ms per 10^6 programs - the loaders are bound by bytes in flight, not by their instruction count).
struct FastLoaderH {
  static constexpr int BATCH = 6;
  using Unit = uint4;
  const uint4* src;
  int valid_units, total;                              // in 8-byte (4-feature) units, like the table
  __device__ __forceinline__ void open(const TcParams& p, int64_t grow0, int lane) {
    constexpr int ROWS = tc::TILE / tc::LOAD_WARPS;
    total = ROWS * (p.F >> 2);
    const int64_t vr = min((int64_t)ROWS, p.n_rows - grow0);
    valid_units = vr > 0 ? (int)vr * (p.F >> 2) : 0;
    src = reinterpret_cast<const uint4*>(reinterpret_cast<const __half*>(p.feats) + (vr > 0 ? grow0 : 0) * p.F) + lane;
  }
  __device__ __forceinline__ void issue(int s0, int lane, uint4 (&x)[BATCH]) const {
#pragma unroll
    for (int j = 0; j < BATCH; ++j) {
      const int u = 2 * ((s0 + j) * 32 + lane);        // first 8-byte unit of this slot
      if (u + 1 < valid_units) x[j] = ptx::ldg_stream_u4(src + (s0 + j) * 32);
      else if (u < valid_units) {
        const uint2 h = ptx::ldg_stream_u2(src + (s0 + j) * 32);
        x[j] = make_uint4(h.x, h.y, 0u, 0u);
      } else x[j] = make_uint4(0u, 0u, 0u, 0u);
    }
  }
  __device__ __forceinline__ uint32_t commit(int s0, int lane, const uint4 (&x)[BATCH], uint32_t A0w,
                                             const uint16_t* tab) const {
    uint32_t mask = 0;
#pragma unroll
    for (int j = 0; j < BATCH; ++j) {
      const int u = 2 * ((s0 + j) * 32 + lane);
      if (u < total) {                                  // total is even: both halves of the slot are in the tile
        const uint32_t o = *reinterpret_cast<const uint32_t*>(tab + u);
        const uint32_t o0 = o & 0xffffu, o1 = o >> 16;
        ptx::sts_v2(A0w + o0, x[j].x, x[j].y);
        ptx::sts_v2(A0w + o1, x[j].z, x[j].w);
        if ((x[j].x | x[j].y) & 0x7fff7fffu) mask |= 1u << ((o0 >> 7) & 15);
        if ((x[j].z | x[j].w) & 0x7fff7fffu) mask |= 1u << ((o1 >> 7) & 15);
      }
    }
    return mask;
  }
};

// ---------------------------------------------------------------------------------------------
// epi 1 building block: 16 accumulator columns (static register index).  m: bit j = column j is
// the last row of a piece, pm: that piece is partial (cut by the sub-block boundary -> atomicAdd).
// a0/a1 are the running even/odd-column sums of the open piece; dst walks the output rows (the
// pieces of a regular sub-block belong to consecutive programs).
__device__ __forceinline__ void flush_piece(float* dst, float s, uint32_t partial) {
  // one pointer, two predicated instructions: keeps the (cold) boundary code small
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0;\n\t@p red.global.add.f32 [%0], %1;\n\t@!p st.global.f32 [%0], %1;\n\t}"
      :: "l"(dst), "f"(s), "r"(partial) : "memory");
}
// Tried and measured (profiles/r01_ncu_segsum_tc_v8_notes.md): four accumulator chains instead of two (no change: the
// `wait` stalls are not the FADD chains), __builtin_expect / goto-ordered cold blocks (ptxas lays the boundary block out as
// the fall-through of its branch regardless of source or PTX order), the boundary path as a __noinline__ call (compact hot
// loop, but 4.08 vs 3.85 ms: call stubs + spills).
template <bool FOLD>
__device__ __forceinline__ void sum_chunk(const uint32_t (&v)[16], uint32_t m, uint32_t pm, float b1, float& a0, float& a1,
                                          float*& dst) {
  auto act = [&](uint32_t u) { return FOLD ? fmaxf(__uint_as_float(u), 0.f) : fmaxf(__uint_as_float(u) + b1, 0.f); };
#pragma unroll
  for (int j = 0; j < 16; j += 4) {
    const float x0 = act(v[j]), x1 = act(v[j + 1]), x2 = act(v[j + 2]), x3 = act(v[j + 3]);
    if ((m >> j) & 15u) {                                 // warp-uniform, ~22 % of the 4-column groups
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float x = i == 0 ? x0 : i == 1 ? x1 : i == 2 ? x2 : x3;
        if (i & 1) a1 += x; else a0 += x;
        if ((m >> (j + i)) & 1u) {
          flush_piece(dst, a0 + a1, (pm >> (j + i)) & 1u);
          dst += tc::HID;
          a0 = 0.f;
          a1 = 0.f;
        }
      }
    } else {
      a0 += x0;
      a1 += x1;
      a0 += x2;
      a1 += x3;
    }
  }
}

// irregular sub-block (an empty program sits between two pieces; rare): per-piece TMEM walk,
// kept small on purpose (8-column loads, then single columns)
__device__ __forceinline__ void sum_irregular(uint32_t tcol, const uint2* __restrict__ d, int nvalid, float b1, float* out_h) {
  int covered = 0, k = 0;
  while (covered < nvalid) {
    const uint2 e = __ldg(d + k);
    ++k;
    uint32_t c = tcol + (e.y & 0xff);
    int n = (e.y >> 8) & 0xff;
    covered += n;
    float s = 0.f;
    for (; n >= 8; n -= 8, c += 8) {
      uint32_t v[8];
      ptx::tmem_ld8(c, v);
      ptx::tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 8; ++j) s += fmaxf(__uint_as_float(v[j]) + b1, 0.f);
    }
    for (; n > 0; --n, ++c) {
      uint32_t v;
      ptx::tmem_ld1(c, v);
      ptx::tmem_ld_wait();
      s += fmaxf(__uint_as_float(v) + b1, 0.f);
    }
    float* dst = out_h + (size_t)e.x * tc::HID;
    if (e.y >> 16) atomicAdd(dst, s); else *dst = s;
  }
}

// ---------------------------------------------------------------------------------------------
// IN: how the feature rows reach the A0 operand.
//   IN_F32   fp32 rows [n_rows, F]: loader warps, LDG.128 -> cvt -> swizzled st.shared
//   IN_F16   fp16 rows [n_rows, F] (the host path's narrowed chunks): loader warps, LDG.128 -> st.shared
//   IN_IMAGE the packed fp16 row image [n_rows, 192] (vaex_rows_pack_image: bias-one columns pre-written, 384-byte
//            pitch): ONE thread per CTA issues three TMA tensor loads per tile (cp.async.bulk.tensor.2d, 128-byte
//            swizzle = the UMMA operand layout) that complete on the leader CTA's mbarrier; no loader warps, no
//            format-conversion instructions, 384 instead of 656 bytes of HBM per row.
enum { IN_F32 = 0, IN_F16 = 1, IN_IMAGE = 2 };
// -DSEGSUM_PROF: clock64 accounting of cluster 0 in shared memory (issuer waits per barrier, epi 0 / epi 1 waits and work of
// their polling warps), dumped to seg_prof at exit; read with vaex_debug_segsum_prof (tools/prof_segsum.py)
#ifdef SEGSUM_PROF
__device__ unsigned long long seg_prof[32];
#define SGP(i, v) do { if (blockIdx.x == 0) sprof[i] += (unsigned long long)(v); } while (0)
#define SGT() clock64()
#else
#define SGP(i, v) do { (void)(v); } while (0)
#define SGT() 0ll
#endif
template <bool FOLD, int IN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(tc::THREADS, 1)
segsum_tc_kernel(const TcParams p, const __grid_constant__ CUtensorMap tmap) {
  using namespace tc;
  constexpr bool HALF_IN = IN == IN_F16;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[B_COUNT];
  __shared__ uint32_t tmem_base_slot;
  __shared__ __align__(16) uint16_t a0_tab[FAST_SLOTS * 32];
#ifdef SEGSUM_PROF
  __shared__ unsigned long long sprof[32];
  if (threadIdx.x < 32) sprof[threadIdx.x] = 0;
#endif

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t sbase = ptx::smem_u32(smem);
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);   // warp-uniform for the compiler
  const int lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const int cluster = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const uint32_t bar0 = ptx::smem_u32(&bars[0]);
  auto bar = [&](int i) { return bar0 + 8u * i; };

  // ---- prologue -------------------------------------------------------------------------------
  if (warp == MMA_WARP && lane == 0) {
    ptx::mbar_init(bar(B_A0_FULL), IN == IN_IMAGE ? 1 : 2 * LOAD_WARPS);   // image: the leader's arrive.expect_tx
    ptx::mbar_init(bar(B_A0_EMPTY), 1);
    ptx::mbar_init(bar(B_ACC0_FULL0), 1);
    ptx::mbar_init(bar(B_ACC0_FULL1), 1);
    ptx::mbar_init(bar(B_A1_FULL0), 2 * 4);
    ptx::mbar_init(bar(B_A1_FULL1), 2 * 4);
    ptx::mbar_init(bar(B_ACC1_FULL0), 1);
    ptx::mbar_init(bar(B_ACC1_FULL1), 1);
    ptx::mbar_init(bar(B_ACC1_EMPTY0), IN == IN_IMAGE ? 2 * 8 : 2 * 4);    // epi-1 warps of the block, both CTAs
    ptx::mbar_init(bar(B_ACC1_EMPTY1), IN == IN_IMAGE ? 2 * 8 : 2 * 4);
    ptx::fence_mbar_init();
  }
  if (warp == EPI1_WARP0) ptx::tmem_alloc_2cta(ptx::smem_u32(&tmem_base_slot), 512);
  {
    // resident weights: already in the exact swizzled byte layout -> straight 16-byte copies
    const uint4* g0 = reinterpret_cast<const uint4*>(p.w0) + (size_t)rank * (W0_BYTES / 16);
    const uint4* g1 = reinterpret_cast<const uint4*>(p.w1) + (size_t)rank * (W1_BYTES / 16);
    uint4* s0 = reinterpret_cast<uint4*>(smem + W0_OFF);
    uint4* s1 = reinterpret_cast<uint4*>(smem + W1_OFF);
    for (int i = threadIdx.x; i < W0_BYTES / 16; i += THREADS) s0[i] = __ldg(g0 + i);
    for (int i = threadIdx.x; i < W1_BYTES / 16; i += THREADS) s1[i] = __ldg(g1 + i);
    if ((p.F & 3) == 0) {                                   // loader fast-path offset table
      const int U = p.F >> 2;
      for (int i = threadIdx.x; i < FAST_SLOTS * 32; i += THREADS) {
        const int row = i / U, col = i - row * U;
        a0_tab[i] = row < TILE / LOAD_WARPS ? (uint16_t)sw128_off(row, col * 4) : (uint16_t)0;
      }
    }
    uint4* a0 = reinterpret_cast<uint4*>(smem + A0_OFF);   // zero once: K padding columns stay 0
    for (int i = threadIdx.x; i < A0_BYTES / 16; i += THREADS) a0[i] = make_uint4(0, 0, 0, 0);
  }
  __syncthreads();
  // constant-1 columns: k = F, F+1 multiply the b0 hi/lo columns of the W0 image (bias inside the
  // MMA); k = 176, 177 multiply the b1 hi/lo columns in the extra K step of layer 1
  if (threadIdx.x < TILE) {
    auto one = [&](int k) { *reinterpret_cast<uint16_t*>(smem + A0_OFF + sw128_off(threadIdx.x, k)) = 0x3C00; };
    one(p.F);
    if (p.F + 1 < K0P) one(p.F + 1);
    if (FOLD) { one(B1_COL); one(B1_COL + 1); }
  }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  ptx::tc_fence_after();
  const uint32_t tmem = tmem_base_slot;

  // ---- roles ----------------------------------------------------------------------------------
  if (IN != IN_IMAGE && warp < LOAD_WARPS) {
    // ===== loaders: fp32 rows -> fp16 swizzled A0; 16 rows per warp =====
    constexpr int ROWS = TILE / LOAD_WARPS;
    const uint32_t A0 = sbase + A0_OFF;
    // HALF_IN: the host checked F % 4 == 0 and 16-byte alignment (a warp's 16-row run then starts 16-byte aligned)
    const bool al16 = HALF_IN || (((reinterpret_cast<uintptr_t>(p.feats) & 15) == 0) && (p.F % 4 == 0));
    const bool al8 = ((reinterpret_cast<uintptr_t>(p.feats) & 7) == 0) && (p.F % 2 == 0);
    auto prefetch = [&](int t) {
      if (!al16 || t >= p.T) return;
      const int64_t r0 = (int64_t)t * PAIR + rank * TILE + warp * ROWS;
      const int64_t nr = min((int64_t)ROWS, p.n_rows - r0);
      if (nr <= 0) return;
      if (HALF_IN) ptx::prefetch_l2_bulk(reinterpret_cast<const __half*>(p.feats) + r0 * p.F, (uint32_t)(nr * p.F * 2) & ~15u);
      else ptx::prefetch_l2_bulk(p.feats + r0 * p.F, (uint32_t)(nr * p.F * 4));
    };
    auto publish = [&](int t, uint32_t nzmask) {
      // 16 flag bits per warp: two warps share one 32-bit word of the row mask
      if (lane == 0) reinterpret_cast<uint16_t*>(p.rowmask)[((size_t)t * 8 + rank * 4) * 2 + warp] = (uint16_t)nzmask;
      ptx::fence_proxy_async();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster_relaxed(bar(B_A0_FULL), 0);
    };
    if (lane == 0) { prefetch(cluster + n_clusters); prefetch(cluster + 2 * n_clusters); }
    if (al16) {
      using Loader = typename std::conditional<HALF_IN, FastLoaderH, FastLoader>::type;
      Loader ld;
      typename Loader::Unit x[Loader::BATCH];
      const uint32_t A0w = A0 + 2048u * warp;
      if (cluster < p.T) {
        ld.open(p, (int64_t)cluster * PAIR + rank * TILE + warp * ROWS, lane);
        ld.issue(0, lane, x);
      }
      int it = 0;
      for (int t = cluster; t < p.T; t += n_clusters, ++it) {
        if (lane == 0) prefetch(t + 3 * n_clusters);
        if (it > 0) ptx::mbar_wait_role(bar(B_A0_EMPTY), (it - 1) & 1, warp == 0, 1, 32 * LOAD_WARPS);   // layer 0 of the previous tile has read A0
        uint32_t nz = ld.commit(0, lane, x, A0w, a0_tab);
        ld.issue(Loader::BATCH, lane, x);
        nz |= ld.commit(Loader::BATCH, lane, x, A0w, a0_tab);
        publish(t, __reduce_or_sync(0xffffffffu, nz));
        if (t + n_clusters < p.T) {                                     // batch 0 of the next tile rides in registers
          ld.open(p, (int64_t)(t + n_clusters) * PAIR + rank * TILE + warp * ROWS, lane);
          ld.issue(0, lane, x);
        }
      }
    } else {
      int it = 0;
      for (int t = cluster; t < p.T; t += n_clusters, ++it) {
        if (it > 0) ptx::mbar_wait_role(bar(B_A0_EMPTY), (it - 1) & 1, warp == 0, 1, 32 * LOAD_WARPS);
        const int64_t row0 = (int64_t)t * PAIR + rank * TILE;
        publish(t, al8 ? load_convert_rows<2>(p, A0, row0, warp * ROWS, lane) : load_convert_rows<1>(p, A0, row0, warp * ROWS, lane));
      }
    }
  } else if (warp >= LOAD_WARPS && warp < EPI1_WARP0) {
    // ===== epi 0: acc0 -> ReLU -> fp16 -> A1 (K-major, 128B swizzle); lane = tile row =====
    const int q = warp & 3;
    const uint32_t lane_base = (uint32_t)(32 * q) << 16;
    const int row = 32 * q + lane;
    const uint32_t a1_row = sbase + A1_OFF + (row >> 3) * 1024 + (row & 7) * 128;
    const int sw = row & 7;
    auto pack32 = [&](const uint32_t (&v)[32], int c) {   // accumulator columns [32c, 32c+32) -> 64 B of A1
      const uint32_t base = a1_row + (c >> 1) * ATOM;
#pragma unroll
      for (int m = 0; m < 4; ++m)
        ptx::sts_v4(base + ((((c & 1) * 4 + m) ^ sw) << 4),
                    ptx::cvt_h2_relu_sat(__uint_as_float(v[8 * m]), __uint_as_float(v[8 * m + 1])),
                    ptx::cvt_h2_relu_sat(__uint_as_float(v[8 * m + 2]), __uint_as_float(v[8 * m + 3])),
                    ptx::cvt_h2_relu_sat(__uint_as_float(v[8 * m + 4]), __uint_as_float(v[8 * m + 5])),
                    ptx::cvt_h2_relu_sat(__uint_as_float(v[8 * m + 6]), __uint_as_float(v[8 * m + 7])));
    };
    int it = 0;
    for (int t = cluster; t < p.T; t += n_clusters, ++it) {
#pragma unroll 1
      for (int g = 0; g < 2; ++g) {                        // hidden half g = acc0 columns [128g, 128g+128)
        const long long p0 = SGT();
        if (q == 0) {
          ptx::mbar_wait(bar(B_ACC0_FULL0 + g), it & 1);
          if (g == 0 && it > 0) ptx::mbar_wait(bar(B_ACC1_FULL1), (it - 1) & 1);   // layer 1 of the previous tile has read A1
        }
        ptx::bar_sync(2, 128);
        ptx::tc_fence_after();
        const long long p1 = SGT();
        if (warp == LOAD_WARPS && lane == 0) SGP(8 + 2 * g, p1 - p0);
        const uint32_t tcol = tmem + lane_base + 128 * g;
        uint32_t va[32], vb[32];
        ptx::tmem_ld32(tcol, va);
        ptx::tmem_ld_wait();
        ptx::tmem_ld32(tcol + 32, vb);
        pack32(va, 4 * g);
        ptx::tmem_ld_wait();
        ptx::tmem_ld32(tcol + 64, va);
        pack32(vb, 4 * g + 1);
        ptx::tmem_ld_wait();
        ptx::tmem_ld32(tcol + 96, vb);
        pack32(va, 4 * g + 2);
        ptx::tmem_ld_wait();
        pack32(vb, 4 * g + 3);
        ptx::fence_proxy_async();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive_cluster_relaxed(bar(B_A1_FULL0 + g), 0);
        if (warp == LOAD_WARPS && lane == 0) SGP(9 + 2 * g, SGT() - p1);
      }
    }
  } else if (warp < MMA_WARP) {
    // ===== epi 1: layer-1 block g = TMEM columns [256+128g, +128) =====
    //   columns [0,64)   <-> rows [64g, 64g+64) of CTA 0  = sub-block 4t + g
    //   columns [64,128) <-> rows [64g, 64g+64) of CTA 1  = sub-block 4t + 2 + g
    // Two groups of four warps (one per TMEM lane quarter); group g drains block g while the
    // tensor pipe fills the other one.  One loop body per 32 columns (two instantiations of the
    // statically indexed 16-column code, ping-pong TMEM loads), four steps per block.
    // IN_IMAGE: the eight former loader warps drain too -- sixteen warps, four per (block g, sub-block h): a warp
    // walks 64 columns instead of 128, which halves the drain latency that bounds the tile period.
    const int q = warp & 3;
    const int e1 = warp < LOAD_WARPS ? warp : warp - (EPI1_WARP0 - LOAD_WARPS);   // image: 0..15; else 8..15
    const int g = IN == IN_IMAGE ? (e1 >> 2) & 1 : (warp - EPI1_WARP0) >> 2;
    const int s_begin = IN == IN_IMAGE ? 2 * (e1 >> 3) : 0, s_end = IN == IN_IMAGE ? s_begin + 2 : 4;
    constexpr int GROUP_THREADS = IN == IN_IMAGE ? 256 : 128;
    const bool poller = IN == IN_IMAGE ? (e1 & 11) == 0 : q == 0;                 // one polling warp per block
    const uint32_t lane_base = (uint32_t)(32 * q) << 16;
    const int hu = rank * TILE + 32 * q + lane;   // hidden unit of this lane
    const float b1 = FOLD ? 0.f : p.b1[hu];
    float* const out_h = p.out + hu;
    // accumulator columns of (block g, 32-column step s).  Two N = 128 row blocks: block g = columns [128 g, +128), CTA h's
    // rows at [64 h, +64) of it.  One N = 256 MMA: column = 128 * CTA + row, so CTA h's rows [64 g, +64) are at 128 h + 64 g.
    auto tcol_of = [&](int s) -> uint32_t {
      return tmem + lane_base + 256 + (L1_N256 ? 128 * (s >> 1) + 64 * g + 32 * (s & 1) : 128 * g + 32 * s);
    };
    uint32_t va[16], vb[16];
    // sub-block metadata of this tile (na, nb: h = 0; nc, nd: h = 1), fetched one tile ahead
    uint4 na = make_uint4(0, 0, 0, 0), nc = na;
    uint2 nb = make_uint2(0, 0), nd = nb;
    auto fetch_meta = [&](int t) {
      if (t >= p.T) return;
      const uint4* m = p.meta + 2 * (size_t)(4 * t + g);
      na = __ldg(m);
      nb = __ldg(reinterpret_cast<const uint2*>(m + 1));
      nc = __ldg(m + 4);
      nd = __ldg(reinterpret_cast<const uint2*>(m + 5));
    };
    fetch_meta(cluster);
    int it = 0;
    for (int t = cluster; t < p.T; t += n_clusters, ++it) {
      const uint4 ma = na, mc = nc;
      const uint2 mb = nb, md = nd;
      fetch_meta(t + n_clusters);
      const long long q0 = SGT();
      ptx::mbar_wait_role(bar(B_ACC1_FULL0 + g), it & 1, poller, 3 + g, GROUP_THREADS);
      ptx::tc_fence_after();
      const long long q1 = SGT();
      if (poller && lane == 0) SGP(12 + 2 * g, q1 - q0);
      ptx::tmem_ld16(tcol_of(s_begin), va);
      float* dst = out_h;
      float a0 = 0.f, a1 = 0.f;
      uint32_t em = 0, pm = 0;
      bool regular = true;
#pragma unroll 1
      for (int s = s_begin; s < s_end; ++s) {              // columns [32s, 32s+32) of the block
        if ((s & 1) == 0) {                                // a new sub-block: 4t + 2h + g, h = s >> 1
          const bool h = s != 0;
          regular = (h ? md.y : mb.y) == 0u;
          dst = out_h + (size_t)(h ? md.x : mb.x) * HID;
          em = h ? mc.x : ma.x;
          pm = h ? mc.z : ma.z;
          a0 = 0.f;
          a1 = 0.f;
        } else {
          const bool h = s == 3;
          em = h ? mc.y : ma.y;
          pm = h ? mc.w : ma.w;
        }
        const uint32_t tcol = tcol_of(s);
        ptx::tmem_ld_wait();
        ptx::tmem_ld16(tcol + 16, vb);
        if (__builtin_expect(regular, 1)) {
          sum_chunk<FOLD>(va, em, pm, b1, a0, a1, dst);
        } else if ((s & 1) == 0) {
          const int sub = 4 * t + (s & 2) + g;
          sum_irregular(tcol, p.desc + (size_t)sub * SUB, min(SUB, (int)p.n_rows - sub * SUB), b1, out_h);
        }
        ptx::tmem_ld_wait();
        if (s + 1 != s_end) ptx::tmem_ld16(tcol_of(s + 1), va);
        if (__builtin_expect(regular, 1)) sum_chunk<FOLD>(vb, em >> 16, pm >> 16, b1, a0, a1, dst);
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster_relaxed(bar(B_ACC1_EMPTY0 + g), 0);
      if (poller && lane == 0) SGP(13 + 2 * g, SGT() - q1);
    }
  } else {
    // ===== MMA issuer: one thread of the leader CTA =====
    // IN_IMAGE: the same thread of each CTA is the TMA producer of its half tile: three [64 halves x 128 rows] boxes =
    // the three swizzle atoms of A0, completing on the LEADER's A0_FULL barrier (cta_group::2).  The leader issues its
    // loads inline, at the point of its sequence where layer 0 of the current tile is known to be complete; the peer
    // CTA's warp 20 has nothing else to do and simply waits for A0_EMPTY.
    const __half* img = reinterpret_cast<const __half*>(p.feats);
    auto tma_tile = [&](int t) {                         // one thread
      const uint32_t full_leader = ptx::mapa_rank(bar(B_A0_FULL), 0);
      const int row0 = t * PAIR + (int)rank * TILE;
#pragma unroll
      for (int a = 0; a < 3; ++a) ptx::tma_load_2d_2cta(sbase + A0_OFF + a * ATOM, &tmap, full_leader, 64 * a, row0);
      if (rank == 0) ptx::mbar_arrive_expect_tx(bar(B_A0_FULL), 2 * A0_BYTES);   // both CTAs' boxes land on this barrier
      const int t2 = t + 2 * n_clusters;                 // the CTA's half tile is one contiguous 48 KB range of the image
      if (t2 < p.T) ptx::prefetch_l2_bulk(img + ((int64_t)t2 * PAIR + rank * TILE) * K0P, (uint32_t)(TILE * K0P * 2));
    };
    if (IN == IN_IMAGE && rank != 0) {
      if (ptx::elect_one()) {
        ptx::tma_prefetch_desc(&tmap);
        int it = 0;
        for (int t = cluster; t < p.T; t += n_clusters, ++it) {
          if (it > 0) ptx::mbar_wait(bar(B_A0_EMPTY), (it - 1) & 1);    // layer 0 of the previous tile has read A0
          tma_tile(t);
        }
      }
    }
    if (rank == 0) {
      if (IN == IN_IMAGE && cluster < p.T && ptx::elect_one()) {
        ptx::tma_prefetch_desc(&tmap);
        const int t1 = cluster + n_clusters;
        if (t1 < p.T) ptx::prefetch_l2_bulk(img + ((int64_t)t1 * PAIR + rank * TILE) * K0P, (uint32_t)(TILE * K0P * 2));
        tma_tile(cluster);
      }
      __syncwarp();
      const uint32_t idesc_n128 = ptx::umma_idesc_f16(256, 128), idesc_n256 = ptx::umma_idesc_f16(256, 256);
      const uint64_t dW0 = ptx::umma_desc_sw128(sbase + W0_OFF), dW1 = ptx::umma_desc_sw128(sbase + W1_OFF);
      const uint64_t dA0 = ptx::umma_desc_sw128(sbase + A0_OFF), dA1 = ptx::umma_desc_sw128(sbase + A1_OFF);
      // +32 B per K=16 step inside an atom (encoded >>4 -> +2), +ATOM per 64 columns
      auto kadv = [](int k) -> uint64_t { return (uint64_t)(((k >> 2) * ATOM + (k & 3) * 32) >> 4); };
      constexpr uint64_t rows64 = (64 * 128) >> 4;      // operand rows 64.. of a CTA (N = 128 halves)
      constexpr int KB1 = B1_COL / 16;                  // the b1 k-step of the W0 / A0 images
      int it = 0;
      for (int t = cluster; t < p.T; t += n_clusters, ++it) {
        // ---- layer 0: two hidden halves ----
        const long long i0 = SGT();
        if (lane == 0) { SGP(0, 1); if (it == 1) SGP(7, 0); }
        ptx::mbar_wait(bar(B_A0_FULL), it & 1);
        ptx::tc_fence_after();
        const long long i1 = SGT();
        if (lane == 0) SGP(1, i1 - i0);
        if (ptx::elect_one()) {
          if (L0_N256) {
            for (int k = 0; k < p.k0_steps; ++k)
              ptx::umma_f16_2cta(tmem + 0, dA0 + kadv(k), dW0 + kadv(k), idesc_n256, k > 0);
            ptx::umma_commit_2cta(bar(B_ACC0_FULL0), 3);
          } else {
            for (int k = 0; k < p.k0_steps; ++k)
              ptx::umma_f16_2cta(tmem + 0, dA0 + kadv(k), dW0 + kadv(k), idesc_n128, k > 0);
            ptx::umma_commit_2cta(bar(B_ACC0_FULL0), 3);
            for (int k = 0; k < p.k0_steps; ++k)
              ptx::umma_f16_2cta(tmem + 128, dA0 + kadv(k), dW0 + rows64 + kadv(k), idesc_n128, k > 0);
          }
          ptx::umma_commit_2cta(bar(B_ACC0_FULL1), 3);
          ptx::umma_commit_2cta(bar(B_A0_EMPTY), 3);
        }
        __syncwarp();
        if (L1_N256) {
          // ---- layer 1, all 256 rows per MMA: K half 0 as soon as epi 0 has delivered it, then K half 1 ----
          if (it > 0) {
            ptx::mbar_wait(bar(B_ACC1_EMPTY0), (it - 1) & 1);
            ptx::mbar_wait(bar(B_ACC1_EMPTY1), (it - 1) & 1);
          }
          ptx::mbar_wait(bar(B_A1_FULL0), it & 1);
          ptx::tc_fence_after();
          if (ptx::elect_one()) {
            if (FOLD) ptx::umma_f16_2cta(tmem + 256, dW0 + kadv(KB1), dA0 + kadv(KB1), idesc_n256, 0);
#pragma unroll
            for (int k = 0; k < 8; ++k)
              ptx::umma_f16_2cta(tmem + 256, dW1 + kadv(k), dA1 + kadv(k), idesc_n256, FOLD || k > 0);
          }
          __syncwarp();
          ptx::mbar_wait(bar(B_A1_FULL1), it & 1);
          ptx::tc_fence_after();
          if (IN == IN_IMAGE && t + n_clusters < p.T) {
            // epi 0 has drained both halves of acc0, so layer 0 is complete and A0_EMPTY has fired: this wait falls through
            ptx::mbar_wait(bar(B_A0_EMPTY), it & 1);
            if (ptx::elect_one()) tma_tile(t + n_clusters);
            __syncwarp();
          }
          if (ptx::elect_one()) {
#pragma unroll
            for (int k = 8; k < 16; ++k) ptx::umma_f16_2cta(tmem + 256, dW1 + kadv(k), dA1 + kadv(k), idesc_n256, 1);
            ptx::umma_commit_2cta(bar(B_ACC1_FULL0), 3);
            ptx::umma_commit_2cta(bar(B_ACC1_FULL1), 3);
          }
          __syncwarp();
          continue;
        }
        // ---- layer 1, row block 0 (rows [0,64) of both CTAs): K half 0 then K half 1 ----
        const long long i2 = SGT();
        if (it > 0) ptx::mbar_wait(bar(B_ACC1_EMPTY0), (it - 1) & 1);
        const long long i3 = SGT();
        ptx::mbar_wait(bar(B_A1_FULL0), it & 1);
        ptx::tc_fence_after();
        const long long i4 = SGT();
        if (lane == 0) { SGP(2, i3 - i2); SGP(3, i4 - i3); }
        if (ptx::elect_one()) {
          if (FOLD) ptx::umma_f16_2cta(tmem + 256, dW0 + kadv(KB1), dA0 + kadv(KB1), idesc_n128, 0);
#pragma unroll
          for (int k = 0; k < 8; ++k)
            ptx::umma_f16_2cta(tmem + 256, dW1 + kadv(k), dA1 + kadv(k), idesc_n128, FOLD || k > 0);
        }
        __syncwarp();
        const long long i5 = SGT();
        ptx::mbar_wait(bar(B_A1_FULL1), it & 1);
        ptx::tc_fence_after();
        if (lane == 0) SGP(4, SGT() - i5);
        if (IN == IN_IMAGE && t + n_clusters < p.T) {
          // epi 0 has drained both halves of acc0, so layer 0 is complete and A0_EMPTY has fired: this wait falls through
          ptx::mbar_wait(bar(B_A0_EMPTY), it & 1);
          if (ptx::elect_one()) tma_tile(t + n_clusters);
          __syncwarp();
        }
        if (ptx::elect_one()) {
#pragma unroll
          for (int k = 8; k < 16; ++k) ptx::umma_f16_2cta(tmem + 256, dW1 + kadv(k), dA1 + kadv(k), idesc_n128, 1);
          ptx::umma_commit_2cta(bar(B_ACC1_FULL0), 3);
        }
        __syncwarp();
        // ---- layer 1, row block 1 (rows [64,128) of both CTAs) ----
        const long long i6 = SGT();
        if (it > 0) ptx::mbar_wait(bar(B_ACC1_EMPTY1), (it - 1) & 1);
        ptx::tc_fence_after();
        if (lane == 0) { SGP(5, SGT() - i6); SGP(6, SGT() - i0); }
        if (ptx::elect_one()) {
          if (FOLD) ptx::umma_f16_2cta(tmem + 384, dW0 + kadv(KB1), dA0 + rows64 + kadv(KB1), idesc_n128, 0);
#pragma unroll
          for (int k = 0; k < 16; ++k)
            ptx::umma_f16_2cta(tmem + 384, dW1 + kadv(k), dA1 + rows64 + kadv(k), idesc_n128, FOLD || k > 0);
          ptx::umma_commit_2cta(bar(B_ACC1_FULL1), 3);
        }
        __syncwarp();
      }
    }
  }

  // ---- teardown -------------------------------------------------------------------------------
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();
  if (warp == EPI1_WARP0) ptx::tmem_dealloc_2cta(tmem, 512);
#ifdef SEGSUM_PROF
  if (blockIdx.x == 0 && threadIdx.x < 32) seg_prof[threadIdx.x] = sprof[threadIdx.x];
#endif
}

// ---------------------------------------------------------------------------------------------
// pre-pass, one thread per 64-row sub-block: bit masks of the last rows of the pieces that tile the
// sub-block and of the partial ones, the first program, and (irregular sub-blocks only) the explicit
// piece list; and one thread per program: zero the output rows that receive atomics (programs cut by a sub-block boundary) or nothing at all (empty programs).
__global__ void tc_prepare_kernel(const int32_t* __restrict__ off, int P, int n_rows, int nsub, uint4* __restrict__ meta,
                                  uint2* __restrict__ desc, float* __restrict__ out) {
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid < nsub) {
    const int r0 = gid * tc::SUB, r1 = min(r0 + tc::SUB, n_rows);
    int lo = 0, hi = P;                           // last p in [0, P) with off[p] <= r0
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (off[mid] <= r0) lo = mid; else hi = mid;
    }
    uint64_t ends = 0, parts = 0;
    int first = -1, last = -1, k = 0;
    for (int pass = 0; pass < 2; ++pass) {        // pass 1 only for irregular sub-blocks: the explicit piece list
      uint2* d = desc + (size_t)gid * tc::SUB;
      int p = lo, r = r0;
      k = 0;
      while (r < r1 && p < P) {                   // p < P: offsets that end before n_rows (caller error) must not walk off the array
        const int s = off[p], e = off[p + 1];
        if (e > r) {                              // skips empty programs
          const int ee = min(e, r1);
          const unsigned partial = (s < r0 || e > r1) ? 1u : 0u;
          if (pass) d[k] = make_uint2((unsigned)p, (unsigned)(r - r0) | ((unsigned)(ee - r) << 8) | (partial << 16));
          ++k;
          ends |= 1ull << (ee - 1 - r0);
          parts |= (uint64_t)partial << (ee - 1 - r0);
          if (first < 0) first = p;
          last = p;
          r = ee;
        }
        ++p;
      }
      if (k == 0 || last - first + 1 == k) break; // regular: pieces belong to consecutive programs
    }
    const unsigned irregular = (k > 0 && last - first + 1 != k) ? 1u : 0u;
    meta[2 * (size_t)gid] = make_uint4((unsigned)ends, (unsigned)(ends >> 32), (unsigned)parts, (unsigned)(parts >> 32));
    meta[2 * (size_t)gid + 1] = make_uint4((unsigned)max(first, 0), irregular, 0u, 0u);
  }
  const int lane = threadIdx.x & 31;
  const int pbase = gid - lane;                   // 32 programs per warp
  int need = 0;
  if (gid < P) {
    const int s = off[gid], e = off[gid + 1];
    need = (e == s) || (s / tc::SUB != (e - 1) / tc::SUB);
  }
  unsigned m = __ballot_sync(0xffffffffu, need);
  while (m) {
    const int b = __ffs(m) - 1;
    m &= m - 1;
    float4* dst = reinterpret_cast<float4*>(out + (size_t)(pbase + b) * tc::HID);
    dst[lane] = make_float4(0.f, 0.f, 0.f, 0.f);
    dst[lane + 32] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// valid[p] = any row of p has a non-zero feature (MLPModel.predict's rule, mlp_model.py:842-844)
__global__ void tc_valid_kernel(const int32_t* __restrict__ off, int P, const uint32_t* __restrict__ rowmask,
                                uint8_t* __restrict__ valid) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  int any = 0;
  for (int r = off[p]; r < off[p + 1] && !any; ++r) any = (rowmask[r >> 5] >> (r & 31)) & 1u;
  valid[p] = (uint8_t)any;
}

// ---------------------------------------------------------------------------------------------
static __half h_clamp(float x) { return __float2half_rn(std::min(std::max(x, -65504.f), 65504.f)); }

int tc_pack(vaex_model* m, const std::vector<float>& w0, cudaStream_t st) {
  using namespace tc;
  m->tc.ready = false;
  if (m->H != HID || m->F >= K0P || m->F < 1) return VAEX_OK;   // shape outside the tcgen05 kernel: SIMT path
  const int F = m->F;
  std::vector<__half> i0(2 * W0_BYTES / 2, __float2half_rn(0.f)), i1(2 * W1_BYTES / 2);
  const std::vector<float>& w1 = m->host["segment_encoder.2.weight"].v;
  const std::vector<float>& b0 = m->host["segment_encoder.0.bias"].v;
  const std::vector<float>& b1 = m->host["segment_encoder.2.bias"].v;
  const bool fold = F + 2 <= B1_COL;                            // b1 as a 17th K step of layer 1
  for (int r = 0; r < 2; ++r)
    for (int row = 0; row < TILE; ++row) {
      const int h = r * TILE + row;
      __half* d0 = i0.data() + (size_t)r * (W0_BYTES / 2);
      __half* d1 = i1.data() + (size_t)r * (W1_BYTES / 2);
      for (int k = 0; k < F; ++k) d0[sw128_off(row, k) / 2] = h_clamp(w0[(size_t)h * F + k]);
      // bias columns (X supplies the 1s): hi + lo fp16 halves keep ~22 bits of the fp32 bias
      const __half b0h = h_clamp(b0[h]);
      d0[sw128_off(row, F) / 2] = b0h;
      if (F + 1 < K0P) d0[sw128_off(row, F + 1) / 2] = h_clamp(b0[h] - __half2float(b0h));
      if (fold) {
        const __half b1h = h_clamp(b1[h]);
        d0[sw128_off(row, B1_COL) / 2] = b1h;
        d0[sw128_off(row, B1_COL + 1) / 2] = h_clamp(b1[h] - __half2float(b1h));
      }
      for (int c = 0; c < HID; ++c) d1[sw128_off(row, c) / 2] = h_clamp(w1[(size_t)h * HID + hidden_of_col(c)]);
    }
  TcPack& t = m->tc;
  if (!t.w0) VAEX_CUDA(cudaMalloc(&t.w0, 2 * W0_BYTES));
  if (!t.w1) VAEX_CUDA(cudaMalloc(&t.w1, 2 * W1_BYTES));
  if (!t.b1) VAEX_CUDA(cudaMalloc(&t.b1, HID * sizeof(float)));
  VAEX_CUDA(cudaMemcpyAsync(t.w0, i0.data(), 2 * W0_BYTES, cudaMemcpyHostToDevice, st));
  VAEX_CUDA(cudaMemcpyAsync(t.w1, i1.data(), 2 * W1_BYTES, cudaMemcpyHostToDevice, st));
  VAEX_CUDA(cudaMemcpyAsync(t.b1, m->host["segment_encoder.2.bias"].v.data(), HID * sizeof(float), cudaMemcpyHostToDevice, st));
  VAEX_CUDA(cudaStreamSynchronize(st));
  VAEX_CUDA(cudaFuncSetAttribute(segsum_tc_kernel<true, IN_F32>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_DYN));
  VAEX_CUDA(cudaFuncSetAttribute(segsum_tc_kernel<false, IN_F32>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_DYN));
  VAEX_CUDA(cudaFuncSetAttribute(segsum_tc_kernel<true, IN_F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_DYN));
  VAEX_CUDA(cudaFuncSetAttribute(segsum_tc_kernel<false, IN_F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_DYN));
  VAEX_CUDA(cudaFuncSetAttribute(segsum_tc_kernel<true, IN_IMAGE>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_DYN));
  t.F = F;
  t.K0P = K0P;
  t.ready = true;
  return VAEX_OK;
}

// ---------------------------------------------------------------------------------------------
// The packed row image (IN_IMAGE): fp16 [n_rows, 192] with a 384-byte pitch -- columns [0, F) the features (saturated,
// round to nearest even: what the loader warps produce on the fly), F and F + 1 the constant 1 that multiplies the b0
// hi / lo columns of the W0 image, 176 and 177 the constant 1 of the b1 k-step, everything else 0 -- followed by the
// per-row "has a non-zero feature" bits (uint32 words; taken from the fp32 values, so a value that rounds to an fp16
// zero still counts: MLPModel.predict's rule, mlp_model.py:842-844).  One 16-byte store per thread.
// One CTA = 32 rows x 24 chunks of 8 halves = one word of the non-zero mask: a 16-byte store per thread, one shared
// atomic per warp-row, one plain store of the mask word (no global atomics, no memset).
constexpr int PACK_THREADS = 32 * (tc::K0P / 8);        // 768
__global__ void __launch_bounds__(PACK_THREADS) rows_pack_image_kernel(const float* __restrict__ feats, int64_t n_rows, int F,
                                                                    uint4* __restrict__ img, uint32_t* __restrict__ rowmask) {
  constexpr int CH = tc::K0P / 8;
  __shared__ uint32_t word;
  if (threadIdx.x == 0) word = 0;
  __syncthreads();
  const int rl = threadIdx.x / CH, c = threadIdx.x - rl * CH;
  const int64_t row = (int64_t)blockIdx.x * 32 + rl;
  const float* src = feats + row * F;
  uint32_t w[4], any = 0;
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    float v[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int col = 8 * c + 2 * e + i;
      float x = 0.f;
      if (col < F) { if (row < n_rows) { x = __ldg(src + col); any |= __float_as_uint(x) << 1; } }
      else if (col == F || col == F + 1 || col == tc::B1_COL || col == tc::B1_COL + 1) x = 1.f;
      v[i] = x;
    }
    w[e] = ptx::cvt_h2_sat(v[0], v[1]);
  }
  img[(size_t)row * CH + c] = make_uint4(w[0], w[1], w[2], w[3]);
  if (any) atomicOr(&word, 1u << rl);
  __syncthreads();
  if (threadIdx.x == 0) rowmask[blockIdx.x] = word;
}

// The image is padded to whole 256-row pair tiles and the pad rows carry the bias-one columns too: a tile load never
// runs out of bounds.  (It must not: layer 1 of tile t still reads the b1 k-step of A0 -- columns 176, 177, the same
// values in every row of every tile -- while the loads of tile t + 1 land; an out-of-bounds row would arrive as zeros and
// take the bias away from the previous tile's row.)
static int64_t image_rows_padded(int64_t n_rows) { return (n_rows + tc::PAIR - 1) / tc::PAIR * tc::PAIR; }
static size_t image_rows_bytes(int64_t n_rows) { return (size_t)image_rows_padded(n_rows) * tc::K0P * sizeof(__half); }
int64_t rows_image_bytes(int64_t n_rows) {
  return (int64_t)(image_rows_bytes(n_rows) + sizeof(uint32_t) * (size_t)((n_rows + 255) / 256) * 8);
}

int rows_pack_image(const float* feats, int64_t n_rows, int F, void* image, cudaStream_t st) {
  using namespace tc;
  VAEX_CHECK(F >= 1 && F + 2 <= B1_COL, "the row image holds 1..174 features");
  VAEX_CHECK(n_rows >= 0 && n_rows < (1ll << 31) - PAIR, "row count exceeds int32 indexing");
  if (n_rows == 0) return VAEX_OK;
  VAEX_CHECK(feats && image && (reinterpret_cast<uintptr_t>(image) & 127) == 0, "image_dev must be 128-byte aligned");
  uint32_t* mask = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(image) + image_rows_bytes(n_rows));
  const int64_t n_pad = image_rows_padded(n_rows);
  rows_pack_image_kernel<<<(unsigned)(n_pad / 32), PACK_THREADS, 0, st>>>(feats, n_rows, F, reinterpret_cast<uint4*>(image), mask);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time dependency on libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static int image_tensor_map(const void* image, int64_t n_rows, CUtensorMap* out) {
  using namespace tc;
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    VAEX_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    VAEX_CHECK(p && q == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled is not available in this driver");
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  const cuuint64_t dims[2] = {(cuuint64_t)K0P, (cuuint64_t)image_rows_padded(n_rows)};   // innermost first; rows padded: no OOB
  const cuuint64_t strides[1] = {(cuuint64_t)K0P * sizeof(__half)};        // bytes between rows
  const cuuint32_t box[2] = {64, (cuuint32_t)TILE};                        // one 128-byte-swizzle atom: 64 halves x 128 rows
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(image), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r)); return VAEX_E_CUDA; }
  return VAEX_OK;
}

int segsum_tc_encode(vaex_model* m, const float* feats, const int32_t* off, int64_t P, int64_t R, float* segsum,
                     uint8_t* valid, cudaStream_t st, int in_mode) {
  using namespace tc;
  VAEX_CHECK(m->tc.ready, "tcgen05 weights not packed");
  VAEX_CHECK(P < (1ll << 31) && R < (1ll << 31) - PAIR, "sizes exceed int32 row indexing");
  const int T = (int)cdiv(R, PAIR);
  const int nsub = 4 * T;
  const int nprep = (int)std::max<int64_t>(P, nsub);
  const size_t meta_bytes = 2 * sizeof(uint4) * (size_t)std::max(nsub, 1);
  const size_t desc_bytes = sizeof(uint2) * (size_t)std::max(nsub, 1) * SUB;
  { int rc = m->tc_ws.ensure(meta_bytes + desc_bytes + sizeof(uint32_t) * (size_t)std::max(T, 1) * 8); if (rc != VAEX_OK) return rc; }
  uint4* meta = m->tc_ws.as<uint4>();
  uint2* desc = reinterpret_cast<uint2*>(reinterpret_cast<uint8_t*>(meta) + meta_bytes);
  uint32_t* rowmask = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(desc) + desc_bytes);
  tc_prepare_kernel<<<(unsigned)cdiv(nprep, 256), 256, 0, st>>>(off, (int)P, (int)R, nsub, meta, desc, segsum);
  VAEX_LAUNCH_CHECK();
  if (T > 0) {
    TcParams prm;
    prm.feats = feats; prm.off = off; prm.desc = desc; prm.meta = meta;
    prm.w0 = m->tc.w0; prm.w1 = m->tc.w1; prm.b1 = m->tc.b1;
    prm.out = segsum; prm.rowmask = rowmask; prm.n_rows = R; prm.P = (int)P; prm.T = T; prm.F = m->F;
    prm.k0_steps = (m->F + (m->F + 1 < K0P ? 2 : 1) + 15) / 16;   // + the bias hi/lo columns
    const bool fold = m->F + 2 <= B1_COL;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int clusters = std::max(1, std::min(sms / 2, T));
    CUtensorMap tmap;
    memset(&tmap, 0, sizeof(tmap));
    if (in_mode == IN_IMAGE) {
      VAEX_CHECK(fold, "the row image needs F <= 174");
      VAEX_CHECK((reinterpret_cast<uintptr_t>(feats) & 127) == 0, "the row image must be 128-byte aligned");
      TRY(image_tensor_map(feats, R, &tmap));
      // the non-zero bits were taken from the fp32 rows when the image was packed
      rowmask = reinterpret_cast<uint32_t*>(const_cast<uint8_t*>(reinterpret_cast<const uint8_t*>(feats)) + image_rows_bytes(R));
      prm.rowmask = rowmask;
      segsum_tc_kernel<true, IN_IMAGE><<<2 * clusters, THREADS, SMEM_DYN, st>>>(prm, tmap);
    } else if (in_mode == IN_F16) {
      VAEX_CHECK(m->F % 4 == 0 && (reinterpret_cast<uintptr_t>(feats) & 15) == 0,
                 "fp16 features need F % 4 == 0 and a 16-byte aligned pointer");
      if (fold) segsum_tc_kernel<true, IN_F16><<<2 * clusters, THREADS, SMEM_DYN, st>>>(prm, tmap);
      else segsum_tc_kernel<false, IN_F16><<<2 * clusters, THREADS, SMEM_DYN, st>>>(prm, tmap);
    } else if (fold) segsum_tc_kernel<true, IN_F32><<<2 * clusters, THREADS, SMEM_DYN, st>>>(prm, tmap);
    else segsum_tc_kernel<false, IN_F32><<<2 * clusters, THREADS, SMEM_DYN, st>>>(prm, tmap);
    VAEX_LAUNCH_CHECK();
  }
  if (valid) {
    tc_valid_kernel<<<(unsigned)cdiv(P, 256), 256, 0, st>>>(off, (int)P, rowmask, valid);
    VAEX_LAUNCH_CHECK();
  }
  return VAEX_OK;
}

}  // namespace vaex

#ifdef SEGSUM_PROF
extern "C" int vaex_debug_segsum_prof(unsigned long long* out) {
  return (int)cudaMemcpyFromSymbol(out, vaex::seg_prof, sizeof(unsigned long long) * 32);
}
#endif
