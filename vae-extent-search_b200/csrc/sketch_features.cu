// Per-statement feature rows of a fixed sketch as a function of the loop extents.
//
// The reference lowers every candidate with TVM on the CPU and walks the TIR (PerStoreFeatureExtractor,
// src/auto_scheduler/feature.cc:609-1048; vector layout GetPerStoreFeature :1053-1188).  Within one sketch only the
// extents differ between candidates, so the walk is run here on a TEMPLATE of the lowered loop nest -- an int32 program
// compiled on the host by vae-extent-search_b200/sketch_features.py -- with one thread per candidate substituting its
// extents: touched regions of every enclosing loop (ComputeRegion :471-499 with the const-int-bound rules of
// src/arith/const_int_bound.cc), strides (CoefficientExtractor / ComputeStride :382-468), reuse (ComputeReuse :502-603),
// the arithmetic-intensity curve (:943-979), allocation and outer-scope features, annotations and thread extents.
// Loops whose extent evaluates to 1 vanish (ScheduleOps drops trivial loops) and index expressions are folded the way
// the Simplify pass in front of the extraction folds them (x * 1, 0 + x, floordiv / floormod below the divisor).
//
// Program layout (int32 words):
//   [0] magic  [1] slots  [2] scalars  [3] buffers  [4] loops  [5] rows  [6] instructions
//   [7] scalar table  [8] buffer table  [9] instruction table  [10] store table  [11] expression pool   (word offsets)
//   scalar: const, n_mul, n_div, add, slot[8]          value = const * prod(extents[mul]) / prod(extents[div]) + add
//   buffer: bytes, ndim, shape[4], row (-1: never stored)
//   instruction: op, a, b, c                            FOR(loop, scalar, kind) FOR_END THREAD(loop, scalar, which) THREAD_END
//                                                       PRAGMA(-, scalar) PRAGMA_END REALIZE_END(buffer, aux) STORE(store)
//   store: buffer, n_access, ops[16], then per access: buffer, acc_type, n_tuples, ndim, expression offset[n_tuples * ndim],
//          then n_loads, expression offset[n_loads] (the indices of every BufferLoad: their arithmetic counts as int ops)
//   expression: n_tokens, token[n]; token = kind | arg << 4: VAR(loop) CONST(scalar) ADD MUL(which operand is a leaf) DIV MOD
//
// Each warp stages its 32 rows in shared memory and writes them out as contiguous 656-byte rows (coalesced).
#include <cstdint>

#include "../../include/vaex_b200.h"
#include "common.cuh"

namespace vaex {
namespace sf {
constexpr int MAX_SLOTS = 64, MAX_SCALARS = 64, MAX_LOOPS = 32, MAX_BUFS = 6, MAX_DIMS = 4, SCALAR_SLOTS = 8;       // an expression has at most 32 tokens
constexpr int MAX_DEPTH = 24, MAX_ACC = 5, MAX_EXPR = 40, ROW = 164, ROW_PAD = 165, WARPS = 2, THREADS = 32 * WARPS;
constexpr int MAX_PROG = 4096;                            // words of program kept in shared memory
constexpr int MAGIC = 0x56465831;
enum { OP_FOR, OP_FOR_END, OP_THREAD, OP_THREAD_END, OP_PRAGMA, OP_PRAGMA_END, OP_REALIZE_END, OP_STORE };
enum { T_VAR, T_CONST, T_ADD, T_MUL, T_DIV, T_MOD };
enum { K_READ, K_WRITE, K_READWRITE };
enum { KIND_FOR, KIND_VEC, KIND_UNROLL, KIND_PAR, KIND_THREAD };

__device__ __forceinline__ float slog(float x) { return x < 0.f ? -log2f(-x + 1.f) : log2f(x + 1.f); }   // feature.cc:1051

struct Walk {
  int64_t S[MAX_SCALARS];
  int32_t ext[MAX_LOOPS];
  int8_t pos[MAX_LOOPS];                                  // stack position of a live loop, -1 otherwise
  int8_t stack[MAX_DEPTH];
  uint8_t kind[MAX_DEPTH];
  // for_touch_regions_ of the loops on the stack, as the aggregates the walk reads (:1028)
  uint8_t cnt[MAX_DEPTH][MAX_BUFS];                       // accesses recorded for (loop, buffer)
  float min_sz[MAX_DEPTH][MAX_BUFS];                      // smallest touched size among them
  double first_sz[MAX_DEPTH][MAX_BUFS];                   // touched size of the first one
  double sum_all[MAX_DEPTH], sum_b1[MAX_DEPTH], sum_read[MAX_DEPTH];   // sum size * bytes; sum bytes; the same over reads
  float row_prod[MAX_BUFS][3];                            // vec / unroll / parallel _prod persist across stores of a buffer
  float row_outer[MAX_BUFS];
  float alloc[MAX_BUFS][4];
  float outer;
  int depth, is_gpu, auto_unroll;
  int tlen[7];
  int saved_unroll[4], n_saved;
};

struct Prog {
  const int32_t* w;
  __device__ __forceinline__ const int32_t* scalar(int i) const { return w + w[7] + i * (4 + SCALAR_SLOTS); }
  __device__ __forceinline__ const int32_t* buffer(int i) const { return w + w[8] + i * (3 + MAX_DIMS); }
  __device__ __forceinline__ const int32_t* instr(int i) const { return w + w[9] + 4 * i; }
  __device__ __forceinline__ const int32_t* store(int off) const { return w + w[10] + off; }
  __device__ __forceinline__ const int32_t* expr(int off) const { return w + w[11] + off; }
};

// Which tokens survive the folds of Simplify (a bit per token; the survivors are again a postfix expression, or none = 0)
__device__ uint32_t simplify(const Prog& P, const Walk& W, const int32_t* e, uint32_t* var_mask) {
  const int n = e[0];
  uint32_t alive = n >= 32 ? 0xffffffffu : ((1u << n) - 1u);
  struct Ent { int64_t fhi, val; int8_t start; bool zero, imm; };
  Ent st[8];
  int sp = 0;
  auto kill = [&](int from, int to) { for (int i = from; i <= to; ++i) alive &= ~(1u << i); };
  for (int i = 0; i < n; ++i) {
    const int tok = e[1 + i], k = tok & 15, arg = tok >> 4;
    if (k == T_VAR) {
      Ent x; x.start = (int8_t)i; x.imm = false; x.val = 0;
      if (W.pos[arg] < 0) { x.zero = true; x.fhi = 0; kill(i, i); }
      else { x.zero = false; x.fhi = W.ext[arg] - 1; }
      st[sp++] = x;
    } else if (k == T_CONST) {
      Ent x; x.start = (int8_t)i; x.val = W.S[arg]; x.fhi = x.val; x.zero = x.val == 0; x.imm = !x.zero;
      if (x.zero) kill(i, i);
      st[sp++] = x;
    } else {
      const Ent b = st[--sp], a = st[--sp];
      Ent r; r.start = a.start; r.imm = false; r.val = 0; r.zero = false; r.fhi = 0;
      const int b_end = i - 1;                            // b's tokens are [b.start, i - 1], a's [a.start, b.start - 1]
      if (k == T_ADD) {
        if (a.zero && b.zero) { r.zero = true; kill(i, i); }
        else if (a.zero) { r = b; r.start = a.start; kill(i, i); }
        else if (b.zero) { r = a; kill(i, i); }
        else r.fhi = a.fhi + b.fhi;
      } else if (k == T_MUL) {
        if (a.zero || b.zero) { r.zero = true; kill(a.start, i); }
        else if (a.imm && a.val == 1) { r = b; r.start = a.start; kill(a.start, b.start - 1); kill(i, i); }
        else if (b.imm && b.val == 1) { r = a; kill(b.start, i); }
        else r.fhi = a.fhi * b.fhi;
      } else {                                            // floordiv / floormod by a positive constant
        const int64_t c = b.val;
        if (a.zero) { r.zero = true; kill(a.start, i); }
        else if (c == 1) {
          if (k == T_DIV) { r = a; kill(b.start, i); } else { r.zero = true; kill(a.start, i); }
        } else if (a.fhi < c) {
          if (k == T_DIV) { r.zero = true; kill(a.start, i); } else { r = a; kill(b.start, i); }
        } else r.fhi = k == T_DIV ? a.fhi / c : min(a.fhi, c - 1);
      }
      (void)b_end;
      st[sp++] = r;
    }
  }
  uint32_t vm = 0;
  for (int i = 0; i < n; ++i)
    if (((alive >> i) & 1u) && (e[1 + i] & 15) == T_VAR) vm |= 1u << (e[1 + i] >> 4);
  *var_mask |= vm;
  return alive;
}

// const_int_bound of the surviving expression with loops at stack positions >= level over their range, the others at 0
__device__ void bound(const Prog& P, const Walk& W, const int32_t* e, uint32_t alive, int level, int64_t* lo_out, int64_t* hi_out) {
  int64_t lo[8], hi[8];
  int sp = 0;
  const int n = e[0];
  for (int i = 0; i < n; ++i) {
    if (!((alive >> i) & 1u)) continue;
    const int tok = e[1 + i], k = tok & 15, arg = tok >> 4;
    if (k == T_VAR) { lo[sp] = 0; hi[sp] = W.pos[arg] >= level ? W.ext[arg] - 1 : 0; ++sp; }
    else if (k == T_CONST) { lo[sp] = hi[sp] = W.S[arg]; ++sp; }
    else {
      --sp;
      const int64_t bl = lo[sp], bh = hi[sp], al = lo[sp - 1], ah = hi[sp - 1];
      int64_t rl, rh;
      if (k == T_ADD) { rl = al + bl; rh = ah + bh; }
      else if (k == T_MUL) {
        const int64_t c0 = al * bl, c1 = al * bh, c2 = ah * bl, c3 = ah * bh;
        rl = min(min(c0, c1), min(c2, c3)); rh = max(max(c0, c1), max(c2, c3));
      } else if (k == T_DIV) {
        auto fd = [](int64_t x, int64_t c) { int64_t q = x / c; return (x % c != 0 && (x < 0)) ? q - 1 : q; };
        rl = fd(al, bl); rh = fd(ah, bl);
      } else {
        if (al >= 0) { if (ah < bl) { rl = al; rh = ah; } else { rl = 0; rh = min(ah, bl - 1); } }
        else { rl = 0; rh = bl - 1; }
      }
      lo[sp - 1] = rl; hi[sp - 1] = rh;
    }
  }
  if (sp == 0) { *lo_out = 0; *hi_out = 0; } else { *lo_out = lo[0]; *hi_out = hi[0]; }
}

// CoefficientExtractor::ExtractCoefficient (:382-443) on the surviving tokens; returns visited_var through *found
__device__ int64_t coefficient(const Prog& P, const Walk& W, const int32_t* e, uint32_t alive, int loop, bool* found) {
  bool var = false, mul = false, add = false;
  int64_t stride = 0;
  // whether an operand is an immediate leaf must be judged on the SURVIVING tree: remember what the last two operands were
  bool leaf[8];
  int64_t val[8];
  int sp = 0;
  const int n = e[0];
  for (int i = 0; i < n; ++i) {
    if (!((alive >> i) & 1u)) continue;
    const int tok = e[1 + i], k = tok & 15, arg = tok >> 4;
    if (k == T_VAR) { if (arg == loop) { var = true; stride = 2; } leaf[sp] = false; val[sp] = 0; ++sp; }
    else if (k == T_CONST) { leaf[sp] = true; val[sp] = W.S[arg]; ++sp; }
    else {
      --sp;
      if (k == T_MUL) {
        if (var && !add) {
          if (leaf[sp - 1]) { mul = true; stride = val[sp - 1]; }
          else if (leaf[sp]) { mul = true; stride = val[sp]; }
        }
      } else if (k == T_ADD) {
        if (var && !mul) { add = true; stride = 1; }
      }
      leaf[sp - 1] = false;
    }
  }
  *found = var;
  return (var && !mul && !add) ? 1 : stride;
}

}  // namespace sf
}  // namespace vaex

#include "sketch_features_walk.inc"
