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
//   scalar: const, n_mul, n_div, add, slot[12]         value = const * prod(extents[mul]) / prod(extents[div]) + add
//   buffer: bytes, ndim, shape[4], row (-1: never stored)
//   instruction: op, a, b, c                            FOR(loop, scalar, kind) FOR_END THREAD(loop, scalar, which) THREAD_END
//                                                       PRAGMA(-, scalar) PRAGMA_END REALIZE_END(buffer, aux) STORE(store)
//   instruction STORE: b = 1 for the last store of its buffer (the one whose features the buffer keeps)
//   store: buffer, n_access, ops[16], then per access: buffer, acc_type, n_tuples, ndim, expression offset[n_tuples * ndim],
//          n_walk (-1: none), walk entry[n_walk] -- the terms of all dimensions sorted by loop, innermost first, constants
//          last: loop id + 1 (0: constant) | dimension << 8 | term << 12; only for one index tuple all in term form;
//          then n_loads, (position in the store's expression list or -1, expression offset)[n_loads]: the indices of
//          every BufferLoad -- their arithmetic counts as int ops
//   expression: n_tokens, ADD mask, MUL mask, DIV|MOD mask (a bit per token), n_terms (-1: none), token[n], term[n_terms][4];
//               term = kind | loop << 4, coefficient scalar (-1: none), c1, c2 -- the stack-free form of a sum of
//               v * c, (v // c1) * c, (v % c1) * c, ((v // c1) % c2) * c and constants;  token = kind | arg << 4: VAR(loop) CONST(scalar) ADD MUL(which operand is a leaf) DIV MOD
//
// One block of 256 threads per SM, a thread per candidate.  Each warp stages its rows through a 32 x 33-word tile of shared
// memory and writes them out in 128-byte segments (fp32 rows, or fp16 rows for the row encoder's VAEX_F_FEATS_F16 input).
#include <algorithm>
#include <cstdint>
#include <mutex>
#include <type_traits>
#include <vector>

#include <cuda_fp16.h>

#include "../../include/vaex_b200.h"
#include "common.cuh"

namespace vaex {
namespace sf {
constexpr int MAX_SLOTS = 64, MAX_SCALARS = 48, MAX_LOOPS = 32, MAX_BUFS = 6, MAX_DIMS = 4, SCALAR_SLOTS = 12;       // an expression has at most 32 tokens
constexpr int MAX_DEPTH = 24, MAX_ACC = 5, MAX_EXPR = 40, ROW = 164, WARPS = 8, THREADS = 32 * WARPS;
constexpr int MAX_PROG = 4096;                            // words of program kept in shared memory
constexpr int MAGIC = 0x56465831;
enum { OP_FOR, OP_FOR_END, OP_THREAD, OP_THREAD_END, OP_PRAGMA, OP_PRAGMA_END, OP_REALIZE_END, OP_STORE };
enum { T_VAR, T_CONST, T_ADD, T_MUL, T_DIV, T_MOD };
enum { K_READ, K_WRITE, K_READWRITE };
enum { KIND_FOR, KIND_VEC, KIND_UNROLL, KIND_PAR, KIND_THREAD };

// feature.cc:1051.  __log2f (the special-function unit: 2^-22 absolute on [0.5, 2], 2 ulp elsewhere) against the reference's
// log2f: below the 2e-6 the parity tests allow, a twentieth of the instructions -- and every row has 100+ of these
__device__ __forceinline__ float slog(float x) { return copysignf(__log2f(fabsf(x) + 1.f), x); }

struct Walk {
  int64_t S[MAX_SCALARS];
  int32_t ext[MAX_LOOPS];
  int8_t pos[MAX_LOOPS];                                  // stack position of a live loop, -1 otherwise
  int8_t stack[MAX_DEPTH];
  uint8_t kind[MAX_DEPTH];
  // for_touch_regions_ of the loops on the stack, as the aggregates the walk reads (:1028)
  uint8_t cnt[MAX_DEPTH][MAX_BUFS];                       // accesses recorded for (loop, buffer)
  float min_sz[MAX_DEPTH][MAX_BUFS];                      // smallest touched size among them
  double first_sz[MAX_BUFS];                              // touched size of the first one, outermost loop only (:863)
  float sum_all[MAX_DEPTH], sum_b1[MAX_DEPTH], sum_read[MAX_DEPTH];    // sum size * bytes; sum bytes; the first over reads
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

// The expression stacks live in registers: four entries (the compiler checks that no index expression needs more), moved
// by hand instead of indexed, so that no evaluation touches local memory.
template <class T>
struct Stack4 {
  T s0, s1, s2, s3;                                       // s0 = top
  __device__ __forceinline__ void push(const T& x) { s3 = s2; s2 = s1; s1 = s0; s0 = x; }
  __device__ __forceinline__ void pop2(T& a, T& b) { b = s0; a = s1; s0 = s2; s1 = s3; }       // a was pushed before b
};
__device__ __forceinline__ int n_tokens(const int32_t* e) { return e[0]; }
__device__ __forceinline__ int token(const int32_t* e, int i) { return e[5 + i]; }               // e[1..3]: ADD / MUL / DIV|MOD masks
__device__ __forceinline__ int n_terms(const int32_t* e) { return e[4]; }                        // -1: no term form
__device__ __forceinline__ const int32_t* term_at(const int32_t* e, int i) { return e + 5 + e[0] + 4 * i; }

// Which tokens survive the folds of Simplify (a bit per token; the survivors are again a postfix expression, or none = 0)
__device__ uint32_t simplify(const Prog& P, const Walk& W, const int32_t* e, uint32_t* var_mask) {
  const int n = n_tokens(e);
  uint32_t alive = n >= 32 ? 0xffffffffu : ((1u << n) - 1u);
  uint32_t vm = 0;
  struct Ent { int64_t fhi, val; int start; bool zero, imm; };
  Stack4<Ent> st;
  auto kill = [&](int from, int to) { alive &= ~((0xffffffffu >> (31 - to)) & (0xffffffffu << from)); };
  for (int i = 0; i < n; ++i) {
    const int tok = token(e, i), k = tok & 15, arg = tok >> 4;
    if (k == T_VAR) {
      Ent x; x.start = i; x.imm = false; x.val = 0;
      if (W.pos[arg] < 0) { x.zero = true; x.fhi = 0; kill(i, i); }
      else { x.zero = false; x.fhi = W.ext[arg] - 1; vm |= 1u << i; }
      st.push(x);
    } else if (k == T_CONST) {
      Ent x; x.start = i; x.val = W.S[arg]; x.fhi = x.val; x.zero = x.val == 0; x.imm = !x.zero;
      if (x.zero) kill(i, i);
      st.push(x);
    } else {
      Ent a, b;
      st.pop2(a, b);
      Ent r; r.start = a.start; r.imm = false; r.val = 0; r.zero = false; r.fhi = 0;
      if (k == T_ADD) {                                   // b's tokens are [b.start, i - 1], a's [a.start, b.start - 1]
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
        const int64_t c = max(b.val, (int64_t)1);
        if (a.zero) { r.zero = true; kill(a.start, i); }
        else if (c == 1) {
          if (k == T_DIV) { r = a; kill(b.start, i); } else { r.zero = true; kill(a.start, i); }
        } else if (a.fhi < c) {
          if (k == T_DIV) { r.zero = true; kill(a.start, i); } else { r = a; kill(b.start, i); }
        } else r.fhi = k == T_DIV ? a.fhi / c : min(a.fhi, c - 1);
      }
      st.push(r);
    }
  }
  vm &= alive;                                            // variables inside a folded subtree are gone
  uint32_t loops = 0;
  while (vm) { const int i = __ffs(vm) - 1; vm &= vm - 1; loops |= 1u << (token(e, i) >> 4); }
  *var_mask |= loops;
  return alive;
}

// const_int_bound of the surviving expression with loops at stack positions >= level over their range, the others at 0
__device__ void bound(const Prog& P, const Walk& W, const int32_t* e, uint32_t alive, int level, int64_t* lo_out, int64_t* hi_out) {
  struct Iv { int64_t lo, hi; };
  Stack4<Iv> st;
  st.s0 = Iv{0, 0};
  const int n = n_tokens(e);
  for (int i = 0; i < n; ++i) {
    if (!((alive >> i) & 1u)) continue;
    const int tok = token(e, i), k = tok & 15, arg = tok >> 4;
    if (k == T_VAR) st.push(Iv{0, W.pos[arg] >= level ? (int64_t)W.ext[arg] - 1 : 0});
    else if (k == T_CONST) st.push(Iv{W.S[arg], W.S[arg]});
    else {
      Iv a, b, r;
      st.pop2(a, b);
      if (k == T_ADD) { r.lo = a.lo + b.lo; r.hi = a.hi + b.hi; }
      else if (k == T_MUL) {
        const int64_t c0 = a.lo * b.lo, c1 = a.lo * b.hi, c2 = a.hi * b.lo, c3 = a.hi * b.hi;
        r.lo = min(min(c0, c1), min(c2, c3)); r.hi = max(max(c0, c1), max(c2, c3));
      } else if (k == T_DIV) {
        auto fd = [](int64_t x, int64_t c) { int64_t q = x / c; return (x % c != 0 && (x < 0)) ? q - 1 : q; };
        const int64_t c = max(b.lo, (int64_t)1);
        r.lo = fd(a.lo, c); r.hi = fd(a.hi, c);
      } else {
        const int64_t c = max(b.lo, (int64_t)1);
        if (a.lo >= 0) { if (a.hi < c) r = a; else { r.lo = 0; r.hi = min(a.hi, c - 1); } }
        else { r.lo = 0; r.hi = c - 1; }
      }
      st.push(r);
    }
  }
  *lo_out = alive ? st.s0.lo : 0;
  *hi_out = alive ? st.s0.hi : 0;
}

// The same bounds for ALL levels in one pass, when every term of the expression depends on one loop variable (sums of
// var * c, floordiv / floormod chains of a single var, constants -- what a split / fuse lowers an index to): a term is
// [zlo, zhi] while its variable is pinned at 0 and [lo, hi] once the variable's loop is bound, so
//   lo(level) = sum zlo + sum over terms whose loop sits at position >= level of (lo - zlo),  likewise hi.
// Interval arithmetic adds term by term, so this equals bound() at every level.  The per-position deltas are ADDED into
// dlo / dhi (the caller clears them).  Returns false (caller falls back to bound() per level) for anything else: a
// product of two variables, floordiv / floormod or a scale of a sum.
__device__ bool bound_all_levels(const Prog& P, const Walk& W, const int32_t* e, uint32_t alive, int64_t* dlo, int64_t* dhi,
                                 int64_t* base_lo, int64_t* base_hi) {
  struct Ent { int64_t lo, hi, zlo, zhi; int pos; };     // pos: stack position of the term's variable, -1 constant, -2 flushed sum
  Stack4<Ent> st;
  st.s0 = Ent{0, 0, 0, 0, -1};
  int64_t bl = 0, bh = 0;
  auto flush = [&](const Ent& x) {
    if (x.pos == -2) return;
    bl += x.zlo; bh += x.zhi;
    if (x.pos >= 0) { dlo[x.pos] += x.lo - x.zlo; dhi[x.pos] += x.hi - x.zhi; }
  };
  auto fd = [](int64_t x, int64_t c) { const int64_t q = x / c; return (x % c != 0 && x < 0) ? q - 1 : q; };
  const int n = n_tokens(e);
  for (int i = 0; i < n; ++i) {
    if (!((alive >> i) & 1u)) continue;
    const int tok = token(e, i), k = tok & 15, arg = tok >> 4;
    if (k == T_VAR) { st.push(Ent{0, (int64_t)W.ext[arg] - 1, 0, 0, (int)W.pos[arg]}); continue; }
    if (k == T_CONST) { const int64_t v = W.S[arg]; st.push(Ent{v, v, v, v, -1}); continue; }
    Ent a, b, r;
    st.pop2(a, b);
    if (k == T_ADD) {
      flush(a); flush(b);
      r = Ent{0, 0, 0, 0, -2};
    } else if (k == T_MUL) {
      const bool ac = a.pos == -1, bc = b.pos == -1;
      if (!(ac || bc)) return false;
      const Ent x = ac ? b : a;
      const int64_t c = ac ? a.lo : b.lo;
      if (x.pos == -2) return false;
      r = x;
      r.lo = min(x.lo * c, x.hi * c); r.hi = max(x.lo * c, x.hi * c);
      r.zlo = min(x.zlo * c, x.zhi * c); r.zhi = max(x.zlo * c, x.zhi * c);
    } else {
      if (a.pos == -2 || b.pos != -1) return false;
      const int64_t c = max(b.lo, (int64_t)1);
      r = a;
      if (k == T_DIV) { r.lo = fd(a.lo, c); r.hi = fd(a.hi, c); r.zlo = fd(a.zlo, c); r.zhi = fd(a.zhi, c); }
      else {
        auto md = [c](int64_t l, int64_t h, int64_t* ol, int64_t* oh) {
          if (l >= 0) { if (h < c) { *ol = l; *oh = h; } else { *ol = 0; *oh = min(h, c - 1); } }
          else { *ol = 0; *oh = c - 1; }
        };
        md(a.lo, a.hi, &r.lo, &r.hi);
        md(a.zlo, a.zhi, &r.zlo, &r.zhi);
      }
    }
    st.push(r);
  }
  if (alive) flush(st.s0);
  *base_lo = bl; *base_hi = bh;
  return true;
}

// CoefficientExtractor::ExtractCoefficient (:382-443) on the surviving tokens; returns visited_var through *found
__device__ int64_t coefficient(const Prog& P, const Walk& W, const int32_t* e, uint32_t alive, int loop, bool* found) {
  bool var = false, mul = false, add = false;
  int64_t stride = 0;
  struct LV { int64_t val; bool leaf; };                  // whether an operand is an immediate leaf is judged on the SURVIVING tree
  Stack4<LV> st;
  const int n = n_tokens(e);
  for (int i = 0; i < n; ++i) {
    if (!((alive >> i) & 1u)) continue;
    const int tok = token(e, i), k = tok & 15, arg = tok >> 4;
    if (k == T_VAR) { if (arg == loop) { var = true; stride = 2; } st.push(LV{0, false}); }
    else if (k == T_CONST) st.push(LV{W.S[arg], true});
    else {
      LV a, b;
      st.pop2(a, b);
      if (k == T_MUL) {
        if (var && !add) {
          if (a.leaf) { mul = true; stride = a.val; }
          else if (b.leaf) { mul = true; stride = b.val; }
        }
      } else if (k == T_ADD) {
        if (var && !mul) { add = true; stride = 1; }
      }
      st.push(LV{0, false});
    }
  }
  *found = var;
  return (var && !mul && !add) ? 1 : stride;
}

// ---- the term form: most indices are a left-to-right sum of  base [* c]  with base = v | v // c1 | v % c1 | (v // c1) % c2, or
// constants.  Each term is folded, bounded and counted on its own -- no expression stack, a handful of instructions per term
// instead of a token loop.  The rules are those of simplify() / bound() / coefficient() applied to that shape of tree.
enum { TERM_PLAIN, TERM_DIV, TERM_MOD, TERM_DIVMOD, TERM_CONST };
struct Term { int64_t w, coef; int pos, loop, n_dm; bool alive, has_mul, is_const; };

__device__ __forceinline__ Term term_eval(const Walk& W, const int32_t* t) {
  Term r;
  const int kind = t[0] & 15;
  r.loop = t[0] >> 4; r.is_const = kind == TERM_CONST; r.n_dm = 0; r.w = 0; r.has_mul = false; r.coef = 1; r.pos = -1; r.alive = false;
  if (r.is_const) { r.w = W.S[t[1]]; r.alive = r.w != 0; return r; }          // w carries the constant's value
  r.pos = W.pos[r.loop];
  if (r.pos < 0) return r;
  int64_t fhi = (int64_t)W.ext[r.loop] - 1;
  if (kind == TERM_DIV || kind == TERM_DIVMOD) {
    const int64_t c = max(W.S[t[2]], (int64_t)1);               // a divisor below 1 only arises from an illegal tiling
    if (c != 1) { if (fhi < c) return r; fhi /= c; ++r.n_dm; }
  }
  if (kind == TERM_MOD || kind == TERM_DIVMOD) {
    const int64_t c = max(W.S[kind == TERM_MOD ? t[2] : t[3]], (int64_t)1);
    if (c == 1) return r;
    if (fhi >= c) { fhi = c - 1; ++r.n_dm; }
  }
  r.alive = true;
  if (t[1] >= 0) { r.coef = W.S[t[1]]; r.has_mul = r.coef != 1; }
  r.w = fhi * r.coef;
  return r;
}

// loops that occur in the folded expression; packed node counts: adds | muls << 8 | floordiv + floormod << 16
__device__ uint32_t terms_summary(const Walk& W, const int32_t* e, uint32_t* var_mask) {
  int n_alive = 0, n_mul = 0, n_dm = 0;
  for (int i = 0; i < n_terms(e); ++i) {
    const Term t = term_eval(W, term_at(e, i));
    if (!t.alive) continue;
    ++n_alive;
    if (!t.is_const) { *var_mask |= 1u << t.loop; n_mul += t.has_mul; n_dm += t.n_dm; }
  }
  return (uint32_t)max(n_alive - 1, 0) | ((uint32_t)n_mul << 8) | ((uint32_t)n_dm << 16);
}

// terms_summary, and in the same pass acc[position of the term's loop] += the width the term adds to the index once that
// loop is bound
__device__ uint32_t terms_summary_widths(const Walk& W, const int32_t* e, uint32_t* var_mask, int64_t* acc) {
  int n_alive = 0, n_mul = 0, n_dm = 0;
  for (int i = 0; i < n_terms(e); ++i) {
    const Term t = term_eval(W, term_at(e, i));
    if (!t.alive) continue;
    ++n_alive;
    if (!t.is_const) { *var_mask |= 1u << t.loop; n_mul += t.has_mul; n_dm += t.n_dm; acc[t.pos] += t.w; }
  }
  return (uint32_t)max(n_alive - 1, 0) | ((uint32_t)n_mul << 8) | ((uint32_t)n_dm << 16);
}

// CoefficientExtractor on ((t0 + t1) + t2) ...: per term the variable, then its Mul node, then (from the second term on) the Add
__device__ int64_t terms_coefficient(const Walk& W, const int32_t* e, int loop, bool* found) {
  bool var = false, mul = false, add = false, first = true;
  int64_t stride = 0;
  for (int i = 0; i < n_terms(e); ++i) {
    const Term t = term_eval(W, term_at(e, i));
    if (!t.alive) continue;
    if (!t.is_const) {
      if (t.loop == loop) { var = true; stride = 2; }
      if (t.has_mul && var && !add) { mul = true; stride = t.coef; }
    }
    if (!first && var && !mul) { add = true; stride = 1; }
    first = false;
  }
  *found = var;
  return (var && !mul && !add) ? 1 : stride;
}

}  // namespace sf
}  // namespace vaex

#include "sketch_features_walk.inc"
