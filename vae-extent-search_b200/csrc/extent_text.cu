// Extent parser and scaler.
//
// Host side: the reference turns printed Ansor states into extent vectors with two regular
// expressions per state (state_to_records, scripts/pre_experiments/model_myself/utils/extent.py:94-136;
// the same loop inlined in vae_extent_search.py:165-176).  Here that is one pass over the bytes of all
// states, spread over the host threads, straight into a (pinned) fp32 [n, d] matrix.
//
// Device side: np.log1p + sklearn StandardScaler.fit / transform (vae_extent_search.py:394-397) as two
// deterministic fp64 column reductions and one elementwise kernel.  The inverse (expm1(x*scale+mean), round,
// clamp) is the snap kernel of elementwise.cu.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "common.cuh"

namespace vaex {
namespace {

// Python's str.isspace() restricted to ASCII: what `\s` (and str.strip) match in a state string.
inline bool py_space(unsigned char c) { return c == ' ' || (c >= 9 && c <= 13) || (c >= 0x1c && c <= 0x1f); }
inline bool digit(unsigned char c) { return c >= '0' && c <= '9'; }

// float(digits) then np.float32: exact through 2^53, correctly rounded (strtod) beyond 15 digits.
inline float digits_to_f32(const char* p, const char* q) {
  if (q - p <= 15) {
    uint64_t v = 0;
    for (const char* s = p; s < q; ++s) v = v * 10 + (uint64_t)(*s - '0');
    return (float)(double)v;
  }
  std::string tmp(p, q);
  return (float)strtod(tmp.c_str(), nullptr);
}

// state.strip().split("Placeholder")[-1][2:]  (extent.py:101-105)
inline void schedule_span(const char* s, const char* e, const char** out_s, const char** out_e) {
  while (s < e && py_space((unsigned char)*s)) ++s;
  while (e > s && py_space((unsigned char)e[-1])) --e;
  static const char key[] = "Placeholder";
  const int kl = 11;
  const char* start = s;
  for (const char* p = e - kl; p >= s; --p)
    if (memcmp(p, key, kl) == 0) { start = p + kl; break; }
  start += 2;
  if (start > e) start = e;
  *out_s = start;
  *out_e = e;
}

// re.findall(r'\(0,\s*(\d+)\)') and re.findall(r'auto_unroll:\s*(\d+)') over one schedule.  emit(kind, value).
template <class Emit>
inline void scan_schedule(const char* s, const char* e, Emit emit) {
  static const char ukey[] = "auto_unroll:";
  const int ul = 12;
  for (const char* p = s; p < e; ++p) {
    if (*p == '(') {
      if (e - p >= 5 && p[1] == '0' && p[2] == ',') {
        const char* q = p + 3;
        while (q < e && py_space((unsigned char)*q)) ++q;
        const char* d0 = q;
        while (q < e && digit((unsigned char)*q)) ++q;
        if (q > d0 && q < e && *q == ')') {
          emit(0, d0, q);
          p = q;  // findall resumes after the match
        }
      }
    } else if (*p == 'a' && e - p > ul && memcmp(p, ukey, ul) == 0) {
      const char* q = p + ul;
      while (q < e && py_space((unsigned char)*q)) ++q;
      const char* d0 = q;
      while (q < e && digit((unsigned char)*q)) ++q;   // \d+ is greedy; no trailing literal to backtrack for
      if (q > d0) {
        emit(1, d0, q);
        p = q - 1;
      }
    }
  }
}

template <class Fn>
void parallel_rows(int64_t n, int64_t bytes, Fn fn) {
  unsigned hw = std::thread::hardware_concurrency();
  int nt = (int)std::min<int64_t>(hw ? hw : 1, std::max<int64_t>(1, bytes >> 20));  // >= 1 MB of text per thread
  nt = (int)std::min<int64_t>(nt, std::max<int64_t>(n, 1));
  if (nt <= 1) { fn(0, n); return; }
  std::vector<std::thread> th;
  for (int t = 0; t < nt; ++t) th.emplace_back(fn, n * t / nt, n * (t + 1) / nt);
  for (auto& x : th) x.join();
}

}  // namespace

// ------------------------------------------------------------------ scaler kernels
constexpr int SC_T = 256;

// partial[b, j] = sum over the block's rows of f(x[i, j]); pass 0: log1p(x); pass 1: (log1p(x) - T_j) and its square.
__global__ void scaler_colsum(const float* __restrict__ x, int64_t n, int d, int64_t rows_per_block,
                              const double* __restrict__ T, double* __restrict__ part0, double* __restrict__ part1) {
  extern __shared__ double sh[];  // [2][SC_T]
  const int cols = d, lanes = SC_T / cols;      // d <= 256
  const int j = threadIdx.x % cols, lane = threadIdx.x / cols;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = min(n, r0 + rows_per_block);
  double a = 0.0, b = 0.0;
  if (lane < lanes) {
    const double t = T ? T[j] : 0.0;
    for (int64_t i = r0 + lane; i < r1; i += lanes) {
      const double v = (double)log1pf(x[i * d + j]);
      if (T) { const double c = v - t; a += c; b += c * c; }
      else a += v;
    }
  }
  sh[threadIdx.x] = a;
  sh[SC_T + threadIdx.x] = b;
  __syncthreads();
  if (threadIdx.x < cols) {
    double sa = 0.0, sb = 0.0;
    for (int l = 0; l < lanes; ++l) { sa += sh[l * cols + threadIdx.x]; sb += sh[SC_T + l * cols + threadIdx.x]; }
    part0[(int64_t)blockIdx.x * d + threadIdx.x] = sa;
    if (part1) part1[(int64_t)blockIdx.x * d + threadIdx.x] = sb;
  }
}

// fixed-order reduction of the block partials; stage 0 -> T = sum / n; stage 1 -> var, scale (sklearn's rules)
__global__ void scaler_finish(const double* __restrict__ part0, const double* __restrict__ part1, int nb, int d, int64_t n,
                              int stage, double* __restrict__ mean, double* __restrict__ var, double* __restrict__ scale) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= d) return;
  double s0 = 0.0, s1 = 0.0;
  for (int b = 0; b < nb; ++b) { s0 += part0[(int64_t)b * d + j]; if (part1) s1 += part1[(int64_t)b * d + j]; }
  if (stage == 0) { mean[j] = s0 / (double)n; return; }
  // _incremental_mean_and_var: unnormalised variance = sum((x-T)^2) - (sum(x-T))^2 / n
  const double v = (s1 - s0 * s0 / (double)n) / (double)n;
  var[j] = v;
  // _is_constant_feature + _handle_zeros_in_scale
  const double eps = 2.220446049250313e-16, m = mean[j];
  const double ub = (double)n * eps * v + ((double)n * m * eps) * ((double)n * m * eps);
  double sc = sqrt(v);
  if (v <= ub || sc == 0.0) sc = 1.0;
  scale[j] = sc;
}

// X = float32(log1p(x)); X -= float32(mean); X /= float32(scale): StandardScaler.transform on fp32 input (scikit-learn
// >= 1.5 casts the statistics to X's dtype first; both steps are single IEEE fp32 operations, no contraction)
__global__ void scaler_transform(const float* __restrict__ x, int64_t count, int d, const double* __restrict__ mean,
                                 const double* __restrict__ scale, float* __restrict__ out) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= count) return;
  const int j = (int)(e % d);
  const float c = __fsub_rn(log1pf(x[e]), (float)mean[j]);
  out[e] = __fdiv_rn(c, (float)scale[j]);
}

}  // namespace vaex

using namespace vaex;

extern "C" {

int vaex_extent_text_scan(const char* text_host, const int64_t* str_offsets_host, int64_t n, int32_t* counts_host) {
  VAEX_CHECK(n >= 0 && (n == 0 || (text_host && str_offsets_host && counts_host)), "bad arguments");
  if (n == 0) return VAEX_OK;
  parallel_rows(n, str_offsets_host[n] - str_offsets_host[0], [=](int64_t a, int64_t b) {
    for (int64_t i = a; i < b; ++i) {
      const char *s, *e;
      schedule_span(text_host + str_offsets_host[i], text_host + str_offsets_host[i + 1], &s, &e);
      int32_t c[2] = {0, 0};
      scan_schedule(s, e, [&](int kind, const char*, const char*) { ++c[kind]; });
      counts_host[2 * i] = c[0];
      counts_host[2 * i + 1] = c[1] ? c[1] : 1;   // no auto_unroll line -> one 0.0 (extent.py:128-129)
    }
  });
  return VAEX_OK;
}

int vaex_extent_text_parse(const char* text_host, const int64_t* str_offsets_host, int64_t n, int32_t d, float* out_host,
                           int32_t* row_len_host) {
  VAEX_CHECK(n >= 0 && d > 0 && (n == 0 || (text_host && str_offsets_host && out_host)), "bad arguments");
  if (n == 0) return VAEX_OK;
  std::vector<int> overflow(1, 0);
  int* ovf = overflow.data();
  parallel_rows(n, str_offsets_host[n] - str_offsets_host[0], [=](int64_t a, int64_t b) {
    std::vector<float> un;
    for (int64_t i = a; i < b; ++i) {
      const char *s, *e;
      schedule_span(text_host + str_offsets_host[i], text_host + str_offsets_host[i + 1], &s, &e);
      float* row = out_host + i * d;
      int32_t k = 0;
      un.clear();
      scan_schedule(s, e, [&](int kind, const char* p, const char* q) {
        const float v = digits_to_f32(p, q);
        if (kind == 0) { if (k < d) row[k] = v; ++k; }
        else un.push_back(v);
      });
      if (un.empty()) un.push_back(0.0f);
      for (float v : un) { if (k < d) row[k] = v; ++k; }   // feature = extents + unrolls (extent.py:131)
      if (row_len_host) row_len_host[i] = k;
      if (k > d) { __atomic_store_n(ovf, 1, __ATOMIC_RELAXED); k = d; }
      for (int32_t j = k; j < d; ++j) row[j] = 0.0f;
    }
  });
  VAEX_CHECK(!overflow[0], "a state has more extents than d (row_len_host holds the lengths)");
  return VAEX_OK;
}

int vaex_scaler_fit(const float* raw_dev, int64_t n, int32_t d, double* mean_dev, double* var_dev, double* scale_dev,
                    double* scratch_dev, int64_t scratch_doubles, void* stream) {
  VAEX_CHECK(raw_dev && mean_dev && var_dev && scale_dev && scratch_dev, "null pointer");
  VAEX_CHECK(n > 0 && d > 0 && d <= SC_T, "need n > 0 and 0 < d <= 256");
  cudaStream_t st = (cudaStream_t)stream;
  const int lanes = SC_T / d;
  int64_t nb = std::min<int64_t>(148 * 4, cdiv(n, (int64_t)lanes * 8));
  nb = std::max<int64_t>(1, std::min<int64_t>(nb, scratch_doubles / (2 * d)));
  VAEX_CHECK(scratch_doubles >= 2 * d, "scratch needs at least 2*d doubles");
  const int64_t rpb = cdiv(n, nb);
  nb = cdiv(n, rpb);
  double* p0 = scratch_dev;
  double* p1 = scratch_dev + nb * d;
  const size_t sh = 2 * SC_T * sizeof(double);
  scaler_colsum<<<(unsigned)nb, SC_T, sh, st>>>(raw_dev, n, d, rpb, nullptr, p0, nullptr);
  VAEX_LAUNCH_CHECK();
  scaler_finish<<<(unsigned)cdiv(d, 64), 64, 0, st>>>(p0, nullptr, (int)nb, d, n, 0, mean_dev, var_dev, scale_dev);
  VAEX_LAUNCH_CHECK();
  scaler_colsum<<<(unsigned)nb, SC_T, sh, st>>>(raw_dev, n, d, rpb, mean_dev, p0, p1);
  VAEX_LAUNCH_CHECK();
  scaler_finish<<<(unsigned)cdiv(d, 64), 64, 0, st>>>(p0, p1, (int)nb, d, n, 1, mean_dev, var_dev, scale_dev);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

int vaex_scaler_transform(const float* raw_dev, int64_t n, int32_t d, const double* mean_dev, const double* scale_dev,
                          float* out_dev, void* stream) {
  VAEX_CHECK(n >= 0 && d > 0, "bad sizes");
  if (n == 0) return VAEX_OK;
  VAEX_CHECK(raw_dev && mean_dev && scale_dev && out_dev, "null pointer");
  scaler_transform<<<(unsigned)cdiv(n * d, 256), 256, 0, (cudaStream_t)stream>>>(raw_dev, n * d, d, mean_dev, scale_dev, out_dev);
  VAEX_LAUNCH_CHECK();
  return VAEX_OK;
}

}  // extern "C"
