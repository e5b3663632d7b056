// Shared helpers for the vaex_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "../../include/vaex_b200.h"

namespace vaex {

void set_error(const std::string& msg);
void count_launch(int n = 1);

#define VAEX_CUDA(expr)                                                                     \
  do {                                                                                      \
    cudaError_t e_ = (expr);                                                                \
    if (e_ != cudaSuccess) {                                                                \
      ::vaex::set_error(std::string(#expr) + ": " + cudaGetErrorString(e_));                \
      return VAEX_E_CUDA;                                                                   \
    }                                                                                       \
  } while (0)

#define VAEX_CHECK(cond, msg)                                                               \
  do {                                                                                      \
    if (!(cond)) {                                                                          \
      ::vaex::set_error(std::string(msg) + " (" #cond ")");                                 \
      return VAEX_E_INVALID;                                                                \
    }                                                                                       \
  } while (0)

#define VAEX_LAUNCH_CHECK()                                                                 \
  do {                                                                                      \
    ::vaex::count_launch();                                                                 \
    VAEX_CUDA(cudaGetLastError());                                                          \
  } while (0)

#define TRY(x) do { int rc_ = (x); if (rc_ != VAEX_OK) return rc_; } while (0)

static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Stream-ordered scratch for the handle-less entry points (top-k, offsets scan, k-center, KLD partials): the device's
// default pool releases freed blocks to the OS at every synchronisation unless a release threshold is set, and the
// next cudaMallocAsync then pays an OS allocation (hundreds of microseconds on a busy host).  The scratch of these
// entry points is a few KB, so keeping up to 64 MB in the pool costs nothing and makes the calls allocation-free in
// steady state.  Set once per device, on first use.
static inline cudaError_t scratch_alloc(void** p, size_t bytes, cudaStream_t st) {
  static bool kept[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && !kept[dev]) {
    cudaMemPool_t pool;
    unsigned long long keep = 64ull << 20, cur = 0;
    if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess &&
        cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &cur) == cudaSuccess && cur < keep)
      (void)cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    (void)cudaGetLastError();
    kept[dev] = true;
  }
  return cudaMallocAsync(p, bytes, st);
}

// epilogues of the SIMT linear kernel
enum Epi {
  EPI_NONE = 0,
  EPI_RELU = 1,
  EPI_RELU_RESIDUAL = 2,  // y = relu(acc+b) + aux[n,m]
  EPI_MUL_MASK = 3,       // y = (acc) * (aux[n,m] > 0)      (ReLU backward)
  EPI_ADD_AUX = 4         // y = acc + b + aux[n,m]          (residual branch of a backward pass)
};

// Y[n, m] = epi(sum_k X[n, k] * W[m, k] + b[m]);  X [N, K] (ldx), W [M, K] row-major.
int linear_simt(const float* X, int64_t ldx, const float* W, const float* b, float* Y, int64_t ldy,
                int64_t N, int M, int K, int epi, const float* aux, int64_t ldaux,
                cudaStream_t st);

}  // namespace vaex
