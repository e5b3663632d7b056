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
