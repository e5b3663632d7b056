#pragma once
#include "common.cuh"

namespace vaex {

int rowdot(const float* X, int64_t ldx, const float* w, const float* b, float* y, int64_t N, int K,
           int sigmoid, const uint8_t* valid, cudaStream_t st);
int segsum_simt(const float* rows, int64_t row_base, const int32_t* off, int64_t p0, int64_t P,
                int H, float* out, cudaStream_t st);
int program_valid(const float* feats, const int32_t* off, int64_t P, int F, uint8_t* valid,
                  cudaStream_t st);

// elementwise.cu
int exclusive_scan_i32(const int32_t* sizes, int64_t n, int32_t* offsets, cudaStream_t st);
int reparam(const float* mu, const float* logvar, const float* eps, uint64_t seed, uint64_t offset,
            float* z, int64_t count, cudaStream_t st);
int snap_extents(const float* x, int64_t n, int d, const float* mean, const float* scale, int mode,
                 const int32_t* slot_extent, const uint8_t* slot_last, int max_innermost,
                 int32_t* out, cudaStream_t st);
int legalise_extents(int32_t* ext, int64_t n, int d, const int32_t* slot_extent, const uint8_t* slot_last, int max_innermost,
                     cudaStream_t st, const int32_t* order = nullptr, int n_order = 0);
int dropout_apply(float* h, int64_t n, int cols, const uint8_t* mask, uint64_t seed, float p,
                  cudaStream_t st);
int relu_mask_rowvec(const float* h, const float* w, float* out, int64_t n, int cols, cudaStream_t st);
int axpy(float* z, const float* g, float lr, int64_t count, cudaStream_t st);
int row_l2norm(const float* g, int64_t n, int cols, float* out, cudaStream_t st);

// topk.cu
size_t topk_workspace_bytes();
int topk(const float* scores, int64_t n, int k, const uint8_t* mask, float* out_vals,
         int64_t* out_idx, int32_t* n_out, void* workspace, cudaStream_t st);

}  // namespace vaex
