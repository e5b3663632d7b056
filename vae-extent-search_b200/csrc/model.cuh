// Host-side model object behind vaex_model_t: fp32 master weights as loaded from the
// reference state_dict, packed device images, grow-only workspaces.
#pragma once
#include <map>
#include <string>
#include <vector>

#include "common.cuh"

namespace vaex {

struct HostTensor {
  int rows = 0, cols = 0;
  std::vector<float> v;
};

// fp16 K-major SWIZZLE_128B images of the two row-encoder layers for the tcgen05 kernel
struct TcPack {
  __half* w0 = nullptr;   // [256][K0P] as 64-wide K atoms: see segsum_tc.cu
  __half* w1 = nullptr;   // [256][256]
  float* b0 = nullptr;    // [256]
  float* b1 = nullptr;    // [256]
  __half* wl0 = nullptr;  // head: [2][128 x 256] per CTA rank
  __half* wl1 = nullptr;
  int F = 0, K0P = 0;
  bool ready = false;
  bool head_ready = false;
};

// packed weights of a fused tcgen05 MLP chain (chain_tc.cu)
struct ChainPack {
  __half* w = nullptr;
  __half* w0 = nullptr;    // layer-0 image when it lives in the operand (encoder with both latent heads)
  float* bias = nullptr;
  void* v3 = nullptr;      // Chain3Pack (chain_tc3.cu: two tiles in flight, streamed weights) when the shape fits it
  void* mc3 = nullptr;     // Mc3Pack (chain_mc3.cu: the MC-dropout passes with layer 0's accumulator kept in TMEM)
  bool v2 = false;         // chain_tc.cu's resident-weight kernel is packed (MC-dropout passes need it)
  int smem_bytes = 0;
  bool ready = false;
  alignas(8) unsigned char params_storage[384] = {};   // shape part of ChainParams
};

struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  int ensure(size_t need);
  void release();
  template <class T> T* as() { return reinterpret_cast<T*>(p); }
};

}  // namespace vaex

struct vaex_model {
  int kind = 0;
  int F = 0, H = 0, L = 0, D = 0, PH = 0, n_pred = 0;
  int feat_dim = 0;                          // extent VAE: width of the optional feature_predictor head (0: none)
  int use_norm = 0, add_sigmoid = 0;
  bool packed = false;
  std::map<std::string, vaex::HostTensor> host;
  std::map<std::string, float*> dev;   // packed fp32 device tensors, by name
  vaex::TcPack tc;
  std::map<std::string, __half*> head_img;   // fp16 images of 256-wide layers for head_tc_linear, by name (segment family)
  vaex::ChainPack dec_chain, pred_chain, enc_mu_chain, enc_lv_chain, enc_both_chain, feat_chain;   // extent-VAE decoder / cost predictor / encoder on tensor cores
  bool head_ready = false;                 // segment family: every latent / decoder / predictor layer has a head_tc image
  void* latent3 = nullptr;                 // R3Pack (reshead3.cu): norm -> l0 -> l1 -> [fc_mean; fc_logvar] in one launch
  int use_tensor = 1;                      // VAEX_OPT_TENSOR_CORES
  int use_fused_ascent = 1;                // VAEX_OPT_FUSED_ASCENT
  vaex::DevBuf ws[6];                  // activations / staging
  vaex::DevBuf topk_ws;
  vaex::DevBuf tc_ws;                  // first_prog + rowmask of the tcgen05 kernel
  struct HostSlot { float* feats = nullptr; int32_t* off = nullptr; float* scores = nullptr; size_t rows = 0, progs = 0; };
  HostSlot slots[2];                   // device staging of vaex_segsum_mlp_predict_host / _blob
  struct PinnedSlot { void* p = nullptr; size_t bytes = 0; };
  PinnedSlot pinned[2];                // host staging of vaex_segsum_mlp_predict_blob
  PinnedSlot pack_stage[4];            // fp16 staging ring of vaex_segsum_mlp_predict_host
  cudaEvent_t pack_events[4] = {nullptr, nullptr, nullptr, nullptr};   // H2D of ring slot i is over
  cudaEvent_t h2d_events[2] = {nullptr, nullptr};                      // feature H2D of device slot i is over
  int64_t pack_stats[2] = {0, 0};      // chunks sent narrowed / raw since the handle was created
  cudaStream_t copy_streams[2] = {nullptr, nullptr};
  cudaEvent_t events[4] = {nullptr, nullptr, nullptr, nullptr};

  const float* w(const std::string& name) const {
    auto it = dev.find(name);
    return it == dev.end() ? nullptr : it->second;
  }
};

namespace vaex {
// segsum_tc.cu
int tc_pack(vaex_model* m, const std::vector<float>& w0_folded, cudaStream_t st);
// head_tc.cu
int head_tc_pack(vaex_model* m, cudaStream_t st);
int head_tc_forward(vaex_model* m, const float* S, int64_t P, const uint8_t* valid, float* scores, cudaStream_t st);
// one 256-wide layer on the head kernel (head_tc.cu): Y = act(X W^T + b) [+ X] -> vector outputs or wd . Y + bd
struct HeadLinear {
  const float* X; int64_t P; int k_in; int64_t pitch;
  const __half* w_img; const float* bias256;
  bool relu, residual;
  const float *nscale, *nshift;          // optional per-column affine on X first (eval-mode BatchNorm), fp32 [256]
  float* out0; int n0; float* out1; int n1;
  float* score; const float* wd256; const float* bd;
};
int head_tc_pack_layer(const std::vector<float>& w, int out_rows, int in_cols, __half** dev, cudaStream_t st);
int head_tc_linear(const HeadLinear& a, cudaStream_t st);
// chain_tc.cu
struct ChainLayer { const std::vector<float>* w; const std::vector<float>* b; int out, in; };   // row-major [out, in] fp32 + bias
int chain_tc_pack(ChainPack& cp, const std::vector<ChainLayer>& layers, int mode, cudaStream_t st, int two_heads = 0);
void chain_tc_release(ChainPack& cp);
// reshead3.cu
int latent_heads3_pack(void** slot, const ChainLayer& l0, const ChainLayer& l1, const ChainLayer& fc, int L, cudaStream_t st);
int latent_heads3_run(void* pack, const float* S, int64_t n, const float* nscale, const float* nshift, float* mean, float* logvar,
                      cudaStream_t st);
void latent_heads3_release(void** slot);
// chain_tc3.cu
int chain3_pack(ChainPack& cp, const std::vector<ChainLayer>& layers, int mode, cudaStream_t st, int two_heads);
int chain3_run(const ChainPack& cp, const float* X, int64_t n, float* out, int32_t* extents, const float* mean, const float* scale,
               int snap_mode, cudaStream_t st, float* out2);
void chain3_release(ChainPack& cp);
// Dropout(p) after the predictor's first ReLU, `passes` times per row on shared layer-0 activations (mc_predict):
// mask uint8 [passes][n, 256] (1 = keep) or, when null, the Philox stream of dropout_kernel with seed + pass (re-seeded
// every chunk_rows rows like the SIMT path); passes == 1: out = cost, else out = mean and var_out = unbiased variance
struct ChainDropout { const uint8_t* mask; uint64_t seed; float p; int passes; int64_t chunk_rows; float* var_out; };
// chain_mc3.cu
int chain_mc3_pack(ChainPack& cp, const std::vector<ChainLayer>& layers, cudaStream_t st);
int chain_mc3_run(const ChainPack& cp, const float* X, int64_t n, float* out, const ChainDropout& drop, cudaStream_t st);
void chain_mc3_release(ChainPack& cp);
int chain_tc_run(const ChainPack& cp, const float* X, int64_t n, float* out, int32_t* extents, const float* mean,
                 const float* scale, int snap_mode, const float* wd, const float* bd, cudaStream_t st,
                 const ChainDropout* drop = nullptr, float* out2 = nullptr);
// segsum_tc.cu
// in_mode: 0 = fp32 rows, 1 = fp16 rows, 2 = the packed row image (rows_pack_image)
int segsum_tc_encode(vaex_model* m, const float* feats, const int32_t* off, int64_t n_programs,
                     int64_t n_rows, float* segsum, uint8_t* valid, cudaStream_t st, int in_mode = 0);
int64_t rows_image_bytes(int64_t n_rows);
int rows_pack_image(const float* feats, int64_t n_rows, int F, void* image, cudaStream_t st);
// ascent_cluster.cu
int ascent_cluster_run(vaex_model* m, float* z, int64_t n, int steps, float lr, float* cost, cudaStream_t st);
// ascent_compose.cu
int ascent_compose_run(vaex_model* vae, vaex_model* mlp, float* z, int64_t n, int steps, float lr, float* score, cudaStream_t st);
// host_pack.cpp
bool host_pack_supported();
bool host_pack_f16(const float* in, uint16_t* out, size_t n);   // true: a non-zero value became zero
}  // namespace vaex
