// extern "C" entry points of libvaex_b200.so (declared in include/vaex_b200.h).
#include <algorithm>
#include <atomic>
#include <cmath>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "common.cuh"
#include "model.cuh"
#include "philox.cuh"
#include "simt_kernels.cuh"

namespace vaex {

static thread_local std::string g_err;
static std::atomic<int64_t> g_launches{0};
void set_error(const std::string& msg) { g_err = msg; }
void count_launch(int n) { g_launches += n; }

int DevBuf::ensure(size_t need) {
  if (need <= bytes) return VAEX_OK;
  if (p) VAEX_CUDA(cudaFree(p));
  p = nullptr; bytes = 0;
  need = (need + (1u << 20) - 1) & ~((size_t)(1u << 20) - 1);
  VAEX_CUDA(cudaMalloc(&p, need));
  bytes = need;
  return VAEX_OK;
}
void DevBuf::release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }

static int upload(vaex_model* m, const std::string& name, const float* src, size_t count, cudaStream_t st) {
  float* d = nullptr;
  auto it = m->dev.find(name);
  if (it != m->dev.end()) { cudaFree(it->second); m->dev.erase(it); }
  VAEX_CUDA(cudaMalloc(&d, std::max<size_t>(count, 1) * sizeof(float)));
  VAEX_CUDA(cudaMemcpyAsync(d, src, count * sizeof(float), cudaMemcpyHostToDevice, st));
  VAEX_CUDA(cudaStreamSynchronize(st));  // src may be a temporary
  m->dev[name] = d;
  return VAEX_OK;
}

static int need(vaex_model* m, const char* name, int rows, int cols) {
  auto it = m->host.find(name);
  if (it == m->host.end()) { set_error(std::string("missing weight ") + name); return VAEX_E_INVALID; }
  if (it->second.rows != rows || it->second.cols != cols) {
    set_error(std::string("bad shape for ") + name + ": got [" + std::to_string(it->second.rows) + "," +
              std::to_string(it->second.cols) + "] want [" + std::to_string(rows) + "," + std::to_string(cols) + "]");
    return VAEX_E_INVALID;
  }
  return VAEX_OK;
}

static std::vector<float> transpose(const HostTensor& t) {
  std::vector<float> o((size_t)t.rows * t.cols);
  for (int r = 0; r < t.rows; ++r)
    for (int c = 0; c < t.cols; ++c) o[(size_t)c * t.rows + r] = t.v[(size_t)r * t.cols + c];
  return o;
}


static int pack_linear(vaex_model* m, const std::string& pre, int out, int in, cudaStream_t st) {
  TRY(need(m, (pre + ".weight").c_str(), out, in));
  TRY(need(m, (pre + ".bias").c_str(), 1, out));
  TRY(upload(m, pre + ".weight", m->host[pre + ".weight"].v.data(), (size_t)out * in, st));
  TRY(upload(m, pre + ".bias", m->host[pre + ".bias"].v.data(), out, st));
  return VAEX_OK;
}

// Linear followed by eval-mode BatchNorm1d: W' = diag(a) W, b' = a*b + (beta - a*mean), a = gamma/sqrt(var+eps)
static int pack_linear_bn(vaex_model* m, const std::string& lin, const std::string& bn, const std::string& out_name,
                          int out, int in, cudaStream_t st) {
  TRY(need(m, (lin + ".weight").c_str(), out, in));
  TRY(need(m, (lin + ".bias").c_str(), 1, out));
  for (const char* k : {".weight", ".bias", ".running_mean", ".running_var"}) TRY(need(m, (bn + k).c_str(), 1, out));
  std::vector<float> w = m->host[lin + ".weight"].v, b = m->host[lin + ".bias"].v;
  for (int j = 0; j < out; ++j) {
    const float a = m->host[bn + ".weight"].v[j] / std::sqrt(m->host[bn + ".running_var"].v[j] + 1e-5f);
    for (int c = 0; c < in; ++c) w[(size_t)j * in + c] *= a;
    b[j] = a * b[j] + (m->host[bn + ".bias"].v[j] - a * m->host[bn + ".running_mean"].v[j]);
  }
  TRY(upload(m, out_name + ".weight", w.data(), w.size(), st));
  TRY(upload(m, out_name + ".bias", b.data(), b.size(), st));
  m->host[out_name + ".weight"] = HostTensor{out, in, w};       // folded copies (the tensor-core images are made from them)
  m->host[out_name + ".bias"] = HostTensor{1, out, b};
  return VAEX_OK;
}

// fp16 image + zero-padded fp32 bias of one <= 256-wide layer for head_tc_linear; `rows` stacks several host tensors
// (fc_mean over fc_logvar)
static int pack_head_layer(vaex_model* m, const std::string& name, const std::vector<std::string>& rows, cudaStream_t st) {
  std::vector<float> w, b;
  int in = 0, out = 0;
  for (const std::string& r : rows) {
    const HostTensor& hw = m->host[r + ".weight"];
    const HostTensor& hb = m->host[r + ".bias"];
    if (in && hw.cols != in) { set_error("stacked layers of " + name + " differ in width"); return VAEX_E_INVALID; }
    in = hw.cols;
    out += hw.rows;
    w.insert(w.end(), hw.v.begin(), hw.v.end());
    b.insert(b.end(), hb.v.begin(), hb.v.end());
  }
  if (out > 256 || in > 256) return VAEX_OK;                     // stays on the SIMT path
  b.resize(256, 0.f);
  TRY(head_tc_pack_layer(w, out, in, &m->head_img[name], st));
  TRY(upload(m, name + ".bias256", b.data(), 256, st));
  return VAEX_OK;
}
static int pack_segment_heads(vaex_model* m, cudaStream_t st) {
  m->head_ready = false;
  if (m->H != 256 || m->L % 32 != 0 || 2 * m->L > 256) return VAEX_OK;
  TRY(pack_head_layer(m, "l0.0", {"l0.0"}, st));
  TRY(pack_head_layer(m, "l1.0", {"l1.0"}, st));
  TRY(pack_head_layer(m, "fc", {"fc_mean", "fc_logvar"}, st));
  if (m->kind == VAEX_KIND_SEGMENT_VAE) {
    TRY(pack_head_layer(m, "dec0", {"dec0"}, st));
    TRY(pack_head_layer(m, "dec1", {"dec1"}, st));
    TRY(pack_head_layer(m, "decoder.7", {"decoder.7"}, st));
  } else {
    if (m->PH != 256) return VAEX_OK;
    for (int i = 0; i < m->n_pred; ++i) TRY(pack_head_layer(m, "pred" + std::to_string(i), {"pred" + std::to_string(i)}, st));
    std::vector<float> wd = m->host["cost_predictor." + std::to_string(4 * m->n_pred) + ".weight"].v;
    wd.resize(256, 0.f);
    TRY(upload(m, "pred.wd256", wd.data(), 256, st));
  }
  m->head_ready = true;
  {
    // the three latent layers as ONE launch (reshead3.cu); the stacked fc weights are the rows of fc_mean over fc_logvar
    HostTensor& fw = m->host["fc_stack.weight"];
    HostTensor& fb = m->host["fc_stack.bias"];
    fw.rows = 2 * m->L; fw.cols = m->H;
    fw.v = m->host["fc_mean.weight"].v;
    fw.v.insert(fw.v.end(), m->host["fc_logvar.weight"].v.begin(), m->host["fc_logvar.weight"].v.end());
    fb.rows = 1; fb.cols = 2 * m->L;
    fb.v = m->host["fc_mean.bias"].v;
    fb.v.insert(fb.v.end(), m->host["fc_logvar.bias"].v.begin(), m->host["fc_logvar.bias"].v.end());
    auto lay = [&](const std::string& pre) {
      HostTensor& w = m->host[pre + ".weight"];
      return ChainLayer{&w.v, &m->host[pre + ".bias"].v, w.rows, w.cols};
    };
    TRY(latent_heads3_pack(&m->latent3, lay("l0.0"), lay("l1.0"), lay("fc_stack"), m->L, st));
  }
  return VAEX_OK;
}

static int pack_segsum_mlp(vaex_model* m, const float* norm, cudaStream_t st) {
  const int F = m->F, H = m->H;
  TRY(need(m, "segment_encoder.0.weight", H, F));
  TRY(need(m, "segment_encoder.0.bias", 1, H));
  // fold SegmentDataLoader.normalize (features /= norm_vec, mlp_model.py:95-105) into W0 columns
  std::vector<float> w0 = m->host["segment_encoder.0.weight"].v;
  if (norm)
    for (int r = 0; r < H; ++r)
      for (int c = 0; c < F; ++c) w0[(size_t)r * F + c] /= norm[c];
  TRY(upload(m, "segment_encoder.0.weight", w0.data(), w0.size(), st));
  TRY(upload(m, "segment_encoder.0.bias", m->host["segment_encoder.0.bias"].v.data(), H, st));
  TRY(pack_linear(m, "segment_encoder.2", H, H, st));
  TRY(pack_linear(m, "l0.0", H, H, st));
  TRY(pack_linear(m, "l1.0", H, H, st));
  if (m->kind == VAEX_KIND_SEGSUM_MLP) TRY(pack_linear(m, "decoder", 1, H, st));
  if (m->use_norm) {  // eval-mode BatchNorm1d -> per-column affine y = s*a + c
    for (const char* k : {"norm.weight", "norm.bias", "norm.running_mean", "norm.running_var"}) TRY(need(m, k, 1, H));
    std::vector<float> a(H), c(H);
    for (int j = 0; j < H; ++j) {
      const float inv = 1.0f / std::sqrt(m->host["norm.running_var"].v[j] + 1e-5f);
      a[j] = m->host["norm.weight"].v[j] * inv;
      c[j] = m->host["norm.bias"].v[j] - m->host["norm.running_mean"].v[j] * a[j];
    }
    TRY(upload(m, "norm.scale", a.data(), H, st));
    TRY(upload(m, "norm.shift", c.data(), H, st));
  }
  TRY(tc_pack(m, w0, st));
  if (m->kind == VAEX_KIND_SEGSUM_MLP) TRY(head_tc_pack(m, st));
  return VAEX_OK;
}

static int pack_encoder(vaex_model* m, cudaStream_t st) {
  const int D = m->D, H = m->H, L = m->L;
  TRY(pack_linear(m, "encoder.0", H, D, st));
  TRY(pack_linear(m, "encoder.2", H, H, st));
  TRY(pack_linear(m, "encoder.4", H, H, st));
  TRY(pack_linear(m, "fc_mu", L, H, st));
  TRY(pack_linear(m, "fc_logvar", L, H, st));
  return VAEX_OK;
}


static ChainLayer chain_layer(vaex_model* m, const std::string& pre) {
  HostTensor& w = m->host[pre + ".weight"];
  return ChainLayer{&w.v, &m->host[pre + ".bias"].v, w.rows, w.cols};
}

// x -> 256 -> 256 -> 256 -> mu (and the same trunk -> logvar): the trunk plus one 64-wide latent head is exactly
// the decoder's shape (156 KB of weights per CTA beside the 64 KB operand); both heads at once would not fit, so
// logvar, when asked for, is a second pass of the same kernel.
static int pack_encoder_chains(vaex_model* m, cudaStream_t st) {
  TRY(chain_tc_pack(m->enc_mu_chain, {chain_layer(m, "encoder.0"), chain_layer(m, "encoder.2"), chain_layer(m, "encoder.4"),
                                      chain_layer(m, "fc_mu")}, 0, st));
  TRY(chain_tc_pack(m->enc_lv_chain, {chain_layer(m, "encoder.0"), chain_layer(m, "encoder.2"), chain_layer(m, "encoder.4"),
                                      chain_layer(m, "fc_logvar")}, 0, st));
  // both heads in ONE pass: [fc_mu; fc_logvar] stacked as a 128-wide output layer, W0 in the operand's free atom
  if (m->L == 64) {
    HostTensor& both = m->host["fc_both.weight"];
    HostTensor& bothb = m->host["fc_both.bias"];
    both.rows = 2 * m->L; both.cols = m->H;
    both.v = m->host["fc_mu.weight"].v;
    both.v.insert(both.v.end(), m->host["fc_logvar.weight"].v.begin(), m->host["fc_logvar.weight"].v.end());
    bothb.rows = 1; bothb.cols = 2 * m->L;
    bothb.v = m->host["fc_mu.bias"].v;
    bothb.v.insert(bothb.v.end(), m->host["fc_logvar.bias"].v.begin(), m->host["fc_logvar.bias"].v.end());
    TRY(chain_tc_pack(m->enc_both_chain, {chain_layer(m, "encoder.0"), chain_layer(m, "encoder.2"), chain_layer(m, "encoder.4"),
                                          chain_layer(m, "fc_both")}, 0, st, 1));
  }
  return VAEX_OK;
}

static int pack_model(vaex_model* m, const float* norm, cudaStream_t st) {
  switch (m->kind) {
    case VAEX_KIND_SEGSUM_MLP: return pack_segsum_mlp(m, norm, st);
    case VAEX_KIND_EXTENT_VAE: {
      TRY(pack_encoder(m, st));
      TRY(pack_linear(m, "decoder.0", m->H, m->L, st));
      TRY(pack_linear(m, "decoder.2", m->H, m->H, st));
      TRY(pack_linear(m, "decoder.4", m->H, m->H, st));
      TRY(pack_linear(m, "decoder.6", m->D, m->H, st));
      TRY(chain_tc_pack(m->dec_chain, {chain_layer(m, "decoder.0"), chain_layer(m, "decoder.2"), chain_layer(m, "decoder.4"),
                                       chain_layer(m, "decoder.6")}, 0, st));
      TRY(pack_encoder_chains(m, st));
      m->feat_dim = 0;
      if (m->host.count("feature_predictor.0.weight") && m->host.count("feature_predictor.4.weight")) {   // optional head
        const int Fd = m->host["feature_predictor.4.weight"].rows;
        TRY(pack_linear(m, "feature_predictor.0", m->H, m->L, st));
        TRY(pack_linear(m, "feature_predictor.2", m->H, m->H, st));
        TRY(pack_linear(m, "feature_predictor.4", Fd, m->H, st));
        m->feat_dim = Fd;
        if (Fd <= 64)
          TRY(chain_tc_pack(m->feat_chain, {chain_layer(m, "feature_predictor.0"), chain_layer(m, "feature_predictor.2"),
                                            chain_layer(m, "feature_predictor.4")}, 0, st));
      }
      return VAEX_OK;
    }
    case VAEX_KIND_COST_PREDICTOR: {
      TRY(pack_encoder(m, st));
      TRY(pack_linear(m, "cost_predictor.0", m->PH, m->L, st));
      TRY(pack_linear(m, "cost_predictor.3", m->PH, m->PH, st));
      TRY(pack_linear(m, "cost_predictor.6", 1, m->PH, st));
      // transposed copies for the backward pass (dcost/dz)
      auto t3 = transpose(m->host["cost_predictor.3.weight"]);
      auto t0 = transpose(m->host["cost_predictor.0.weight"]);
      TRY(upload(m, "cost_predictor.3.weight_t", t3.data(), t3.size(), st));
      TRY(upload(m, "cost_predictor.0.weight_t", t0.data(), t0.size(), st));
      TRY(chain_tc_pack(m->pred_chain, {chain_layer(m, "cost_predictor.0"), chain_layer(m, "cost_predictor.3"),
                                        chain_layer(m, "cost_predictor.6")}, 1, st));
      TRY(pack_encoder_chains(m, st));
      return VAEX_OK;
    }
    case VAEX_KIND_SEGMENT_VAE: {
      TRY(pack_segsum_mlp(m, norm, st));                      // row encoder, norm, l0, l1 (+ tcgen05 images)
      TRY(pack_linear(m, "fc_mean", m->L, m->H, st));
      TRY(pack_linear(m, "fc_logvar", m->L, m->H, st));
      TRY(pack_linear_bn(m, "decoder.0", "decoder.1", "dec0", m->H, m->L, st));
      TRY(pack_linear_bn(m, "decoder.4", "decoder.5", "dec1", m->H, m->H, st));
      TRY(pack_linear(m, "decoder.7", m->H, m->H, st));
      TRY(pack_segment_heads(m, st));
      return VAEX_OK;
    }
    case VAEX_KIND_SEGMENT_COST: {
      TRY(pack_segsum_mlp(m, norm, st));
      TRY(pack_linear(m, "fc_mean", m->L, m->H, st));
      TRY(pack_linear(m, "fc_logvar", m->L, m->H, st));
      for (int i = 0; i < m->n_pred; ++i)
        TRY(pack_linear_bn(m, "cost_predictor." + std::to_string(4 * i), "cost_predictor." + std::to_string(4 * i + 1),
                           "pred" + std::to_string(i), m->PH, i == 0 ? m->L : m->PH, st));
      TRY(pack_linear(m, "cost_predictor." + std::to_string(4 * m->n_pred), 1, m->PH, st));
      TRY(pack_segment_heads(m, st));
      // the whole (BN-folded) predictor as ONE resident-weight chain: L <= 128 -> 256 x n_pred -> 1
      if (m->PH == 256 && m->n_pred >= 1 && m->n_pred <= 3 && m->L <= 128) {
        std::vector<ChainLayer> ls;
        for (int i = 0; i < m->n_pred; ++i) ls.push_back(chain_layer(m, "pred" + std::to_string(i)));
        ls.push_back(chain_layer(m, "cost_predictor." + std::to_string(4 * m->n_pred)));
        TRY(chain_tc_pack(m->pred_chain, ls, 1, st));
      }
      return VAEX_OK;
    }
  }
  set_error("unknown model kind");
  return VAEX_E_INVALID;
}

}  // namespace vaex

using namespace vaex;

extern "C" {

const char* vaex_last_error(void) { return g_err.c_str(); }
int vaex_abi_version(void) { return VAEX_ABI_VERSION; }
int64_t vaex_launch_count(void) { return g_launches.load(); }

int vaex_device_count(void) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) { set_error(std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e)); return VAEX_E_CUDA; }
  return n;
}

int vaex_model_create(int kind, const int32_t* dims, int ndims, vaex_model_t* out) {
  VAEX_CHECK(out && dims, "null argument");
  vaex_model* m = new vaex_model();
  m->kind = kind;
  int rc = VAEX_OK;
  if (kind == VAEX_KIND_SEGSUM_MLP && ndims == 4) {
    m->F = dims[0]; m->H = dims[1]; m->use_norm = dims[2]; m->add_sigmoid = dims[3];
    if (m->F < 1 || m->H < 1) rc = VAEX_E_INVALID;
  } else if (kind == VAEX_KIND_EXTENT_VAE && ndims == 3) {
    m->D = dims[0]; m->H = dims[1]; m->L = dims[2];
    if (m->D < 1 || m->H < 1 || m->L < 1) rc = VAEX_E_INVALID;
  } else if (kind == VAEX_KIND_COST_PREDICTOR && ndims == 4) {
    m->D = dims[0]; m->H = dims[1]; m->L = dims[2]; m->PH = dims[3];
    if (m->D < 1 || m->H < 1 || m->L < 1 || m->PH < 1) rc = VAEX_E_INVALID;
  } else if (kind == VAEX_KIND_SEGMENT_VAE && ndims == 3) {
    m->F = dims[0]; m->H = dims[1]; m->L = dims[2]; m->use_norm = 1;
    if (m->F < 1 || m->H < 1 || m->L < 1) rc = VAEX_E_INVALID;
  } else if (kind == VAEX_KIND_SEGMENT_COST && ndims == 5) {
    m->F = dims[0]; m->H = dims[1]; m->L = dims[2]; m->PH = dims[3]; m->n_pred = dims[4]; m->use_norm = 1;
    if (m->F < 1 || m->H < 1 || m->L < 1 || m->PH < 1 || m->n_pred < 1 || m->n_pred > 8) rc = VAEX_E_INVALID;
  } else {
    rc = VAEX_E_INVALID;
  }
  if (rc != VAEX_OK) { delete m; set_error("bad kind/dims for vaex_model_create"); return rc; }
  *out = m;
  return VAEX_OK;
}

int vaex_model_set_option(vaex_model_t h, int option, int value) {
  VAEX_CHECK(h, "null handle");
  if (option == VAEX_OPT_TENSOR_CORES) { h->use_tensor = value < 0 ? 0 : value > 2 ? 2 : value; return VAEX_OK; }
  if (option == VAEX_OPT_FUSED_ASCENT) { h->use_fused_ascent = value ? 1 : 0; return VAEX_OK; }
  set_error("unknown option");
  return VAEX_E_INVALID;
}

int vaex_model_load_weights(vaex_model_t h, const char* name, const float* data, int32_t rows, int32_t cols) {
  VAEX_CHECK(h && name && data, "null argument");
  VAEX_CHECK(rows >= 1 && cols >= 1, "bad shape");
  HostTensor t;
  t.rows = rows; t.cols = cols;
  t.v.assign(data, data + (size_t)rows * cols);
  h->host[name] = std::move(t);
  h->packed = false;
  return VAEX_OK;
}

int vaex_model_pack(vaex_model_t h, const float* fea_norm_vec_host, void* stream) {
  VAEX_CHECK(h, "null handle");
  int n = vaex_device_count();
  if (n <= 0) { if (n == 0) set_error("no CUDA device: vaex_b200 has no CPU fallback"); return VAEX_E_CUDA; }
  TRY(pack_model(h, fea_norm_vec_host, (cudaStream_t)stream));
  h->packed = true;
  return VAEX_OK;
}

int vaex_model_destroy(vaex_model_t h) {
  if (!h) return VAEX_OK;
  for (auto& kv : h->dev) cudaFree(kv.second);
  for (auto& kv : h->head_img) if (kv.second) cudaFree(kv.second);
  latent_heads3_release(&h->latent3);
  for (auto& b : h->ws) b.release();
  h->topk_ws.release();
  h->tc_ws.release();
  if (h->tc.w0) cudaFree(h->tc.w0);
  if (h->tc.w1) cudaFree(h->tc.w1);
  if (h->tc.b0) cudaFree(h->tc.b0);
  if (h->tc.b1) cudaFree(h->tc.b1);
  if (h->tc.wl0) cudaFree(h->tc.wl0);
  if (h->tc.wl1) cudaFree(h->tc.wl1);
  for (auto& sl : h->slots) { if (sl.feats) cudaFree(sl.feats); if (sl.off) cudaFree(sl.off); if (sl.scores) cudaFree(sl.scores); }
  for (ChainPack* cp : {&h->dec_chain, &h->pred_chain, &h->enc_mu_chain, &h->enc_lv_chain, &h->enc_both_chain, &h->feat_chain}) chain_tc_release(*cp);
  for (auto& pn : h->pinned) if (pn.p) cudaFreeHost(pn.p);
  for (auto& pn : h->pack_stage) if (pn.p) cudaFreeHost(pn.p);
  for (auto& s : h->copy_streams) if (s) cudaStreamDestroy(s);
  for (auto& e : h->events) if (e) cudaEventDestroy(e);
  delete h;
  return VAEX_OK;
}

}  // extern "C"

#include "abi_segsum.inc"
#include "abi_vae.inc"
#include "abi_segvae.inc"
