/* vaex_b200 -- C ABI of the B200-native hot path of Jiw00n/vae-extent-search.
 *
 * This is the drop-in boundary.  The reference has no FFI of its own
 * on this path -- it is PyTorch eager code -- so each entry point names the reference
 * *method* whose body it replaces (paths relative to the reference checkout).  The Python
 * mirror of the reference classes (vae-extent-search_b200/*.py) binds these with ctypes;
 * INTEGRATION.md shows the stub a reference maintainer would add.
 *
 * Conventions
 *   - every function returns 0 on success, a negative VAEX_E_* code on failure; the message
 *     is available from vaex_last_error() (thread-local).  No C++ exception crosses.
 *   - pointers named *_dev are device pointers owned by the caller (e.g. torch tensors);
 *     *_host are host pointers.  `stream` is a cudaStream_t passed as void* (NULL = default).
 *   - device entry points are asynchronous with respect to the host; *_host entry points are
 *     synchronous (they return when the host output buffer is filled).
 *   - a handle may be used from one thread at a time.
 *   - there is no CPU fallback: without a CUDA device every compute call returns
 *     VAEX_E_CUDA.
 */
#ifndef VAEX_B200_H_
#define VAEX_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VAEX_ABI_VERSION 1

enum {
  VAEX_OK = 0,
  VAEX_E_INVALID = -1, /* bad argument / shape / state */
  VAEX_E_CUDA = -2,    /* CUDA runtime error (message has the cudaError string) */
  VAEX_E_NOMEM = -3,
  VAEX_E_UNSUPPORTED = -4
};

/* model kinds (vaex_model_create) */
enum {
  /* SegmentSumMLPModule, python/tvm/auto_scheduler/cost_model/mlp_model.py:147-221.
   * dims = {in_dim F, hidden H (=256), use_norm(0/1), add_sigmoid(0/1)} */
  VAEX_KIND_SEGSUM_MLP = 1,
  /* VAE_feature_head, scripts/pre_experiments/model_myself/utils/model.py:6-81.
   * dims = {input_dim D, hidden H, latent L} */
  VAEX_KIND_EXTENT_VAE = 2,
  /* VAECostPredictor (extent family), utils/model.py:86-273.
   * dims = {input_dim D, hidden H, latent L, predictor_hidden PH} */
  VAEX_KIND_COST_PREDICTOR = 3,
  /* SegmentVAE, scripts/vae_experiments/models/vae.py:16-137.  dims = {input_dim F, hidden H, latent L} */
  VAEX_KIND_SEGMENT_VAE = 4,
  /* VAECostPredictor (segment family), scripts/vae_experiments/models/regression.py:11-185.
   * dims = {input_dim F, hidden H, latent L, predictor_hidden PH, predictor_layers} */
  VAEX_KIND_SEGMENT_COST = 5
};

/* flags for vaex_segsum_mlp_forward */
enum {
  VAEX_F_MASK_INVALID = 1, /* all-zero program -> -inf (MLPModel.predict, mlp_model.py:842-844) */
  VAEX_F_FORCE_SIMT = 2,   /* run the fp32 SIMT check kernels instead of the tcgen05 kernel */
  VAEX_F_FEATS_F16 = 4,    /* feats_dev holds IEEE fp16 (saturated to +-65504) instead of fp32: the tcgen05 row encoder
                              rounds to fp16 first in any case, so the scores are bit-identical; needs F % 4 == 0 */
  VAEX_F_NO_HOST_PACK = 8, /* vaex_segsum_mlp_predict_host: copy the fp32 rows as they are (no fp16 narrowing on the host) */
  VAEX_F_FEATS_IMAGE = 16  /* feats_dev is a packed row image made by vaex_rows_pack_image (fp16, 384-byte pitch, bias-one
                              columns pre-written): the row encoder fetches its operand tiles with TMA tensor loads, no
                              loader warps.  Scores are bit-identical with fp32 / fp16 rows.  F <= 174 */
};

/* snap modes for vaex_snap_extents / vaex_decode_snap */
enum {
  VAEX_SNAP_EXP = 0,   /* clamp(round(exp(x*(scale+1e-8)+mean)),1): vae_sch_extent.ipynb:1374-1383,1457 */
  VAEX_SNAP_EXPM1 = 1, /* inverse of log1p+StandardScaler: vae_extent_search.py:394-397 */
  VAEX_SNAP_DIVISOR = 4 /* OR-able: then legalise to tile factors (utils.cc:433-463 semantics) */
};

/* losses of MLPModelInternal (mlp_model.py:369-392) */
enum {
  VAEX_LOSS_MSE = 0,        /* 'rmse': torch.nn.MSELoss on sigmoid outputs */
  VAEX_LOSS_RANKNET = 1,    /* RankNetLoss, mlp_model.py:869-876 */
  VAEX_LOSS_LAMBDARANK = 2, /* LambdaRankLoss, :879-925 (the default) */
  VAEX_LOSS_LISTNET = 3     /* ListNetLoss, :928-944 */
};

typedef struct vaex_model* vaex_model_t;
typedef struct vaex_trainer* vaex_trainer_t;
typedef struct vaex_segvae_trainer* vaex_segvae_trainer_t;
typedef struct vaex_segreg_trainer* vaex_segreg_trainer_t;

const char* vaex_last_error(void);
int vaex_abi_version(void);
/* number of CUDA devices visible, or a negative error */
int vaex_device_count(void);
/* kernels launched by this library since load (for bench.py's gpu_launches) */
int64_t vaex_launch_count(void);

/* ---- model lifetime ------------------------------------------------------------------ */
int vaex_model_create(int kind, const int32_t* dims, int ndims, vaex_model_t* out);
/* `name` is the reference state_dict key (e.g. "segment_encoder.0.weight", "fc_mu.bias",
 * "norm.running_var"); data is host fp32 row-major [rows, cols] (bias: rows=1). */
int vaex_model_load_weights(vaex_model_t h, const char* name, const float* data_host,
                            int32_t rows, int32_t cols);
/* Uploads + packs weights for the kernels: folds eval-mode BatchNorm, folds 1/fea_norm_vec
 * (may be NULL) into the first layer's columns (SegmentDataLoader.normalize,
 * mlp_model.py:95-105), builds fp16 K-major swizzled images for tcgen05 and fp32 transposed
 * copies for the backward pass. */
int vaex_model_pack(vaex_model_t h, const float* fea_norm_vec_host, void* stream);
int vaex_model_destroy(vaex_model_t h);
/* per-model options.  VAEX_OPT_TENSOR_CORES (default 1): extent-VAE decode / predict_cost run the fused
 * tcgen05 chains (fp16 operands, fp32 accumulate: 1e-3 relative) when the shapes allow; 0 = fp32 SIMT
 * kernels (1e-5).  The SegmentSumMLP entry points take VAEX_F_FORCE_SIMT per call instead. */
/* Segment-VAE family: 1 also puts the latent heads (norm, l0, l1, fc_mean | fc_logvar) and the cost predictor on the tcgen05
 * head kernel, one launch per layer; the SegmentVAE *decoder* joins only at value 2 (on the shipped checkpoints its output
 * is a cancellation that fp16 operands resolve to ~7 % only). */
enum { VAEX_OPT_TENSOR_CORES = 1,
       /* (default 1) vaex_latent_ascent runs all steps in one cluster kernel (L = 64, PH = 256); 0 = one launch per layer */
       VAEX_OPT_FUSED_ASCENT = 2 };
int vaex_model_set_option(vaex_model_t h, int option, int value);

/* ---- SegmentSumMLP scoring (replaces SegmentSumMLPModule.forward, mlp_model.py:179-221,
 *      and the batch loop of MLPModelInternal._predict_a_task, :669-681) --------------------- */
/* exclusive scan of int32 sizes[n] -> int32 offsets[n+1] (SegmentDataLoader, :91-92) */
int vaex_segment_offsets(const int32_t* sizes_dev, int64_t n, int32_t* offsets_dev, void* stream);
/* Allocation-free variant of vaex_segment_offsets.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments and bit-identical output; the per-block sums of the
 * scan live in a file-scope `static __device__` array of 64 KiB that every GPU loading the library holds until
 * exit, so n is limited to VAEX_SEGMENT_OFFSETS_NOALLOC_MAX sizes (larger n: VAEX_E_INVALID, use
 * vaex_segment_offsets).  The call makes no cudaMalloc / cudaFree / cudaMallocAsync and calls no allocating library
 * routine; the first call on a device resolves the array's device address (cudaGetSymbolAddress, once per device,
 * with that device current) and should therefore be made outside stream capture.  All calls on a device share the
 * one array, which keeps the previous call's contents (nothing needs to start at zero: each call overwrites the
 * entries it reads): calls that overlap on the GPU (different streams, no ordering between them) race. */
#define VAEX_SEGMENT_OFFSETS_NOALLOC_MAX ((int64_t)16384 * 2048) /* 33 554 432 sizes */
int vaex_segment_offsets_noalloc(const int32_t* sizes_dev, int64_t n, int32_t* offsets_dev, void* stream);
/* feats_dev: fp32 [n_rows, F] row-major, rows of one program contiguous; seg_offsets_dev:
 * int32[n_programs+1]; scores_dev: fp32[n_programs]. */
int vaex_segsum_mlp_forward(vaex_model_t h, const float* feats_dev, const int32_t* seg_offsets_dev,
                            int64_t n_programs, int64_t n_rows, float* scores_dev, int flags,
                            void* stream);
/* per-program 256-d segment sums of the encoded rows (SegmentVAE.segment_sum,
 * scripts/vae_experiments/models/vae.py:68-85); segsum_dev fp32 [n_programs, H] */
int vaex_segsum_encode(vaex_model_t h, const float* feats_dev, const int32_t* seg_offsets_dev,
                       int64_t n_programs, int64_t n_rows, float* segsum_dev, int flags,
                       void* stream);
/* The packed row image: the device-resident form of SegmentDataLoader's flattened feature tensor (mlp_model.py:72-85)
 * for a pool that is scored more than once -- the active-learning loop re-scores the same candidates after every
 * re-fit (vae_extent_search.py:1720-1889), and the rows never change.  vaex_rows_image_bytes(n_rows) bytes, 128-byte
 * aligned: fp16 [n_rows, 192] (features saturated to +-65504, round to nearest even -- what the tcgen05 row encoder
 * does to fp32 rows on the fly -- plus the constant-1 columns that carry the biases inside the MMAs), then the per-row
 * "any non-zero feature" bits taken from the fp32 values (the -inf rule, mlp_model.py:842-844, stays exact).  The
 * image depends on in_dim only, not on a model.  Pass it as feats_dev with VAEX_F_FEATS_IMAGE. */
int64_t vaex_rows_image_bytes(int64_t n_rows);
int vaex_rows_pack_image(const float* feats_dev, int64_t n_rows, int32_t in_dim, void* image_dev, void* stream);
/* The residual head on its own (SegmentSumMLPModule.forward after the scatter_add, mlp_model.py:210-221): segsum_dev
 * fp32 [n_programs, H] (as written by vaex_segsum_encode) -> norm -> h += relu(l0 h) -> h += relu(l1 h) -> decoder
 * (-> sigmoid) -> scores_dev fp32 [n_programs]; valid_dev_or_null uint8 [n_programs]: 0 -> -inf.
 * flags: VAEX_F_FORCE_SIMT. */
int vaex_segsum_head(vaex_model_t h, const float* segsum_dev, int64_t n_programs, const uint8_t* valid_dev_or_null,
                     float* scores_dev, int flags, void* stream);
/* End-to-end host call (what MLPModel.predict costs a reference user, mlp_model.py:832-846):
 * host fp32 features + int32 sizes in, host fp32 scores out; chunked H2D on two streams
 * overlapped with compute.  Synchronous.  When the tcgen05 kernel serves the model, F % 4 == 0 and the host has
 * AVX2 + F16C, the host threads narrow each 128 MiB chunk to fp16 while the previous one crosses PCIe (half the
 * bytes on the bus; the kernel rounds to fp16 first in any case, so the scores do not change by one bit; a chunk
 * holding a non-zero value that would round to zero travels as fp32 so that the all-zero-program rule stays exact).
 * VAEX_F_NO_HOST_PACK switches that off. */
int vaex_segsum_mlp_predict_host(vaex_model_t h, const float* feats_host, const int32_t* sizes_host,
                                 int64_t n_programs, int64_t n_rows, float* scores_host, int flags);
/* The narrowing on its own, for callers that keep fp16 rows resident in HBM (VAEX_F_FEATS_F16): out_host[i] =
 * fp16(clamp(feats_host[i], +-65504)), round to nearest even = cvt.rn.satfinite.f16.f32.  *lossy_zero_or_null = 1 when
 * some non-zero input became a zero.  Host only (AVX2 + F16C). */
int vaex_pack_features_f16(const float* feats_host, uint16_t* out_host, int64_t count, int32_t* lossy_zero_or_null);
/* counters of a handle: which = 0 chunks vaex_segsum_mlp_predict_host / _blob sent narrowed to fp16, 1 = sent as fp32
 * (only calls large enough to start the narrowing threads count) */
int vaex_model_get_counter(vaex_model_t h, int which, int64_t* value);

/* ---- Ansor's packed per-store feature blob (SerializeFeatures, src/auto_scheduler/feature.cc:1562-1622;
 *      replaces unpack_feature, python/tvm/auto_scheduler/feature.py:49-158).  Host functions. ---- */
/* validates the blob and returns counts[5] = {records, feature rows (a record that failed lowering counts as
 * one zero row, feature.py:114-116), throughputs, task ids, min costs} */
int vaex_feature_blob_scan(const void* blob_host, int64_t nbytes, int32_t vec_len, int64_t* counts);
/* feats_host fp32 [rows, vec_len], sizes_host int32[records] (rows per record), lowered_host uint8[records]
 * (0 where the record had size 0; nullable), trailers nullable.  Same errors as the reference's assertions. */
int vaex_feature_blob_unpack(const void* blob_host, int64_t nbytes, int32_t vec_len, float* feats_host,
                             int32_t* sizes_host, uint8_t* lowered_host_or_null, float* throughputs_host_or_null,
                             int32_t* task_ids_host_or_null, float* min_costs_host_or_null);
/* MLPModel.predict from the bytes TVM's GetPerStoreFeaturesFromStates returns (mlp_model.py:832-846 with
 * feature.py:224-246): parse -> pinned staging -> chunked H2D overlapped with the kernels -> scores_host
 * [records].  The model's input width must equal the blob's vec_len (a constant task embedding is folded into
 * the first bias by the caller).  Above 64 MiB of rows the staging pass also narrows the rows to fp16 on the host
 * cores (same rule and same bits as vaex_segsum_mlp_predict_host; VAEX_F_NO_HOST_PACK keeps fp32 staging; counted
 * by vaex_model_get_counter).  Synchronous. */
int vaex_segsum_mlp_predict_blob(vaex_model_t h, const void* blob_host, int64_t nbytes, float* scores_host,
                                 int64_t n_scores, int flags);

/* ---- extent-VAE family ----------------------------------------------------------------- */
/* VAE_feature_head.encode / VAECostPredictor.encode (utils/model.py:52-56, :145-160) */
int vaex_vae_encode(vaex_model_t h, const float* x_dev, int64_t n, float* mu_dev,
                    float* logvar_dev_or_null, void* stream);
/* reparameterize (utils/model.py:58-61): z = mu + eps*exp(0.5*logvar); eps_dev may be NULL,
 * then eps ~ N(0,1) from Philox4x32-10(seed, offset). */
int vaex_reparam(const float* mu_dev, const float* logvar_dev, const float* eps_dev_or_null,
                 uint64_t seed, uint64_t offset, float* z_dev, int64_t n, int32_t latent,
                 void* stream);
/* VAE_feature_head.decode (utils/model.py:63-64) -> x_cont fp32 [n, D] */
int vaex_vae_decode(vaex_model_t h, const float* z_dev, int64_t n, float* x_dev, void* stream);
/* VAE_feature_head.predict_feature (scripts/pre_experiments/model_myself/utils/model.py:43-50,66-67): the optional
 * head z [n, L] -> relu(Linear) -> relu(Linear) -> Linear -> features [n, feature_dim], present when the state_dict
 * carries feature_predictor.{0,2,4}.  feature_dim <= 64 runs as one tcgen05 chain, wider heads on the fp32 kernels.
 * VAEX_E_INVALID when the model was packed without the head. */
int vaex_vae_predict_feature(vaex_model_t h, const float* z_dev, int64_t n, float* feat_dev, void* stream);
/* width of that head (0: none) */
int vaex_vae_feature_dim(vaex_model_t h, int32_t* feature_dim);
/* postprocess + clamp (vae_sch_extent.ipynb:1374-1383,1457) and optional divisor legalise.
 * mean/scale: fp32[D] (StandardScaler.mean_/scale_); slot_extent int32[D] (0 = free slot),
 * slot_last uint8[D] (1 = innermost length of a split) only for VAEX_SNAP_DIVISOR. */
int vaex_snap_extents(const float* x_dev, int64_t n, int32_t d, const float* mean_dev,
                      const float* scale_dev, int mode, const int32_t* slot_extent_dev,
                      const uint8_t* slot_last_dev, int32_t max_innermost, int32_t* extents_dev,
                      void* stream);
/* Divisor legalisation on its own, in place on int32 extents [n, d] (GetFactors / tile legality,
 * src/auto_scheduler/search_policy/utils.cc:433-463; SplitStep lengths must multiply to a divisor of the
 * step's extent, transform_step.cc:1015-1023).  slot_order_dev_or_null int32[n_order]: the order in which
 * the slots are walked, so that the lengths of one SplitStep are consecutive even though the printed loop
 * nest interleaves them (NULL = slot order 0..d-1); slots not listed are left as they are. */
int vaex_legalise_extents(int32_t* extents_dev, int64_t n, int32_t d, const int32_t* slot_extent_dev,
                          const uint8_t* slot_last_dev, const int32_t* slot_order_dev_or_null, int32_t n_order,
                          int32_t max_innermost, void* stream);
/* decode + snap in one call (the scoring path of BASELINE configs 3/5): z [n, L] -> continuous extents
 * x_dev_or_null fp32 [n, D] and integer extents_dev_or_null int32 [n, D]; with tensor cores enabled the
 * snap runs in the decoder kernel's epilogue (x never touches HBM unless asked for).  Arguments as in
 * vaex_vae_decode / vaex_snap_extents. */
int vaex_decode_snap(vaex_model_t h, const float* z_dev, int64_t n, float* x_dev_or_null,
                     int32_t* extents_dev_or_null, const float* mean_dev, const float* scale_dev, int mode,
                     const int32_t* slot_extent_dev, const uint8_t* slot_last_dev, int32_t max_innermost,
                     void* stream);
/* VAECostPredictor.predict_cost (utils/model.py:168-170); drop_mask_dev (uint8 [n, PH], 1=keep)
 * or drop_seed != 0 activates Dropout(p) after the first ReLU (mc_predict, :241-273). */
int vaex_predict_cost(vaex_model_t h, const float* z_dev, int64_t n, float* cost_dev,
                      const uint8_t* drop_mask_dev_or_null, uint64_t drop_seed, float p_drop,
                      void* stream);
/* VAECostPredictor.mc_predict's T dropout-active passes (utils/model.py:241-273) on latents z [n, L] (the encoder is
 * deterministic in eval mode, so the reference's T encodes are one): mean_dev fp32 [n] = mean over the passes,
 * var_dev_or_null fp32 [n] = unbiased variance (predictions.var(dim=0)).  drop_masks_dev_or_null uint8 [T][n, PH]
 * (1 = keep) replays recorded Bernoulli draws; otherwise pass t uses the Philox stream of
 * vaex_predict_cost(drop_seed + t).  With tensor cores enabled all T passes of a 256-row tile run in ONE kernel: layer
 * 0 once, its activations kept in registers, per pass one masked copy into the operand + one 256 x 256 GEMM + the
 * scalar head, mean / variance by Welford updates -- no [T, n] intermediate in HBM. */
int vaex_mc_predict_cost(vaex_model_t h, const float* z_dev, int64_t n, int32_t T, const uint8_t* drop_masks_dev_or_null,
                         uint64_t drop_seed, float p_drop, float* mean_dev, float* var_dev_or_null, void* stream);
/* select_topk_z_grad's autograd.grad(cost.sum(), z) (utils/select.py:128-135) */
int vaex_cost_grad_z(vaex_model_t h, const float* z_dev, int64_t n, float* cost_dev,
                     float* grad_z_dev, float* grad_norm_dev_or_null, void* stream);
/* select_topk_latent_diversity (utils/select.py:185-270): k-center greedy on the L2-normalised rows
 * z[cand] with z[used] as the initial centres; out_sel_dev int64[k] = the chosen entries of cand_dev in
 * selection order.  z_dev fp32 [n, L]; cand_dev int64[M]; used_dev int64[U] (U may be 0).  Returns after the
 * selection is complete (it reads one status word back). */
int vaex_kcenter_select(const float* z_dev, int64_t n, int32_t L, const int64_t* cand_dev, int64_t M,
                        const int64_t* used_dev, int64_t U, int32_t k, float eps, int64_t* out_sel_dev,
                        void* stream);
/* Farthest-point sampling on the RAW rows z[cand] (no row normalisation) from a given first pick: out_sel_dev[0] =
 * cand[first_pos]; every further pick is the candidate with the largest minimum Euclidean distance to the picks so far
 * (lowest candidate position on ties).  Replaces the host loops of select_features
 * (scripts/pre_experiments/for_inference/tuning_improved.py:138-170, first pick = the best acquisition score) and of
 * select_init_latent_diversity (scripts/pre_experiments/model_myself/utils/select.py:413-428, first pick random).
 * cand_dev_or_null int64[M] (null: rows [0, M) of z).  Returns after the selection is complete. */
int vaex_fps_select(const float* z_dev, int64_t n, int32_t L, const int64_t* cand_dev_or_null, int64_t M, int64_t first_pos,
                    int32_t k, int64_t* out_sel_dev, void* stream);
/* out_d_dev[j] = min_u ||zc_j - zu_u||_2 over raw rows (+inf when U = 0): the distance of every unlabeled latent to the
 * labeled set, tuning_improved.py:96-112 (which forms it through ||a||^2 + ||b||^2 - 2 a.b; here the direct difference). */
int vaex_latent_min_dist(const float* zc_dev, int64_t M, const float* zu_dev, int64_t U, int32_t L, float* out_d_dev,
                         void* stream);
/* The 164-d per-statement feature rows of a fixed sketch, evaluated from the loop extents of n candidates:
 * what GetPerStoreFeaturesWorkerFunc + PerStoreFeatureExtractor + GetPerStoreFeature (src/auto_scheduler/feature.cc:1296-1368,
 * :609-1048, :1053-1188) produce by lowering each candidate with TVM on the CPU, here from a template of the sketch's
 * lowered loop nest.  program_dev: the int32 program of vae-extent-search_b200/sketch_features.py SketchTemplate.compile()
 * (layout in csrc/sketch_features.cu); extents_dev int32 [n, slots]; rows_dev fp32 [n * rows, 164], candidate-major, i.e.
 * a SegmentSum-MLP batch whose segments all have `rows` statements.  The first call with a given program (device address,
 * length) reads the program back and checks every offset, count and id in it (one sync; VAEX_E_INVALID for a damaged program); later calls launch without touching the host, so the call is
 * capturable into a CUDA graph after a warm-up call.  A buffer that is reused for a DIFFERENT program must not keep its
 * address and length (the Python wrapper allocates one tensor per compiled template). */
int vaex_sketch_features(const int32_t* program_dev, int32_t program_words, const int32_t* extents_dev, int64_t n,
                         float* rows_dev, void* stream);
/* The same rows narrowed to IEEE fp16 with cvt.rn.satfinite (rows_f16_dev [n * rows, 164] halves): the input the tcgen05 row
 * encoder takes under VAEX_F_FEATS_F16 -- it rounds fp32 rows to fp16 in exactly this way before its first GEMM, so the
 * scores are bit-identical while half the bytes are written and read. */
int vaex_sketch_features_f16(const int32_t* program_dev, int32_t program_words, const int32_t* extents_dev, int64_t n,
                             uint16_t* rows_f16_dev, void* stream);
/* multi-step ascent z += lr * dcost/dz (new surface, BASELINE config 3) */
int vaex_latent_ascent(vaex_model_t h, float* z_inout_dev, int64_t n, int32_t steps, float lr,
                       float* cost_dev, void* stream);

/* The north-star ascent graph (BASELINE.json north_star (3), config 3): score(z) = head(decoder_trunk(z)) -- the decoder's
 * three hidden layers (VAE_feature_head.decoder.{0,2,4}, utils/model.py:31-39) feeding the SegmentSumMLP's residual head
 * (l0, l1, decoder; mlp_model.py:210-221) on the 256-d vector -- and `steps` plain ascent steps z += lr * d score / d z
 * with the fused backward pass through both models, all steps in one cluster kernel (eight CTAs per sixteen restarts,
 * fp32).  vae: kind VAEX_KIND_EXTENT_VAE (L = 64, H = 256); mlp: kind VAEX_KIND_SEGSUM_MLP (H = 256, no norm / sigmoid).
 * z_inout_dev fp32 [n, 64], score_dev fp32 [n] = the score of the final z.  The single-step precedent in the reference is
 * select_topk_z_grad (utils/select.py:128-135). */
int vaex_latent_ascent_decoder_head(vaex_model_t vae, vaex_model_t mlp, float* z_inout_dev, int64_t n, int32_t steps,
                                    float lr, float* score_dev, void* stream);

/* ---- segment-VAE family (eval mode: BatchNorm folded to affines, Dropout = identity) -------- */
/* SegmentVAE.forward steps 1-3 (vae.py:113-128) / VAECostPredictor.encode (regression.py:82-101):
 * row encoder -> segment sum -> norm -> residual l0, l1 -> fc_mean, fc_logvar.
 * mean_dev, logvar_dev fp32 [n_programs, L]; segsum_dev_or_null fp32 [n_programs, H] receives the
 * raw segment sums (the reference returns them as `segment_sum_vec`).  Kinds 4 and 5. */
int vaex_segvae_encode(vaex_model_t h, const float* feats_dev, const int32_t* seg_offsets_dev,
                       int64_t n_programs, int64_t n_rows, float* mean_dev, float* logvar_dev_or_null,
                       float* segsum_dev_or_null, int flags, void* stream);
/* SegmentVAE.decode (vae.py:109-111): z [n, L] -> reconstructed segment-sum vector [n, H].  Kind 4. */
int vaex_segvae_decode(vaex_model_t h, const float* z_dev, int64_t n, float* recon_dev, void* stream);
/* VAECostPredictor.predict_cost (regression.py:109-111): z [n, L] -> cost [n].  Kind 5. */
int vaex_segcost_predict(vaex_model_t h, const float* z_dev, int64_t n, float* cost_dev, void* stream);
/* T dropout-active passes of the segment cost predictor on latents z [n, L] (the model API select_features calls,
 * tuning_improved.py:89: model.mc_predict(seg_sizes, feats, T, use_mean=True) -- the segment VAECostPredictor of
 * scripts/vae_experiments/models/regression.py:49-60 has Dropout(p) after every predictor block but the last; the
 * method itself is absent from the reference, its extent-family namesake is utils/model.py:241-273).  mean_dev fp32 [n],
 * var_dev_or_null fp32 [n] = unbiased variance over the passes.  drop_masks_dev_or_null uint8
 * [T][predictor_layers - 1][n, PH] (1 = keep) replays recorded draws; otherwise pass t, dropout layer j uses the Philox
 * stream drop_seed + t (predictor_layers - 1) + j over the elements of the [n, PH] activation matrix.  Block 0 is
 * computed once; per pass one dropout + one single-layer launch per remaining block (tcgen05 head kernel when enabled). */
int vaex_segcost_mc_predict(vaex_model_t h, const float* z_dev, int64_t n, int32_t T, const uint8_t* drop_masks_dev_or_null,
                            uint64_t drop_seed, float p_drop, float* mean_dev, float* var_dev_or_null, void* stream);

/* ---- training of the SegmentSumMLP (MLPModelInternal._fit_a_model / _fine_tune_a_model,
 *      mlp_model.py:512-600,609-651).  Parameters, gradients and the Adam moments are flat fp32
 *      device buffers owned by the caller, laid out in state_dict order (vaex_trainer_param_offset);
 *      data-parallel training all-reduces `grads` between the two calls. ------------------------ */
int vaex_trainer_create(int32_t in_dim, int32_t hidden, int32_t add_sigmoid, vaex_trainer_t* out);
int vaex_trainer_destroy(vaex_trainer_t t);
int64_t vaex_trainer_param_count(vaex_trainer_t t);
/* offset / element count of a reference state_dict key inside the flat buffers */
int vaex_trainer_param_offset(vaex_trainer_t t, const char* name, int64_t* offset, int64_t* count);
/* shuffled mini-batch assembly on the device (SegmentDataLoader._fetch_indices, :129-141): packs the
 * rows (and labels) of programs indices_dev[0..n) of a resident data set; out_offsets_dev int32[n+1],
 * out_feats_dev fp32 [sum of their sizes, F] (the caller knows that sum from its host copy of the sizes). */
int vaex_gather_segments(const float* feats_dev, int32_t in_dim, const int32_t* seg_offsets_dev,
                         const int64_t* indices_dev, int64_t n, int32_t* out_offsets_dev,
                         float* out_feats_dev, const float* labels_dev_or_null,
                         float* out_labels_dev_or_null, void* stream);
/* one batch (<= 4096 programs): forward, loss (VAEX_LOSS_*) and the full backward pass.
 * grads_dev is overwritten; *loss_dev receives the loss; preds_dev_or_null the predictions. */
int vaex_trainer_forward_backward(vaex_trainer_t t, const float* params_dev, float* grads_dev,
                                  const float* feats_dev, const int32_t* seg_offsets_dev,
                                  const float* labels_dev, int64_t n_programs, int64_t n_rows,
                                  int loss_type, float* loss_dev, float* preds_dev_or_null, void* stream);
/* loss value (and optionally d loss / d pred) of n <= 4096 predictions (_validate, :653-661) */
int vaex_loss_eval(int loss_type, const float* preds_dev, const float* labels_dev, int64_t n,
                   float* loss_dev, float* dpred_dev_or_null, void* stream);
/* g = grad_scale * grads; clip_grad_norm_(clip_norm) if clip_norm > 0 (pre-clip norm -> *gnorm_dev);
 * torch.optim.Adam step `step` (1-based) with L2 weight decay, in place. */
int vaex_adam_step(float* params_dev, const float* grads_dev, float* m_dev, float* v_dev, int64_t n,
                   int64_t step, float lr, float beta1, float beta2, float eps, float weight_decay,
                   float clip_norm, float grad_scale, float* gnorm_dev_or_null, void* stream);
/* torch.optim.SGD step without momentum (fine-tuning, :626) */
int vaex_sgd_step(float* params_dev, const float* grads_dev, int64_t n, float lr, float weight_decay,
                  float grad_scale, void* stream);

/* One epoch of the loop above in a single call (MLPModelInternal._fit_a_model's inner loop, mlp_model.py:545-570; the SGD
 * fine-tuning loop :628-640 with optimizer = 1): for every batch of `batch` (<= 4096) programs order[s : s + batch) of the
 * device-resident set -- vaex_gather_segments, vaex_trainer_forward_backward, vaex_allreduce_grads over comm_or_null (data
 * parallel: every rank passes its own shard and order; the mean gradient is applied), clip_grad_norm_ + Adam (optimizer 0;
 * betas 0.9 / 0.999, eps 1e-8, L2 weight decay) or SGD (1) -- and the reference's moving-average loss
 * (avg = 0.95 avg + 0.05 loss; the first batch initialises it when first_batch_initialises_avg != 0) in *avg_loss_dev.
 * sizes_host int32 [programs of the set] (rows per program, the host copy of the loader's segment_sizes), order_host /
 * order_dev: the same int64 permutation on both sides.  *step_inout is Adam's step counter.  Asynchronous on `stream`
 * except for host-side buffer growth. */
typedef struct vaex_comm* vaex_comm_t;
int vaex_trainer_epoch(vaex_trainer_t t, vaex_comm_t comm_or_null, float* params_dev, float* grads_dev, float* m_dev,
                       float* v_dev, const float* feats_dev, const int32_t* seg_offsets_dev, const int32_t* sizes_host,
                       const float* labels_dev, const int64_t* order_host, const int64_t* order_dev, int64_t n,
                       int32_t batch, int loss_type, int optimizer, float lr, float weight_decay, float clip_norm,
                       int64_t* step_inout, float* avg_loss_dev, int first_batch_initialises_avg, int64_t* batches_out,
                       void* stream);

/* ---- training of the extent-family VAECostPredictor (train_regression, vae_extent_search.py:1413-1440; losses
 *      utils/training.py:133-226) ---------------------------------------------------------------------------- */
/* pair_loss_fn (training.py:144-166): mean over i < j of max(0, -sign(t_j - t_i) (p_j - p_i) + margin); n <= 4096 */
int vaex_pair_loss(const float* pred_dev, const float* target_dev, int64_t n, float margin, float* loss_dev,
                   float* dpred_dev_or_null, void* stream);
/* kld_loss_fn (training.py:181-186): -0.5 mean(1 + logvar - mu^2 - exp(logvar)) over `count` elements */
int vaex_kld_loss(const float* mu_dev, const float* logvar_dev, int64_t count, float* loss_dev, float* dmu_dev_or_null,
                  float* dlogvar_dev_or_null, void* stream);
typedef struct vaex_reg_trainer* vaex_reg_trainer_t;
/* flat fp32 parameter buffer in state_dict order of VAECostPredictor(input D, hidden H, latent L, predictor_hidden PH,
 * predictor_layers 2) (utils/model.py:108-131); "@encoder_group" / "@predictor_group" name the two AdamW groups */
int vaex_reg_trainer_create(int32_t D, int32_t H, int32_t L, int32_t PH, vaex_reg_trainer_t* out);
int vaex_reg_trainer_destroy(vaex_reg_trainer_t t);
int64_t vaex_reg_trainer_param_count(vaex_reg_trainer_t t);
int vaex_reg_trainer_param_offset(vaex_reg_trainer_t t, const char* name, int64_t* offset, int64_t* count);
/* One step's forward(use_mean=True) + compute_total_loss + backward.  x_dev [n, D], labels_dev [n], n <= 4096;
 * drop_mask_dev uint8 [n, PH] (1 = keep) makes the main forward train-mode (NULL = eval); noise_dev [n, L] is
 * noise_std * randn of smooth_loss_fn (NULL = 0); cfg_host[6] = {lambda_reg, lambda_pair, margin, gamma, beta, p_drop};
 * losses_dev[5] = {total, reg, pair, smooth, kld}; grads_dev is overwritten; cost_dev_or_null [n] = cost_pred. */
int vaex_reg_trainer_forward_backward(vaex_reg_trainer_t t, const float* params_dev, float* grads_dev, const float* x_dev,
                                      const float* labels_dev, int64_t n, const uint8_t* drop_mask_dev_or_null,
                                      const float* noise_dev_or_null, const float* cfg_host, float* losses_dev,
                                      float* cost_dev_or_null, void* stream);
/* VAE pre-training (vae_train's inner loop + vae_feature_loss, utils/training.py:14-32,70-88): flat parameter buffer in
 * state_dict order of VAE_feature_head(input D, latent L, hidden H) without the feature head (utils/model.py:16-50). */
typedef struct vaex_vae_trainer* vaex_vae_trainer_t;
int vaex_vae_trainer_create(int32_t D, int32_t H, int32_t L, vaex_vae_trainer_t* out);
int vaex_vae_trainer_destroy(vaex_vae_trainer_t t);
int64_t vaex_vae_trainer_param_count(vaex_vae_trainer_t t);
int vaex_vae_trainer_param_offset(vaex_vae_trainer_t t, const char* name, int64_t* offset, int64_t* count);
/* forward(x, use_mean = (eps_dev == NULL)) + alpha_recon * mse(x_recon, x) + beta * kl + backward; eps_dev [n, L] is the
 * randn of reparameterize (utils/model.py:58-61); losses_dev[3] = {total, recon, kl}; n <= 4096.  The optimizer is
 * vaex_adam_step (torch.optim.Adam, no clipping, :65). */
int vaex_vae_trainer_forward_backward(vaex_vae_trainer_t t, const float* params_dev, float* grads_dev, const float* x_dev,
                                      int64_t n, const float* eps_dev_or_null, float alpha_recon, float beta,
                                      float* losses_dev, void* stream);
/* ---- SegmentVAE training step (segment family; VAE_Trainer.train_vae, scripts/vae_experiments/trainer.py:23-118) ----
 * Flat fp32 parameter / gradient buffers in SegmentVAE.parameters() order (scripts/vae_experiments/models/vae.py:32-66);
 * vaex_segvae_trainer_param_offset maps a state_dict key to its slice. */
int vaex_segvae_trainer_create(int32_t F, int32_t H, int32_t L, vaex_segvae_trainer_t* out);
int vaex_segvae_trainer_destroy(vaex_segvae_trainer_t t);
int64_t vaex_segvae_trainer_param_count(vaex_segvae_trainer_t t);
int vaex_segvae_trainer_param_offset(vaex_segvae_trainer_t t, const char* name, int64_t* offset, int64_t* count);
/* SegmentVAE.forward in train mode (vae.py:113-137: batch-statistics BatchNorm x3, decoder Dropout, reparameterisation)
 * + vae_loss (trainer.py:270-280: sum-reduced MSE against the LIVE segment sums + beta * KL sum) + backward.
 * feats_dev fp32 [n_rows, F], seg_offsets_dev int32 [n_programs + 1], 2 <= n_programs <= 4096.
 * eps_dev_or_null fp32 [n_programs, L]: the reparameterisation noise (NULL = use_mean: logvar := 0, z = mean).
 * Dropout: drop_mask_dev_or_null uint8 [n_programs, H] (1 = keep), else drop_seed != 0 (Philox), else none.
 * bn_running_dev_or_null fp32 [3][2][H]: running_mean / running_var of norm, decoder.1, decoder.5, updated in place
 * (momentum 0.1, unbiased variance).  losses_dev fp32 [3] = {total, recon, kld}.  mean / recon outputs optional. */
int vaex_segvae_trainer_forward_backward(vaex_segvae_trainer_t t, const float* params_dev, float* grads_dev,
                                         const float* feats_dev, const int32_t* seg_offsets_dev, int64_t n_programs,
                                         int64_t n_rows, const float* eps_dev_or_null, const uint8_t* drop_mask_dev_or_null,
                                         uint64_t drop_seed, float p_drop, float beta, float* bn_running_dev_or_null,
                                         float* losses_dev, float* mean_out_dev_or_null, float* recon_out_dev_or_null,
                                         void* stream);
/* ---- segment-family VAECostPredictor training step (Regression_Trainer, scripts/vae_experiments/trainer.py:296-568) ----
 * Flat buffers in VAECostPredictor.parameters() order (scripts/vae_experiments/models/regression.py:34-61);
 * vaex_segreg_trainer_param_offset maps a state_dict key, "@encoder_group" or "@predictor_group" to its slice. */
int vaex_segreg_trainer_create(int32_t F, int32_t H, int32_t L, int32_t PH, int32_t n_layers, vaex_segreg_trainer_t* out);
int vaex_segreg_trainer_destroy(vaex_segreg_trainer_t t);
int64_t vaex_segreg_trainer_param_count(vaex_segreg_trainer_t t);
int vaex_segreg_trainer_param_offset(vaex_segreg_trainer_t t, const char* name, int64_t* offset, int64_t* count);
/* forward(use_mean=True) in train mode + compute_total_loss (trainer.py:538-556: mse + lambda_pair * pair + gamma * smooth
 * + beta * kld) + backward.  smooth_loss_fn runs predict_cost twice more in train mode, so a step is three predictor passes
 * (z, z, z + noise), each with its own batch statistics, dropout draws and running-statistics update, all live in z.
 * cfg_host fp32 [5] = {lambda_pair, margin, gamma, beta, p_drop}; noise_dev fp32 [P, L] = noise_std * randn;
 * drop_masks_dev_or_null uint8 [3][n_layers - 1][P, PH] (1 = keep), else drop_seed != 0 (Philox), else no dropout;
 * bn_running_dev_or_null fp32 = norm {mean, var}[H] then per predictor block {mean, var}[PH];
 * losses_dev fp32 [5] = {total, reg, pair, smooth, kld}; cost_dev_or_null fp32 [P]. */
int vaex_segreg_trainer_forward_backward(vaex_segreg_trainer_t t, const float* params_dev, float* grads_dev,
                                         const float* feats_dev, const int32_t* seg_offsets_dev, const float* labels_dev,
                                         int64_t n_programs, int64_t n_rows, const float* noise_dev,
                                         const uint8_t* drop_masks_dev_or_null, uint64_t drop_seed, const float* cfg_host,
                                         float* bn_running_dev_or_null, float* losses_dev, float* cost_dev_or_null,
                                         void* stream);
/* torch.optim.AdamW on the group params[off, off + cnt); the clip coefficient uses the norm of the whole gradient
 * buffer (clip_grad_norm_ over all parameters), recomputed into gnorm_dev when compute_norm != 0 */
int vaex_adamw_step(float* params_dev, const float* grads_dev, float* m_dev, float* v_dev, int64_t n_total, int64_t off,
                    int64_t cnt, int64_t step, float lr, float beta1, float beta2, float eps, float weight_decay,
                    float clip_norm, float grad_scale, float* gnorm_dev_or_null, int compute_norm, void* stream);

/* ---- extent parser + scaler ------------------------------------------------------ */
/* state_to_records' numeric part (scripts/pre_experiments/model_myself/utils/extent.py:94-136, inlined again at
 * vae_extent_search.py:165-176): text_host holds n printed Ansor states back to back, state i =
 * bytes [str_offsets_host[i], str_offsets_host[i+1]).  Per state: schedule = strip(state) after its last
 * "Placeholder" + 2 characters; every match of \(0,\s*(\d+)\) is an extent, every match of auto_unroll:\s*(\d+) an
 * unroll value (one 0.0 when there is none); the row is extents followed by unrolls.  Host only, all host threads.
 * scan: counts_host int32 [n, 2] = (#extents, #unrolls incl. the default). */
int vaex_extent_text_scan(const char* text_host, const int64_t* str_offsets_host, int64_t n, int32_t* counts_host);
/* parse: out_host fp32 [n, d] (rows zero-padded to d; a pinned buffer makes the H2D copy asynchronous),
 * row_len_host_or_null int32[n] = values found per state.  VAEX_E_INVALID if a state holds more than d values
 * (its first d are still written). */
int vaex_extent_text_parse(const char* text_host, const int64_t* str_offsets_host, int64_t n, int32_t d,
                           float* out_host, int32_t* row_len_host_or_null);
/* np.log1p + sklearn StandardScaler.fit (vae_extent_search.py:394-397; sklearn's _incremental_mean_and_var and
 * _handle_zeros_in_scale rules) on raw extents raw_dev fp32 [n, d], d <= 256: mean_dev, var_dev, scale_dev fp64 [d].
 * Deterministic (block partials reduced in a fixed order).  scratch_dev: at least 2*d*592 doubles. */
int vaex_scaler_fit(const float* raw_dev, int64_t n, int32_t d, double* mean_dev, double* var_dev, double* scale_dev,
                    double* scratch_dev, int64_t scratch_doubles, void* stream);
/* StandardScaler.transform(np.log1p(raw)) with sklearn's rounding (statistics cast to fp32, one fp32 subtract,
 * one fp32 divide): out_dev fp32 [n, d].  The inverse is vaex_snap_extents(VAEX_SNAP_EXPM1). */
int vaex_scaler_transform(const float* raw_dev, int64_t n, int32_t d, const double* mean_dev, const double* scale_dev,
                          float* out_dev, void* stream);

/* ---- top-k (util_select_topk, utils/select.py:54-81) -------------------------------------- */
/* k largest of scores[n] restricted to mask (uint8, 1 = eligible; NULL = all), ordered by
 * (score desc, index asc).  out_vals fp32[k], out_idx int64[k]; *n_out_dev (int32) = number
 * written (min(k, eligible)).  k <= 1024. */
int vaex_topk(const float* scores_dev, int64_t n, int32_t k, const uint8_t* mask_dev_or_null,
              float* out_vals_dev, int64_t* out_idx_dev, int32_t* n_out_dev, void* stream);

/* ---- multi-GPU: one process per GPU, NCCL over NVLink / NVSwitch.  The reference is single-GPU (its
 *      selection sorts one numpy array, utils/select.py:54-81; its training loop is one process, mlp_model.py:512-600);
 *      these entry points are what replaces that sort / that optimizer step when the pool or the batch is sharded.
 *      NCCL is bound at run time (dlopen libnccl.so.2; VAEX_NCCL_LIB overrides the path); world == 1 needs none. ---- */
typedef struct vaex_comm* vaex_comm_t;
#define VAEX_NCCL_ID_BYTES 128
/* rank 0: a fresh ncclUniqueId (128 bytes) that the host side hands to every rank (torch.distributed store, MPI, a file) */
int vaex_nccl_unique_id(uint8_t* id_out_host);
/* with the rank's device current; id_host may be NULL when world == 1 */
int vaex_nccl_init(int32_t rank, int32_t world, const uint8_t* id_host, vaex_comm_t* out);
/* The collectives below may be stream-captured into a CUDA graph; NCCL then keeps the communicator alive for the
 * graph, and ncclCommDestroy waits for it: destroy (or reset) every such graph BEFORE vaex_nccl_destroy, or the call
 * does not return (seen: a 2-rank bench that held its step graphs at exit).  bench.py keeps NCCL work out of its graph. */
int vaex_nccl_destroy(vaex_comm_t c);
int vaex_comm_info(vaex_comm_t c, int32_t* rank_or_null, int32_t* world_or_null);
/* Candidate merge of the sharded scoring path: local_vals_dev fp32[k] / local_idx_dev int64[k] are this rank's winners in
 * vaex_topk order (score desc, index asc; rank-local indices), *n_local_dev_or_null how many of them are real (NULL = k);
 * index_base = global index of the rank's first program (ranks own ascending contiguous blocks).  One ncclAllGather of
 * k 16-byte records per rank + one single-CTA merge: out_vals_dev fp32[k], out_idx_dev int64[k] = the k best over all
 * ranks in (score desc, global index asc) order, identical on every rank; entries past *n_out are (-inf, -1).
 * k <= 1024 and world * k <= 8192.  Asynchronous on `stream`. */
int vaex_topk_allgather_merge(vaex_comm_t c, const float* local_vals_dev, const int64_t* local_idx_dev,
                              const int32_t* n_local_dev_or_null, int32_t k, int64_t index_base, float* out_vals_dev,
                              int64_t* out_idx_dev, int32_t* n_out_dev_or_null, void* stream);
/* data-parallel training: sum all-reduce of the flat fp32 gradient buffer, in place (the 1 / world factor is the
 * grad_scale argument of vaex_adam_step / vaex_adamw_step / vaex_sgd_step) */
int vaex_allreduce_grads(vaex_comm_t c, float* grads_dev, int64_t n, void* stream);
/* recv_dev [world][bytes_per_rank] <- every rank's send_dev (the final exchange of the sharded latent ascent) */
int vaex_allgather(vaex_comm_t c, const void* send_dev, void* recv_dev, int64_t bytes_per_rank, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VAEX_B200_H_ */
