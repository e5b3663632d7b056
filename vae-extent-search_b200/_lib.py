"""ctypes binding of libvaex_b200.so (include/vaex_b200.h).  There is no CPU fallback: if the
library is missing or no CUDA device is present, calls fail loudly."""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvaex_b200.so")

KIND_SEGSUM_MLP, KIND_EXTENT_VAE, KIND_COST_PREDICTOR, KIND_SEGMENT_VAE, KIND_SEGMENT_COST = 1, 2, 3, 4, 5
F_MASK_INVALID, F_FORCE_SIMT, F_FEATS_F16, F_NO_HOST_PACK, F_FEATS_IMAGE = 1, 2, 4, 8, 16
OPT_TENSOR_CORES, OPT_FUSED_ASCENT = 1, 2
SNAP_EXP, SNAP_EXPM1, SNAP_DIVISOR = 0, 1, 4
LOSS_IDS = {"rmse": 0, "rankNetLoss": 1, "lambdaRankLoss": 2, "listNetLoss": 3}

c_i32p = ctypes.POINTER(ctypes.c_int32)
c_f32p = ctypes.POINTER(ctypes.c_float)
vp, i64, i32, u64, f32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32, ctypes.c_uint64, ctypes.c_float

# name -> argtypes; every symbol declared in include/vaex_b200.h (tests check the export list)
SIGNATURES = {
    "vaex_last_error": ([], ctypes.c_char_p),
    "vaex_abi_version": ([], ctypes.c_int),
    "vaex_device_count": ([], ctypes.c_int),
    "vaex_launch_count": ([], i64),
    "vaex_model_create": ([ctypes.c_int, c_i32p, ctypes.c_int, ctypes.POINTER(vp)], ctypes.c_int),
    "vaex_model_load_weights": ([vp, ctypes.c_char_p, vp, i32, i32], ctypes.c_int),
    "vaex_model_pack": ([vp, vp, vp], ctypes.c_int),
    "vaex_model_destroy": ([vp], ctypes.c_int),
    "vaex_model_set_option": ([vp, ctypes.c_int, ctypes.c_int], ctypes.c_int),
    "vaex_segment_offsets": ([vp, i64, vp, vp], ctypes.c_int),
    "vaex_segment_offsets_noalloc": ([vp, i64, vp, vp], ctypes.c_int),
    "vaex_segsum_mlp_forward": ([vp, vp, vp, i64, i64, vp, ctypes.c_int, vp], ctypes.c_int),
    "vaex_segsum_encode": ([vp, vp, vp, i64, i64, vp, ctypes.c_int, vp], ctypes.c_int),
    "vaex_rows_image_bytes": ([i64], i64),
    "vaex_rows_pack_image": ([vp, i64, i32, vp, vp], ctypes.c_int),
    "vaex_segsum_head": ([vp, vp, i64, vp, vp, ctypes.c_int, vp], ctypes.c_int),
    "vaex_segsum_mlp_predict_host": ([vp, vp, vp, i64, i64, vp, ctypes.c_int], ctypes.c_int),
    "vaex_pack_features_f16": ([vp, vp, i64, vp], ctypes.c_int),
    "vaex_model_get_counter": ([vp, ctypes.c_int, ctypes.POINTER(i64)], ctypes.c_int),
    "vaex_vae_encode": ([vp, vp, i64, vp, vp, vp], ctypes.c_int),
    "vaex_reparam": ([vp, vp, vp, u64, u64, vp, i64, i32, vp], ctypes.c_int),
    "vaex_vae_decode": ([vp, vp, i64, vp, vp], ctypes.c_int),
    "vaex_vae_predict_feature": ([vp, vp, i64, vp, vp], ctypes.c_int),
    "vaex_vae_feature_dim": ([vp, vp], ctypes.c_int),
    "vaex_feature_blob_scan": ([vp, i64, i32, vp], ctypes.c_int),
    "vaex_feature_blob_unpack": ([vp, i64, i32, vp, vp, vp, vp, vp, vp], ctypes.c_int),
    "vaex_segsum_mlp_predict_blob": ([vp, vp, i64, vp, i64, ctypes.c_int], ctypes.c_int),
    "vaex_decode_snap": ([vp, vp, i64, vp, vp, vp, vp, ctypes.c_int, vp, vp, i32, vp], ctypes.c_int),
    "vaex_snap_extents": ([vp, i64, i32, vp, vp, ctypes.c_int, vp, vp, i32, vp, vp], ctypes.c_int),
    "vaex_legalise_extents": ([vp, i64, i32, vp, vp, vp, i32, i32, vp], ctypes.c_int),
    "vaex_extent_text_scan": ([vp, vp, i64, vp], ctypes.c_int),
    "vaex_extent_text_parse": ([vp, vp, i64, i32, vp, vp], ctypes.c_int),
    "vaex_scaler_fit": ([vp, i64, i32, vp, vp, vp, vp, i64, vp], ctypes.c_int),
    "vaex_scaler_transform": ([vp, i64, i32, vp, vp, vp, vp], ctypes.c_int),
    "vaex_predict_cost": ([vp, vp, i64, vp, vp, u64, f32, vp], ctypes.c_int),
    "vaex_mc_predict_cost": ([vp, vp, i64, i32, vp, u64, f32, vp, vp, vp], ctypes.c_int),
    "vaex_cost_grad_z": ([vp, vp, i64, vp, vp, vp, vp], ctypes.c_int),
    "vaex_latent_ascent": ([vp, vp, i64, i32, f32, vp, vp], ctypes.c_int),
    "vaex_latent_ascent_decoder_head": ([vp, vp, vp, i64, i32, f32, vp, vp], ctypes.c_int),
    "vaex_segvae_encode": ([vp, vp, vp, i64, i64, vp, vp, vp, ctypes.c_int, vp], ctypes.c_int),
    "vaex_segvae_decode": ([vp, vp, i64, vp, vp], ctypes.c_int),
    "vaex_segcost_predict": ([vp, vp, i64, vp, vp], ctypes.c_int),
    "vaex_trainer_create": ([i32, i32, i32, ctypes.POINTER(vp)], ctypes.c_int),
    "vaex_trainer_destroy": ([vp], ctypes.c_int),
    "vaex_trainer_param_count": ([vp], i64),
    "vaex_trainer_param_offset": ([vp, ctypes.c_char_p, ctypes.POINTER(i64), ctypes.POINTER(i64)], ctypes.c_int),
    "vaex_gather_segments": ([vp, i32, vp, vp, i64, vp, vp, vp, vp, vp], ctypes.c_int),
    "vaex_trainer_forward_backward": ([vp, vp, vp, vp, vp, vp, i64, i64, ctypes.c_int, vp, vp, vp], ctypes.c_int),
    "vaex_trainer_epoch": ([vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, ctypes.c_int, ctypes.c_int, f32, f32, f32,
                            ctypes.POINTER(i64), vp, ctypes.c_int, ctypes.POINTER(i64), vp], ctypes.c_int),
    "vaex_loss_eval": ([ctypes.c_int, vp, vp, i64, vp, vp, vp], ctypes.c_int),
    "vaex_adam_step": ([vp, vp, vp, vp, i64, i64, f32, f32, f32, f32, f32, f32, f32, vp, vp], ctypes.c_int),
    "vaex_sgd_step": ([vp, vp, i64, f32, f32, f32, vp], ctypes.c_int),
    "vaex_pair_loss": ([vp, vp, i64, f32, vp, vp, vp], ctypes.c_int),
    "vaex_kld_loss": ([vp, vp, i64, vp, vp, vp, vp], ctypes.c_int),
    "vaex_reg_trainer_create": ([i32, i32, i32, i32, ctypes.POINTER(vp)], ctypes.c_int),
    "vaex_reg_trainer_destroy": ([vp], ctypes.c_int),
    "vaex_reg_trainer_param_count": ([vp], i64),
    "vaex_reg_trainer_param_offset": ([vp, ctypes.c_char_p, ctypes.POINTER(i64), ctypes.POINTER(i64)], ctypes.c_int),
    "vaex_reg_trainer_forward_backward": ([vp, vp, vp, vp, vp, i64, vp, vp, vp, vp, vp, vp], ctypes.c_int),
    "vaex_vae_trainer_create": ([i32, i32, i32, ctypes.POINTER(vp)], ctypes.c_int),
    "vaex_vae_trainer_destroy": ([vp], ctypes.c_int),
    "vaex_vae_trainer_param_count": ([vp], i64),
    "vaex_vae_trainer_param_offset": ([vp, ctypes.c_char_p, ctypes.POINTER(i64), ctypes.POINTER(i64)], ctypes.c_int),
    "vaex_vae_trainer_forward_backward": ([vp, vp, vp, vp, i64, vp, f32, f32, vp, vp], ctypes.c_int),
    "vaex_segvae_trainer_create": ([i32, i32, i32, ctypes.POINTER(vp)], ctypes.c_int),
    "vaex_segvae_trainer_destroy": ([vp], ctypes.c_int),
    "vaex_segvae_trainer_param_count": ([vp], i64),
    "vaex_segvae_trainer_param_offset": ([vp, ctypes.c_char_p, ctypes.POINTER(i64), ctypes.POINTER(i64)], ctypes.c_int),
    "vaex_segvae_trainer_forward_backward": ([vp, vp, vp, vp, vp, i64, i64, vp, vp, u64, f32, f32, vp, vp, vp, vp, vp], ctypes.c_int),
    "vaex_segreg_trainer_create": ([i32, i32, i32, i32, i32, ctypes.POINTER(vp)], ctypes.c_int),
    "vaex_segreg_trainer_destroy": ([vp], ctypes.c_int),
    "vaex_segreg_trainer_param_count": ([vp], i64),
    "vaex_segreg_trainer_param_offset": ([vp, ctypes.c_char_p, ctypes.POINTER(i64), ctypes.POINTER(i64)], ctypes.c_int),
    "vaex_segreg_trainer_forward_backward": ([vp, vp, vp, vp, vp, vp, i64, i64, vp, vp, u64, vp, vp, vp, vp, vp], ctypes.c_int),
    "vaex_adamw_step": ([vp, vp, vp, vp, i64, i64, i64, i64, f32, f32, f32, f32, f32, f32, f32, vp, ctypes.c_int, vp], ctypes.c_int),
    "vaex_kcenter_select": ([vp, i64, i32, vp, i64, vp, i64, i32, f32, vp, vp], ctypes.c_int),
    "vaex_fps_select": ([vp, i64, i32, vp, i64, i64, i32, vp, vp], ctypes.c_int),
    "vaex_sketch_features": ([vp, i32, vp, i64, vp, vp], ctypes.c_int),
    "vaex_sketch_features_f16": ([vp, i32, vp, i64, vp, vp], ctypes.c_int),
    "vaex_latent_min_dist": ([vp, i64, vp, i64, i32, vp, vp], ctypes.c_int),
    "vaex_segcost_mc_predict": ([vp, vp, i64, i32, vp, u64, f32, vp, vp, vp], ctypes.c_int),
    "vaex_nccl_unique_id": ([vp], ctypes.c_int),
    "vaex_nccl_init": ([i32, i32, vp, ctypes.POINTER(vp)], ctypes.c_int),
    "vaex_nccl_destroy": ([vp], ctypes.c_int),
    "vaex_comm_info": ([vp, ctypes.POINTER(i32), ctypes.POINTER(i32)], ctypes.c_int),
    "vaex_topk_allgather_merge": ([vp, vp, vp, vp, i32, i64, vp, vp, vp, vp], ctypes.c_int),
    "vaex_allreduce_grads": ([vp, vp, i64, vp], ctypes.c_int),
    "vaex_allgather": ([vp, vp, vp, i64, vp], ctypes.c_int),
    "vaex_topk": ([vp, i64, i32, vp, vp, vp, vp, vp], ctypes.c_int),
}


class VaexError(RuntimeError):
    pass


_lib = None


def load():
    """Returns the CDLL with typed signatures.  Raises if the library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise VaexError(
            "%s not found: build it with vae-extent-search_b200/csrc/build.sh "
            "(or __graft_entry__.build()).  vaex_b200 has no CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (argtypes, restype) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = restype
    if lib.vaex_abi_version() != 1:
        raise VaexError("ABI version mismatch")
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load().vaex_last_error()
        raise VaexError("vaex_b200 error %d: %s" % (rc, msg.decode() if msg else "?"))


def require_cuda(what):
    """Every compute path is CUDA-only: fail with a VaexError instead of whatever torch raises first."""
    import torch
    if not torch.cuda.is_available():
        raise VaexError("%s needs a CUDA device (vaex_b200 has no CPU fallback)" % what)


def ptr(t):
    """Device (or host) pointer of a torch tensor / numpy array / None."""
    if t is None:
        return None
    if isinstance(t, np.ndarray):
        return t.ctypes.data
    return t.data_ptr()


def stream_ptr():
    import torch
    return torch.cuda.current_stream().cuda_stream


def check_device(t):
    """The library packs onto, and launches on, the CURRENT device's stream: a tensor that lives on another GPU would be
    dereferenced from the wrong context.  Fail loudly instead (wrap the call in `with torch.cuda.device(t.device):`)."""
    import torch
    if t.is_cuda and t.device.index is not None and t.device.index != torch.cuda.current_device():
        raise VaexError("tensor on cuda:%d while the current device is cuda:%d: wrap the call in torch.cuda.device(tensor.device)"
                        % (t.device.index, torch.cuda.current_device()))
    return t


class Model:
    """Owns a vaex_model_t.  weights: dict name -> array-like (reference state_dict keys)."""

    def __init__(self, kind, dims):
        self.lib = load()
        self.kind = kind
        self.h = vp()
        arr = (ctypes.c_int32 * len(dims))(*[int(d) for d in dims])
        check(self.lib.vaex_model_create(kind, arr, len(dims), ctypes.byref(self.h)))

    def load_state(self, state, norm_vec=None):
        for name, t in state.items():
            if hasattr(t, "detach"):
                t = t.detach().to("cpu").float().numpy()
            a = np.ascontiguousarray(np.asarray(t, dtype=np.float32))
            if a.ndim == 0 or "num_batches_tracked" in name:
                continue
            rows, cols = (1, a.shape[0]) if a.ndim == 1 else a.shape
            check(self.lib.vaex_model_load_weights(self.h, name.encode(), a.ctypes.data, rows, cols))
        nv = None
        if norm_vec is not None:
            if hasattr(norm_vec, "detach"):
                norm_vec = norm_vec.detach().to("cpu").float().numpy()
            nv = np.ascontiguousarray(np.asarray(norm_vec, dtype=np.float32))
        check(self.lib.vaex_model_pack(self.h, nv.ctypes.data if nv is not None else None, None))

    def set_option(self, option, value):
        check(self.lib.vaex_model_set_option(self.h, int(option), int(value)))

    def counter(self, which):
        v = i64(0)
        check(self.lib.vaex_model_get_counter(self.h, int(which), ctypes.byref(v)))
        return int(v.value)

    def close(self):
        if self.h:
            self.lib.vaex_model_destroy(self.h)
            self.h = vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
