"""Drop-in mirror of the segment-VAE family
(reference: scripts/vae_experiments/models/vae.py `SegmentVAE`, models/regression.py
`VAECostPredictor`).  Same constructors, attribute / state_dict key names (the shipped
model_final/checkpoints/*.pt and for_inference/*.pth load with load_state_dict), method names and
return tuples.  Inference (eval mode) runs in libvaex_b200.so: row encoder + segment sum on the
tcgen05 kernel, BatchNorm folded into affines at pack time, Dropout = identity.  CUDA tensors
only -- no CPU fallback; training mode raises until the backward kernels land."""
import numpy as np
import torch
import torch.nn as nn

from . import _lib
from .vae_model import _need_cuda


class _SegmentBase(nn.Module):
    _kind = None

    def _dims(self):
        raise NotImplementedError

    def packed(self, fea_norm_vec=None):
        nk = None
        if fea_norm_vec is not None:
            nv = fea_norm_vec.detach().cpu().numpy() if hasattr(fea_norm_vec, "detach") else fea_norm_vec
            nk = hash(np.ascontiguousarray(nv, dtype=np.float32).tobytes())
        key = tuple((p.data_ptr(), p._version) for p in self.state_dict(keep_vars=True).values())
        cache = self.__dict__.setdefault("_packed_cache", {})           # one packed copy per fea_norm_vec (select_features
        cached = cache.get(nk)                                           # alternates between a folded and a raw one)
        if cached is not None and cached[1] == key:
            return cached[0]
        m = _lib.Model(self._kind, self._dims())
        m.load_state(self.state_dict(), fea_norm_vec)
        m.set_option(_lib.OPT_TENSOR_CORES, int(self.__dict__.get("_tensor_cores", True)))
        for k in [k for k, v in cache.items() if v[1] != key]:           # copies of weights that have since changed
            del cache[k]
        while len(cache) >= 2:
            del cache[next(iter(cache))]
        cache[nk] = (m, key)
        return m

    def invalidate_packed(self):
        """Drops the packed copy.  The cache key is (storage address, autograd version) per tensor -- no device read, so a
        call costs nothing and is legal under CUDA-graph capture -- and writes through .data do not bump the version."""
        self.__dict__.pop("_packed_cache", None)

    def set_tensor_cores(self, on=True):
        """True (default): the latent heads, the decoder and the cost predictor run as single-layer launches of the tcgen05
        head kernel (fp16 operands, fp32 accumulate and residuals: ~1e-3 of the reference's fp32); False: fp32 SIMT (1e-5).
        The row encoder takes VAEX_F_FORCE_SIMT per call (flags=)."""
        self.__dict__["_tensor_cores"] = bool(on)
        for cached in self.__dict__.get("_packed_cache", {}).values():
            cached[0].set_option(_lib.OPT_TENSOR_CORES, int(bool(on)))
        return self

    def __getstate__(self):
        d = self.__dict__.copy()
        d.pop("_packed_cache", None)
        return d

    def _check_eval(self):
        if self.training:
            raise NotImplementedError("the CUDA path implements eval-mode inference (BatchNorm running stats, "
                                      "Dropout off); call .eval() -- training lands with the backward kernels")

    def _offsets(self, segment_sizes, device):
        sizes = segment_sizes.detach().to(device=device, dtype=torch.int32).contiguous()
        offsets = torch.empty(sizes.shape[0] + 1, dtype=torch.int32, device=device)
        _lib.check(_lib.load().vaex_segment_offsets(_lib.ptr(sizes), sizes.shape[0], _lib.ptr(offsets), _lib.stream_ptr()))
        return offsets

    def _encode_segments(self, segment_sizes, features, flags=0, fea_norm_vec=None, want_segsum=True):
        """-> (mean, logvar, segment_sum_vec)"""
        self._check_eval()
        feats = _need_cuda(features)
        n = segment_sizes.shape[0]
        offsets = self._offsets(segment_sizes, feats.device)
        mean = torch.empty(n, self.latent_dim, device=feats.device)
        logvar = torch.empty(n, self.latent_dim, device=feats.device)
        segsum = torch.empty(n, self.hidden_dim, device=feats.device) if want_segsum else None
        _lib.check(_lib.load().vaex_segvae_encode(self.packed(fea_norm_vec).h, _lib.ptr(feats), _lib.ptr(offsets), n,
                                                 feats.shape[0], _lib.ptr(mean), _lib.ptr(logvar), _lib.ptr(segsum),
                                                 flags, _lib.stream_ptr()))
        return mean, logvar, segsum

    def reparameterize(self, mean, logvar, eps=None, seed=None):
        """vae.py:103-107 / regression.py:103-107; eps=None draws N(0,1) from the kernel's Philox stream."""
        mean, logvar = _need_cuda(mean), _need_cuda(logvar)
        z = torch.empty_like(mean)
        if seed is None:
            seed = int(torch.randint(1, 2 ** 62, (1,)).item())
        _lib.check(_lib.load().vaex_reparam(_lib.ptr(mean), _lib.ptr(logvar), _lib.ptr(_need_cuda(eps)) if eps is not None else None,
                                           seed, 0, _lib.ptr(z), mean.shape[0], mean.shape[1], _lib.stream_ptr()))
        return z


class SegmentVAE(_SegmentBase):
    """scripts/vae_experiments/models/vae.py:16-137."""
    _kind = _lib.KIND_SEGMENT_VAE

    def __init__(self, input_dim, hidden_dim=256, latent_dim=64, dropout=0.1):
        super().__init__()
        self.input_dim, self.hidden_dim, self.latent_dim = input_dim, hidden_dim, latent_dim
        self.use_mean = False
        self.segment_encoder = nn.Sequential(nn.Linear(input_dim, hidden_dim), nn.ReLU(),
                                             nn.Linear(hidden_dim, hidden_dim), nn.ReLU())
        self.norm = nn.BatchNorm1d(hidden_dim)
        self.l0 = nn.Sequential(nn.Linear(hidden_dim, hidden_dim), nn.ReLU())
        self.l1 = nn.Sequential(nn.Linear(hidden_dim, hidden_dim), nn.ReLU())
        self.fc_mean = nn.Linear(hidden_dim, latent_dim)
        self.fc_logvar = nn.Linear(hidden_dim, latent_dim)
        self.decoder = nn.Sequential(
            nn.Linear(latent_dim, hidden_dim), nn.BatchNorm1d(hidden_dim), nn.ReLU(), nn.Dropout(dropout),
            nn.Linear(hidden_dim, hidden_dim), nn.BatchNorm1d(hidden_dim), nn.ReLU(),
            nn.Linear(hidden_dim, hidden_dim))

    def _dims(self):
        return [self.input_dim, self.hidden_dim, self.latent_dim]

    def segment_sum(self, segment_sizes, features, flags=0):
        """vae.py:68-85 takes the *encoded* rows; here the row encoder and the sum are one kernel,
        so this takes the raw feature rows and returns sum_r segment_encoder(features)[r]."""
        return self._encode_segments(segment_sizes, features, flags)[2]

    def encode_segments(self, segment_sizes, features, flags=0, fea_norm_vec=None):
        """vae.py:118-128: (mean, logvar, segment_sum_vec); logvar = 0 when self.use_mean (vae.py:99-100)."""
        mean, logvar, segsum = self._encode_segments(segment_sizes, features, flags, fea_norm_vec)
        if self.use_mean:
            logvar = torch.zeros_like(mean)
        return mean, logvar, segsum

    def decode(self, z):
        """vae.py:109-111."""
        self._check_eval()
        z = _need_cuda(z)
        recon = torch.empty(z.shape[0], self.hidden_dim, device=z.device)
        _lib.check(_lib.load().vaex_segvae_decode(self.packed().h, _lib.ptr(z), z.shape[0], _lib.ptr(recon), _lib.stream_ptr()))
        return recon

    def forward(self, segment_sizes, features, use_mean=False, eps=None, flags=0):
        """vae.py:113-137 -> (mean, logvar, z, recon, segment_sum_vec).  Like the reference,
        z = mean + eps * exp(0.5 logvar) is sampled even when use_mean (logvar is then 0)."""
        self.use_mean = use_mean
        mean, logvar, segsum = self.encode_segments(segment_sizes, features, flags)
        z = self.reparameterize(mean, logvar, eps)
        return mean, logvar, z, self.decode(z), segsum


class VAECostPredictor(_SegmentBase):
    """scripts/vae_experiments/models/regression.py:11-185 (segment family)."""
    _kind = _lib.KIND_SEGMENT_COST

    def __init__(self, input_dim, hidden_dim=256, latent_dim=128, predictor_hidden=256, predictor_layers=3, dropout=0.1):
        super().__init__()
        self.input_dim, self.hidden_dim, self.latent_dim = input_dim, hidden_dim, latent_dim
        self.predictor_hidden, self.predictor_layers = predictor_hidden, predictor_layers
        self.segment_encoder = nn.Sequential(nn.Linear(input_dim, hidden_dim), nn.ReLU(),
                                             nn.Linear(hidden_dim, hidden_dim), nn.ReLU())
        self.norm = nn.BatchNorm1d(hidden_dim)
        self.l0 = nn.Sequential(nn.Linear(hidden_dim, hidden_dim), nn.ReLU())
        self.l1 = nn.Sequential(nn.Linear(hidden_dim, hidden_dim), nn.ReLU())
        self.fc_mean = nn.Linear(hidden_dim, latent_dim)
        self.fc_logvar = nn.Linear(hidden_dim, latent_dim)
        mods, cur = [], latent_dim
        for i in range(predictor_layers):
            mods.extend([nn.Linear(cur, predictor_hidden), nn.BatchNorm1d(predictor_hidden), nn.ReLU(),
                         nn.Dropout(dropout) if i < predictor_layers - 1 else nn.Identity()])
            cur = predictor_hidden
        mods.append(nn.Linear(predictor_hidden, 1))
        self.cost_predictor = nn.Sequential(*mods)

    def _dims(self):
        return [self.input_dim, self.hidden_dim, self.latent_dim, self.predictor_hidden, self.predictor_layers]

    def encode(self, segment_sizes, features, flags=0, fea_norm_vec=None):
        """regression.py:82-101 -> (mean, logvar, segment_sum_vec)."""
        return self._encode_segments(segment_sizes, features, flags, fea_norm_vec)

    def predict_cost(self, z):
        """regression.py:109-111."""
        self._check_eval()
        z = _need_cuda(z)
        cost = torch.empty(z.shape[0], device=z.device)
        _lib.check(_lib.load().vaex_segcost_predict(self.packed().h, _lib.ptr(z), z.shape[0], _lib.ptr(cost), _lib.stream_ptr()))
        return cost

    def forward(self, segment_sizes, features, use_mean=False, eps=None, flags=0, fea_norm_vec=None):
        """regression.py:113-135 -> (cost_pred, mean, logvar, z)."""
        mean, logvar, _ = self._encode_segments(segment_sizes, features, flags, fea_norm_vec, want_segsum=False)
        z = mean if use_mean else self.reparameterize(mean, logvar, eps)
        return self.predict_cost(z), mean, logvar, z

    def get_latent_embedding(self, segment_sizes, features, use_mean=True, fea_norm_vec=None, flags=0):
        """The latent of every program, as select_features asks the model for it
        (scripts/pre_experiments/for_inference/tuning_improved.py:74): the encoder mean, or a reparameterised draw."""
        mean, logvar, _ = self._encode_segments(segment_sizes, features, flags, fea_norm_vec, want_segsum=False)
        return mean if use_mean else self.reparameterize(mean, logvar)

    def mc_predict(self, segment_sizes, features, T=20, use_mean=True, masks=None, seed=None, fea_norm_vec=None, flags=0):
        """T dropout-active passes of the cost predictor -> (mean, unbiased variance) over the passes: the call
        select_features makes (tuning_improved.py:89); the reference's segment model does not define it, the extent
        family's is utils/model.py:241-273.  The encoder runs once (eval mode: deterministic).  masks (uint8
        [T, predictor_layers - 1, N, predictor_hidden], 1 = keep) replays a given Bernoulli draw."""
        z = self.get_latent_embedding(segment_sizes, features, use_mean, fea_norm_vec, flags)
        n = z.shape[0]
        if seed is None:
            seed = int(torch.randint(1, 2 ** 62, (1,)).item())
        mk = None
        if masks is not None:
            mk = masks.to(device=z.device, dtype=torch.uint8).contiguous()
            want = (self.predictor_layers - 1, n, self.predictor_hidden)
            if mk.dim() != 4 or mk.shape[0] < T or tuple(mk.shape[1:]) != want:
                raise ValueError("masks must be [>= %d, %d, %d, %d], got %s" % ((T,) + want + (tuple(mk.shape),)))
        mean, var = torch.empty(n, device=z.device), torch.empty(n, device=z.device)
        _lib.check(_lib.load().vaex_segcost_mc_predict(self.packed(fea_norm_vec).h, _lib.ptr(z), n, int(T), _lib.ptr(mk),
                                                      0 if mk is not None else int(seed), float(getattr(self.cost_predictor[3], "p", 0.0)),
                                                      _lib.ptr(mean), _lib.ptr(var), _lib.stream_ptr()))
        return mean, var

    def get_encoder_params(self):
        """regression.py:137-146."""
        ps = []
        for mod in (self.segment_encoder, self.norm, self.l0, self.l1, self.fc_mean, self.fc_logvar):
            ps.extend(mod.parameters())
        return ps

    def get_predictor_params(self):
        return self.cost_predictor.parameters()

    def load_pretrained_encoder(self, checkpoint):
        """regression.py:152-176: copies the matching encoder tensors of a SegmentVAE checkpoint."""
        vae_state = checkpoint["model_state_dict"] if "model_state_dict" in checkpoint else checkpoint
        keys = ["segment_encoder", "norm", "l0", "l1", "fc_mean", "fc_logvar"]
        own = self.state_dict()
        loaded = []
        with torch.no_grad():
            for name, param in vae_state.items():
                if any(name.startswith(k) for k in keys) and name in own and own[name].shape == param.shape:
                    own[name].copy_(param)
                    loaded.append(name)
        print(f"Loaded {len(loaded)} parameters from pretrained VAE")
        return (checkpoint.get("config", {}), checkpoint.get("metrics", {})) if isinstance(checkpoint, dict) else ({}, {})
