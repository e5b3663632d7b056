"""Drop-in mirror of the notebook's extent-VAE classes
(reference: scripts/pre_experiments/model_myself/utils/model.py, inlined in
vae_extent_search.py:258-339,515-702).  Same constructors, method names, return tuples and
state_dict keys; encode / decode / predict_cost / mc_predict / the z-gradient run in the
sm_100a kernels behind libvaex_b200.so.  CUDA tensors only -- no CPU fallback."""
import numpy as np
import torch
import torch.nn as nn

from . import _lib


def _mlp(dims, last_act=True):
    mods = []
    for i in range(len(dims) - 1):
        mods.append(nn.Linear(dims[i], dims[i + 1]))
        if last_act or i + 2 < len(dims):
            mods.append(nn.ReLU())
    return nn.Sequential(*mods)


def _need_cuda(t):
    if not t.is_cuda:
        raise _lib.VaexError("vaex_b200 needs CUDA tensors (no CPU fallback)")
    return _lib.check_device(t).detach().contiguous().float()


class _Packed:
    """Mixin: lazily (re)packs the module's state_dict into a kernel-side model."""
    _kind = None

    def _dims(self):
        raise NotImplementedError

    def packed(self):
        key = tuple((p.data_ptr(), p._version)
                    for p in self.state_dict(keep_vars=True).values())
        cached = self.__dict__.get("_packed_cache")
        if cached is not None and cached[1] == key:
            return cached[0]
        m = _lib.Model(self._kind, self._dims())
        m.load_state(dict(self.state_dict()))
        m.set_option(_lib.OPT_TENSOR_CORES, int(self.__dict__.get("_tensor_cores", True)))
        self.__dict__["_packed_cache"] = (m, key)
        return m

    def invalidate_packed(self):
        """Drops the packed copy.  The cache key is (storage address, autograd version) per tensor -- no device read, so a
        call costs nothing and is legal under CUDA-graph capture -- and writes through .data do not bump the version."""
        self.__dict__.pop("_packed_cache", None)

    def set_tensor_cores(self, on=True):
        """True (default): decoder / cost predictor on the fused tcgen05 chain (fp16 operands, fp32
        accumulate: ~1e-3 of the reference's fp32); False: the fp32 SIMT kernels (1e-5)."""
        self.__dict__["_tensor_cores"] = bool(on)
        cached = self.__dict__.get("_packed_cache")
        if cached is not None:
            cached[0].set_option(_lib.OPT_TENSOR_CORES, int(bool(on)))
        return self

    def __getstate__(self):
        d = self.__dict__.copy()
        d.pop("_packed_cache", None)
        return d

    def _encode(self, x, want_logvar=True):
        x = _need_cuda(x)
        n, L = x.shape[0], self.fc_mu.out_features
        mu = torch.empty(n, L, device=x.device)
        logvar = torch.empty(n, L, device=x.device) if want_logvar else None
        _lib.check(_lib.load().vaex_vae_encode(self.packed().h, _lib.ptr(x), n, _lib.ptr(mu), _lib.ptr(logvar),
                                              _lib.stream_ptr()))
        return mu, logvar

    def reparameterize(self, mu, logvar, eps=None, seed=None):
        """model.py:58-61.  eps=None draws N(0,1) from the kernel's Philox stream."""
        mu, logvar = _need_cuda(mu), _need_cuda(logvar)
        z = torch.empty_like(mu)
        if seed is None:
            seed = int(torch.randint(1, 2 ** 62, (1,)).item())
        _lib.check(_lib.load().vaex_reparam(_lib.ptr(mu), _lib.ptr(logvar), _lib.ptr(_need_cuda(eps)) if eps is not None else None,
                                           seed, 0, _lib.ptr(z), mu.shape[0], mu.shape[1], _lib.stream_ptr()))
        return z


class VAE_feature_head(_Packed, nn.Module):
    """utils/model.py:6-81."""
    _kind = _lib.KIND_EXTENT_VAE

    def __init__(self, input_dim, feature_dim=None, latent_dim=16, hidden_dim=128):
        super().__init__()
        self.encoder = _mlp([input_dim, hidden_dim, hidden_dim, hidden_dim])
        self.fc_mu = nn.Linear(hidden_dim, latent_dim)
        self.fc_logvar = nn.Linear(hidden_dim, latent_dim)
        self.decoder = _mlp([latent_dim, hidden_dim, hidden_dim, hidden_dim, input_dim], last_act=False)
        self.use_feature = feature_dim is not None
        if self.use_feature:
            self.feature_predictor = _mlp([latent_dim, hidden_dim, hidden_dim, feature_dim], last_act=False)

    def _dims(self):
        return [self.encoder[0].in_features, self.encoder[0].out_features, self.fc_mu.out_features]

    def encode(self, x, use_mean=False):
        mu, logvar = self._encode(x, want_logvar=not use_mean)
        return mu if use_mean else (mu, logvar)

    def decode(self, z):
        z = _need_cuda(z)
        x = torch.empty(z.shape[0], self.decoder[-1].out_features, device=z.device)
        _lib.check(_lib.load().vaex_vae_decode(self.packed().h, _lib.ptr(z), z.shape[0], _lib.ptr(x), _lib.stream_ptr()))
        return x

    def decode_snap(self, z, mean, scale, log1p=False, slot_extent=None, slot_last=None, max_innermost=64):
        """decode -> postprocess (vae_sch_extent.ipynb:1374-1383) -> clamp(min=1) (:1457), and,
        when slot tables are given, legalise to tile factors (utils.cc:433-463 semantics)."""
        return self.decode_and_snap(z, mean, scale, log1p, slot_extent, slot_last, max_innermost, want_x=False)[1]

    def decode_and_snap(self, z, mean, scale, log1p=False, slot_extent=None, slot_last=None, max_innermost=64, want_x=True):
        """One launch: (x fp32 or None, extents int32).  The snap runs in the decoder kernel's epilogue."""
        z = _need_cuda(z)
        dev, n, D = z.device, z.shape[0], self.decoder[-1].out_features
        mean = torch.as_tensor(mean, dtype=torch.float32).to(dev).contiguous()
        scale = torch.as_tensor(scale, dtype=torch.float32).to(dev).contiguous()
        x = torch.empty(n, D, device=dev) if want_x else None
        out = torch.empty(n, D, dtype=torch.int32, device=dev)
        mode = _lib.SNAP_EXPM1 if log1p else _lib.SNAP_EXP
        se = sl = None
        if slot_extent is not None:
            mode |= _lib.SNAP_DIVISOR
            se = torch.as_tensor(slot_extent, dtype=torch.int32).to(dev).contiguous()
            sl = torch.as_tensor(slot_last, dtype=torch.uint8).to(dev).contiguous()
        _lib.check(_lib.load().vaex_decode_snap(self.packed().h, _lib.ptr(z), n, _lib.ptr(x), _lib.ptr(out), _lib.ptr(mean),
                                               _lib.ptr(scale), mode, _lib.ptr(se), _lib.ptr(sl), max_innermost, _lib.stream_ptr()))
        return x, out

    def latent_ascent_through(self, cost_model, z0, steps=200, lr=1e-2):
        """BASELINE config 3, north-star graph: `steps` ascent steps on score(z) = head(decoder_trunk(z)) where head is
        cost_model's (a SegmentSumMLPModule) residual head.  -> (z, score of the final z)."""
        z = _need_cuda(z0).clone()
        score = torch.empty(z.shape[0], device=z.device)
        _lib.check(_lib.load().vaex_latent_ascent_decoder_head(self.packed().h, cost_model.packed().h, _lib.ptr(z), z.shape[0],
                                                              int(steps), float(lr), _lib.ptr(score), _lib.stream_ptr()))
        return z, score

    def predict_feature(self, z):
        """utils/model.py:66-67: z -> feature_predictor(z) (only when the module was built with feature_dim)."""
        if not self.use_feature:
            raise AttributeError("this VAE_feature_head was built without feature_dim: it has no feature_predictor")
        z = _need_cuda(z)
        out = torch.empty(z.shape[0], self.feature_predictor[-1].out_features, device=z.device)
        _lib.check(_lib.load().vaex_vae_predict_feature(self.packed().h, _lib.ptr(z), z.shape[0], _lib.ptr(out), _lib.stream_ptr()))
        return out

    def forward(self, x, use_mean=True):
        mu, logvar = self._encode(x)
        z = mu if use_mean else self.reparameterize(mu, logvar)
        return self.decode(z), mu, logvar, z, (self.predict_feature(z) if self.use_feature else None)


def snap_extents(x, mean, scale, log1p=False, slot_extent=None, slot_last=None, max_innermost=64):
    x = _need_cuda(x)
    dev = x.device
    mean = torch.as_tensor(mean, dtype=torch.float32).to(dev).contiguous()
    scale = torch.as_tensor(scale, dtype=torch.float32).to(dev).contiguous()
    out = torch.empty(x.shape, dtype=torch.int32, device=dev)
    mode = _lib.SNAP_EXPM1 if log1p else _lib.SNAP_EXP
    se = sl = None
    if slot_extent is not None:
        mode |= _lib.SNAP_DIVISOR
        se = torch.as_tensor(slot_extent, dtype=torch.int32).to(dev).contiguous()
        sl = torch.as_tensor(slot_last, dtype=torch.uint8).to(dev).contiguous()
    _lib.check(_lib.load().vaex_snap_extents(_lib.ptr(x), x.shape[0], x.shape[1], _lib.ptr(mean), _lib.ptr(scale), mode,
                                            _lib.ptr(se), _lib.ptr(sl), max_innermost, _lib.ptr(out), _lib.stream_ptr()))
    return out


class _PredictCostFn(torch.autograd.Function):
    """predict_cost with a CUDA backward so that torch.autograd.grad(cost.sum(), z)
    (utils/select.py:128-135) keeps working: d cost/d z comes from vaex_cost_grad_z."""

    @staticmethod
    def forward(ctx, z, model):
        zc = _need_cuda(z)
        n = zc.shape[0]
        cost = torch.empty(n, device=zc.device)
        grad = torch.empty_like(zc)
        _lib.check(_lib.load().vaex_cost_grad_z(model.packed().h, _lib.ptr(zc), n, _lib.ptr(cost), _lib.ptr(grad), None,
                                               _lib.stream_ptr()))
        ctx.save_for_backward(grad)
        return cost

    @staticmethod
    def backward(ctx, gout):
        (grad,) = ctx.saved_tensors
        return gout.unsqueeze(1) * grad, None


class VAECostPredictor(_Packed, nn.Module):
    """utils/model.py:86-273 (extent family: 2 predictor layers, Dropout after the first)."""
    _kind = _lib.KIND_COST_PREDICTOR

    def __init__(self, input_dim, feature_dim=None, hidden_dim=256, latent_dim=64, predictor_hidden=256,
                 predictor_layers=2, dropout=0.1, use_feature=False):
        super().__init__()
        if predictor_layers != 2:
            raise NotImplementedError("the extent-family kernels implement predictor_layers=2 (the notebook's setting)")
        self.input_dim, self.feature_dim, self.hidden_dim, self.latent_dim = input_dim, feature_dim, hidden_dim, latent_dim
        self.dropout_p = dropout
        self.encoder = _mlp([input_dim, hidden_dim, hidden_dim, hidden_dim])
        self.fc_mu = nn.Linear(hidden_dim, latent_dim)
        self.fc_logvar = nn.Linear(hidden_dim, latent_dim)
        self.cost_predictor = nn.Sequential(
            nn.Linear(latent_dim, predictor_hidden), nn.ReLU(), nn.Dropout(dropout),
            nn.Linear(predictor_hidden, predictor_hidden), nn.ReLU(), nn.Identity(),
            nn.Linear(predictor_hidden, 1))
        self.use_feature = use_feature
        if use_feature:
            self.feature_predictor = _mlp([latent_dim, hidden_dim, hidden_dim, feature_dim], last_act=False)

    def _dims(self):
        return [self.encoder[0].in_features, self.encoder[0].out_features, self.fc_mu.out_features,
                self.cost_predictor[0].out_features]

    def encode(self, input_data, use_mean=False):
        mu, logvar = self._encode(input_data, want_logvar=not use_mean)
        return mu if use_mean else (mu, logvar)

    def predict_cost(self, z, drop_mask=None, drop_seed=0):
        """utils/model.py:168-170.  Inference only: Dropout is active only when a mask or a seed is given (mc_predict
        passes them), whatever self.training says, and the backward returns d cost / d z but no parameter gradients --
        training runs through training.RegressionTrainer (vaex_reg_trainer_forward_backward)."""
        if self.training and torch.is_grad_enabled() and any(p.requires_grad for p in self.cost_predictor.parameters()) \
                and drop_mask is None and not drop_seed:
            raise _lib.VaexError("VAECostPredictor.predict_cost in train mode builds no parameter gradients and applies no "
                                 "Dropout: call .eval() / torch.no_grad() to score, or train with training.RegressionTrainer")
        if z.requires_grad and torch.is_grad_enabled():
            return _PredictCostFn.apply(z, self)
        zc = _need_cuda(z)
        cost = torch.empty(zc.shape[0], device=zc.device)
        mk = None
        if drop_mask is not None:
            mk = drop_mask.to(device=zc.device, dtype=torch.uint8).contiguous()
        _lib.check(_lib.load().vaex_predict_cost(self.packed().h, _lib.ptr(zc), zc.shape[0], _lib.ptr(cost), _lib.ptr(mk),
                                                int(drop_seed), float(self.cost_predictor[2].p), _lib.stream_ptr()))
        return cost

    def forward(self, input_data, use_mean=True):
        mean, logvar = self._encode(input_data)
        z = mean if use_mean else self.reparameterize(mean, logvar)
        return self.predict_cost(z), mean, logvar, z

    def get_encoder_params(self):
        return list(self.encoder.parameters()) + list(self.fc_mu.parameters()) + list(self.fc_logvar.parameters())

    def get_cost_predictor_params(self):
        return self.cost_predictor.parameters()

    def get_feature_predictor_params(self):
        return self.feature_predictor.parameters()

    def load_pretrained_encoder(self, checkpoint):
        """utils/model.py:215-233: copy encoder / fc_mu / fc_logvar tensors whose shapes match."""
        own = self.state_dict()
        with torch.no_grad():
            for name, param in checkpoint.items():
                if name.startswith(("encoder", "fc_mu", "fc_logvar")) and name in own and own[name].shape == param.shape:
                    own[name].copy_(param)

    def mc_predict(self, input_tensor, T=20, masks=None, seed=None):
        """utils/model.py:241-273: T dropout-active passes -> (mean, unbiased var) over T.
        masks (uint8 [T, N, PH]) reproduces a given Bernoulli draw exactly; otherwise each pass
        uses the kernel's Philox stream."""
        self.eval()
        mu, _ = self._encode(input_tensor, want_logvar=False)
        if seed is None:
            seed = int(torch.randint(1, 2 ** 62, (1,)).item())
        n = mu.shape[0]
        mean, var = torch.empty(n, device=mu.device), torch.empty(n, device=mu.device)
        mk = None
        if masks is not None:
            mk = masks.to(device=mu.device, dtype=torch.uint8).contiguous()
            if mk.shape[0] < T:
                raise ValueError("masks holds %d passes, T = %d" % (mk.shape[0], T))
        _lib.check(_lib.load().vaex_mc_predict_cost(self.packed().h, _lib.ptr(mu), n, int(T), _lib.ptr(mk), 0 if mk is not None else int(seed),
                                                   float(self.cost_predictor[2].p), _lib.ptr(mean), _lib.ptr(var), _lib.stream_ptr()))
        return mean, var

    def cost_grad_z(self, z):
        """(cost, d cost.sum()/d z, ||grad||_2 per row) in one call (select_topk_z_grad, select.py:111-148)."""
        zc = _need_cuda(z)
        n = zc.shape[0]
        cost, norm, grad = torch.empty(n, device=zc.device), torch.empty(n, device=zc.device), torch.empty_like(zc)
        _lib.check(_lib.load().vaex_cost_grad_z(self.packed().h, _lib.ptr(zc), n, _lib.ptr(cost), _lib.ptr(grad), _lib.ptr(norm),
                                               _lib.stream_ptr()))
        return cost, grad, norm

    def latent_ascent(self, z0, steps=200, lr=1e-2):
        """BASELINE config 3: `steps` plain ascent steps z += lr * dcost/dz on every restart."""
        z = _need_cuda(z0).clone()
        cost = torch.empty(z.shape[0], device=z.device)
        _lib.check(_lib.load().vaex_latent_ascent(self.packed().h, _lib.ptr(z), z.shape[0], int(steps), float(lr),
                                                 _lib.ptr(cost), _lib.stream_ptr()))
        return z, cost


TOPK_MAX = 1024      # winners per vaex_topk call (csrc/topk.cu)


def topk(scores, k, mask=None, sync=True):
    """GPU top-k with the deterministic (score desc, index asc) order -> (values, int64 indices).
    sync=False skips the host read of the winner count and returns (values[k], indices[k],
    count) with count a device int32 tensor; entries past count are unspecified."""
    s = _need_cuda(scores)
    k = int(min(k, s.shape[0]))
    if k > TOPK_MAX:
        # vaex_topk selects at most 1024 winners per call (one CTA sorts them); larger k -- select_programs' pool of
        # topk_factor * num_select candidates for num_select >= 205, utils/select.py:329-334 -- is taken in rounds of 1024
        # with the winners masked out: the concatenation is still (score desc, index asc)
        mk = torch.ones(s.shape[0], dtype=torch.uint8, device=s.device) if mask is None else \
            mask.to(device=s.device, dtype=torch.uint8).clone()
        vs, ix, left = [], [], k
        while left > 0:
            v, i = topk(s, min(left, TOPK_MAX), mk, sync=True)
            if v.numel() == 0:
                break
            vs.append(v)
            ix.append(i)
            mk[i] = 0
            left -= v.numel()
        vals = torch.cat(vs) if vs else torch.empty(0, device=s.device)
        idx = torch.cat(ix) if ix else torch.empty(0, dtype=torch.int64, device=s.device)
        if sync:
            return vals, idx
        return vals, idx, torch.tensor([vals.numel()], dtype=torch.int32, device=s.device)
    vals = torch.empty(k, device=s.device)
    idx = torch.empty(k, dtype=torch.int64, device=s.device)
    n_out = torch.zeros(1, dtype=torch.int32, device=s.device)
    mk = None if mask is None else mask.to(device=s.device, dtype=torch.uint8).contiguous()
    _lib.check(_lib.load().vaex_topk(_lib.ptr(s), s.shape[0], k, _lib.ptr(mk), _lib.ptr(vals), _lib.ptr(idx), _lib.ptr(n_out),
                                    _lib.stream_ptr()))
    if not sync:
        return vals, idx, n_out
    m = int(n_out.item())
    return vals[:m], idx[:m]
