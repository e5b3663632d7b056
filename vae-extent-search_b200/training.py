"""GPU training of the extent-family VAECostPredictor (reference: scripts/pre_experiments/model_myself/
utils/training.py:133-275 and the notebook's train_regression, vae_extent_search.py:1389-1440).

`reg_loss_fn / pair_loss_fn / kld_loss_fn / smooth_loss_fn / compute_total_loss` keep the reference's names and
arguments and return device scalars (values only -- what the validation loop prints); the differentiable path is
`RegressionTrainer`: flat fp32 parameter / gradient / AdamW-moment buffers, one CUDA call for
forward(use_mean=True) + total loss + backward, one per parameter group for clip_grad_norm_(1.0) + AdamW.
Data-parallel training all-reduces the flat gradient buffer (NCCL) between the two, as train.Trainer does."""
import ctypes

import numpy as np
import torch

from . import _lib, shard


def _f(t):
    if not t.is_cuda:
        raise _lib.VaexError("vaex_b200 needs CUDA tensors (no CPU fallback)")
    return t.detach().contiguous().float()


def reg_loss_fn(cost_pred, cost_true, loss_type='mse'):
    """training.py:133-141."""
    p, y = _f(cost_pred).view(-1), _f(cost_true).view(-1)
    if loss_type != 'mse':
        return (p - y).abs().mean()
    out = torch.zeros(1, device=p.device)
    _lib.check(_lib.load().vaex_loss_eval(_lib.LOSS_IDS["rmse"], _lib.ptr(p), _lib.ptr(y), p.shape[0], _lib.ptr(out), None,
                                         _lib.stream_ptr()))
    return out[0]


def pair_loss_fn(cost_pred, cost_true, margin=0.1, return_grad=False):
    """training.py:144-166: margin ranking over all pairs i < j (no [B, B] tensors: one CTA walks the pairs)."""
    p, y = _f(cost_pred).view(-1), _f(cost_true).view(-1)
    out = torch.zeros(1, device=p.device)
    grad = torch.empty_like(p) if return_grad else None
    _lib.check(_lib.load().vaex_pair_loss(_lib.ptr(p), _lib.ptr(y), p.shape[0], float(margin), _lib.ptr(out), _lib.ptr(grad),
                                         _lib.stream_ptr()))
    return (out[0], grad) if return_grad else out[0]


def kld_loss_fn(mean, logvar, return_grad=False):
    """training.py:181-186."""
    m, l = _f(mean), _f(logvar)
    out = torch.zeros(1, device=m.device)
    dm, dl = (torch.empty_like(m), torch.empty_like(l)) if return_grad else (None, None)
    _lib.check(_lib.load().vaex_kld_loss(_lib.ptr(m), _lib.ptr(l), m.numel(), _lib.ptr(out), _lib.ptr(dm), _lib.ptr(dl),
                                        _lib.stream_ptr()))
    return (out[0], dm, dl) if return_grad else out[0]


def smooth_loss_fn(model, z, noise_std=0.1, noise=None):
    """training.py:169-178 (eval-mode predictor on z and z + noise_std * randn)."""
    model.eval()
    z = _f(z)
    if noise is None:
        noise = noise_std * torch.randn_like(z)
    return reg_loss_fn(model.predict_cost(z), model.predict_cost(z + noise))


def compute_total_loss(model, cost_pred, mean, logvar, z, labels, feature, config, return_components=True, noise=None):
    """training.py:200-226, values only (the differentiable step is RegressionTrainer.step)."""
    reg = reg_loss_fn(cost_pred, labels, loss_type=config.get('loss_type', 'mse'))
    pair = pair_loss_fn(cost_pred, labels, margin=config.get('margin', 0.1))
    smooth = smooth_loss_fn(model, z, noise_std=config.get('noise_std', 0.1), noise=noise)
    kld = kld_loss_fn(mean, logvar)
    total = config['lambda_reg'] * reg + config['lambda_pair'] * pair + config['gamma'] * smooth + config['beta'] * kld
    if return_components:
        return total, {'reg_loss': reg.item(), 'pair_loss': pair.item(), 'smooth_loss': smooth.item(),
                       'kld_loss': kld.item(), 'feature_loss': 0.0}
    return total


def lambda_pair_warmup(epoch, warmup_epochs, lambda_pair_max):
    """vae_extent_search.py:1406-1410."""
    if warmup_epochs <= 0:
        return lambda_pair_max
    return lambda_pair_max * (min(max(epoch, 0), warmup_epochs) / warmup_epochs)


class RegressionTrainer:
    """optimizer = AdamW([{encoder params, encoder_lr}, {predictor params, cost_predictor_lr}], weight_decay=1e-5)
    (vae_extent_search.py:1389-1398) + the step of train_regression (:1424-1438)."""
    LOSS_NAMES = ("total", "reg_loss", "pair_loss", "smooth_loss", "kld_loss")

    def __init__(self, model, encoder_lr, cost_predictor_lr, weight_decay=1e-5, encoder_freeze=False, max_norm=1.0,
                 device="cuda:0", group=None):
        if getattr(model, "use_feature", False):
            raise NotImplementedError("feature_predictor head is outside the hot path")
        self.lib, self.model, self.group = _lib.load(), model, group
        self.device = torch.device(device)
        D, H, L, PH = model._dims()
        self.L, self.PH, self.p_drop = L, PH, float(model.cost_predictor[2].p)
        self.h = _lib.vp()
        _lib.check(self.lib.vaex_reg_trainer_create(D, H, L, PH, ctypes.byref(self.h)))
        self.n = int(self.lib.vaex_reg_trainer_param_count(self.h))
        self.slices = {name: self._offset(name) for name in model.state_dict()}
        self.groups = {"encoder": self._offset("@encoder_group"), "predictor": self._offset("@predictor_group")}
        self.lrs = {"encoder": float(encoder_lr), "predictor": float(cost_predictor_lr)}
        self.weight_decay, self.encoder_freeze, self.max_norm = float(weight_decay), bool(encoder_freeze), float(max_norm)
        z = lambda: torch.zeros(self.n, device=self.device)   # noqa: E731
        self.params, self.grads, self.m, self.v = z(), z(), z(), z()
        self.losses = torch.zeros(5, device=self.device)
        self.gnorm = torch.zeros(1, device=self.device)
        self.t = 0
        self.import_params()

    def _offset(self, name):
        o, c = _lib.i64(), _lib.i64()
        _lib.check(self.lib.vaex_reg_trainer_param_offset(self.h, name.encode(), ctypes.byref(o), ctypes.byref(c)))
        return o.value, c.value

    def import_params(self):
        with torch.no_grad():
            for name, t in self.model.state_dict().items():
                o, c = self.slices[name]
                assert c == t.numel(), name
                self.params[o:o + c].copy_(t.detach().reshape(-1).to(self.device, torch.float32))

    def export_params(self):
        with torch.no_grad():
            sd = self.model.state_dict()
            for name, (o, c) in self.slices.items():
                sd[name].copy_(self.params[o:o + c].reshape(sd[name].shape).to(sd[name].device))
        return self.model

    def forward_backward(self, x, labels, config, drop_mask=None, noise=None, cost_out=None):
        """Fills self.grads and self.losses = [total, reg, pair, smooth, kld] (device; nothing is read back)."""
        x, labels = _f(x), _f(labels).view(-1)
        cfg = (ctypes.c_float * 6)(config['lambda_reg'], config['lambda_pair'], config.get('margin', 0.1), config['gamma'],
                                   config['beta'], self.p_drop)
        mk = None if drop_mask is None else drop_mask.to(device=x.device, dtype=torch.uint8).contiguous()
        nz = None if noise is None else _f(noise)
        _lib.check(self.lib.vaex_reg_trainer_forward_backward(
            self.h, _lib.ptr(self.params), _lib.ptr(self.grads), _lib.ptr(x), _lib.ptr(labels), x.shape[0], _lib.ptr(mk),
            _lib.ptr(nz), cfg, _lib.ptr(self.losses), _lib.ptr(cost_out), _lib.stream_ptr()))
        return self.losses

    def optimizer_step(self, grad_scale=1.0, betas=(0.9, 0.999), eps=1e-8):
        """clip_grad_norm_(parameters, max_norm) then AdamW.step() over the two groups."""
        self.t += 1
        if self.encoder_freeze:                      # frozen parameters have no .grad: they do not enter the norm
            o, c = self.groups["encoder"]
            self.grads[o:o + c].zero_()
        first = True
        for name in ("encoder", "predictor"):
            if name == "encoder" and self.encoder_freeze:
                continue
            o, c = self.groups[name]
            _lib.check(self.lib.vaex_adamw_step(_lib.ptr(self.params), _lib.ptr(self.grads), _lib.ptr(self.m), _lib.ptr(self.v),
                                                self.n, o, c, self.t, self.lrs[name], betas[0], betas[1], eps, self.weight_decay,
                                                self.max_norm, grad_scale, _lib.ptr(self.gnorm), int(first), _lib.stream_ptr()))
            first = False

    def step(self, x, labels, config, drop_mask="auto", noise=None):
        """One iteration of train_regression's inner loop.  drop_mask="auto": a fresh Bernoulli(1 - p) keep mask while
        the module is in train mode (the reference's smooth_loss_fn leaves it in eval mode for the rest of the epoch,
        training.py:173), none otherwise; noise=None draws noise_std * randn."""
        x = _f(x)
        if isinstance(drop_mask, str):
            drop_mask = (torch.rand(x.shape[0], self.PH, device=x.device) >= self.p_drop) if self.model.training else None
        if noise is None:
            noise = config.get('noise_std', 0.1) * torch.randn(x.shape[0], self.L, device=x.device)
        self.forward_backward(x, labels, config, drop_mask, noise)
        self.model.eval()                            # smooth_loss_fn's side effect
        scale = shard.reduce_gradients(self.grads, self.group, getattr(self, "comm", None))
        self.optimizer_step(grad_scale=scale)
        return self.losses

    def components(self):
        vals = self.losses.cpu().tolist()
        return dict(zip(self.LOSS_NAMES, vals))

    def close(self):
        if self.h:
            self.lib.vaex_reg_trainer_destroy(self.h)
            self.h = _lib.vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def train_regression(vae_cost_model, trainer, train_loader, config, warmup_epochs=0):
    """The training part of train_regression (vae_extent_search.py:1413-1440): `trainer` is a RegressionTrainer
    (it plays the optimizer's role).  Returns the model with the trained parameters written back."""
    lambda_pair_max = config['lambda_pair']
    cfg = dict(config)
    for epoch in range(1, config['epochs'] + 1):
        vae_cost_model.train()
        for x_batch, labels in train_loader:
            cfg['lambda_pair'] = lambda_pair_warmup(epoch, warmup_epochs, lambda_pair_max) if warmup_epochs else lambda_pair_max
            trainer.step(x_batch.to(trainer.device), labels.to(trainer.device).squeeze(-1), cfg)
    return trainer.export_params()


def vae_feature_loss(x_recon, x, mu, logvar, feature_pred=None, feature=None, alpha_recon=0, alpha_feature=0, beta=1.0):
    """training.py:14-32, values only (no feature head on the hot path) -> (loss, recon_loss, kl, feature_loss)."""
    xr, xx = _f(x_recon).reshape(-1), _f(x).reshape(-1)
    recon = ((xr - xx) ** 2).mean()
    kl = kld_loss_fn(mu, logvar)
    return alpha_recon * recon + beta * kl, recon, kl, 0.0


class VAETrainer:
    """optimizer = torch.optim.Adam(vae.parameters(), lr) + the inner loop of vae_train (training.py:65,70-88):
    forward(x, use_mean=False) -> vae_feature_loss -> backward -> Adam.step, as two CUDA calls."""

    def __init__(self, vae, lr=1e-3, alpha_recon=1.0, beta=0.01, device="cuda:0", group=None):
        if getattr(vae, "use_feature", False):
            raise NotImplementedError("feature_predictor head (alpha_feature experiments) is outside the hot path")
        self.lib, self.model, self.group = _lib.load(), vae, group
        self.device = torch.device(device)
        D, H, L = vae._dims()
        self.L, self.lr, self.alpha_recon, self.beta = L, float(lr), float(alpha_recon), float(beta)
        self.h = _lib.vp()
        _lib.check(self.lib.vaex_vae_trainer_create(D, H, L, ctypes.byref(self.h)))
        self.n = int(self.lib.vaex_vae_trainer_param_count(self.h))
        self.slices = {}
        for name in vae.state_dict():
            o, c = _lib.i64(), _lib.i64()
            _lib.check(self.lib.vaex_vae_trainer_param_offset(self.h, name.encode(), ctypes.byref(o), ctypes.byref(c)))
            self.slices[name] = (o.value, c.value)
        z = lambda: torch.zeros(self.n, device=self.device)   # noqa: E731
        self.params, self.grads, self.m, self.v = z(), z(), z(), z()
        self.losses = torch.zeros(3, device=self.device)
        self.t = 0
        self.import_params()

    import_params = RegressionTrainer.import_params
    export_params = RegressionTrainer.export_params

    def forward_backward(self, x, eps=None):
        """eps [n, L]: the randn of reparameterize; None = use_mean.  Fills self.grads and self.losses = [total, recon, kl]."""
        x = _f(x)
        e = None if eps is None else _f(eps)
        _lib.check(self.lib.vaex_vae_trainer_forward_backward(self.h, _lib.ptr(self.params), _lib.ptr(self.grads), _lib.ptr(x),
                                                              x.shape[0], _lib.ptr(e), self.alpha_recon, self.beta,
                                                              _lib.ptr(self.losses), _lib.stream_ptr()))
        return self.losses

    def step(self, x, eps="auto"):
        x = _f(x)
        if isinstance(eps, str):
            eps = torch.randn(x.shape[0], self.L, device=x.device)
        self.forward_backward(x, eps)
        scale = shard.reduce_gradients(self.grads, self.group, getattr(self, "comm", None))
        self.t += 1
        _lib.check(self.lib.vaex_adam_step(_lib.ptr(self.params), _lib.ptr(self.grads), _lib.ptr(self.m), _lib.ptr(self.v), self.n,
                                           self.t, self.lr, 0.9, 0.999, 1e-8, 0.0, 0.0, scale, None, _lib.stream_ptr()))
        return self.losses

    def close(self):
        if self.h:
            self.lib.vaex_vae_trainer_destroy(self.h)
            self.h = _lib.vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
