"""Training of the segment-family SegmentVAE on the GPU (segment family).

Mirror of VAE_Trainer (scripts/vae_experiments/trainer.py:23-118, loss :270-280): the train-mode forward
(batch-statistics BatchNorm x3, decoder Dropout, reparameterisation), the sum-reduced loss, the backward pass and
clip + AdamW run as two calls into libvaex_b200.so (vaex_segvae_trainer_forward_backward, vaex_adamw_step); the
epoch loop, the cosine warm-restart schedule, validation R^2 and early stopping are host logic as in the reference.
CUDA tensors only -- no CPU fallback."""
import copy
import ctypes
import math

import numpy as np
import torch

from . import _lib, shard
from .seg_models import SegmentVAE


def _f(t):
    if not t.is_cuda:
        raise _lib.VaexError("vaex_b200 needs CUDA tensors (no CPU fallback)")
    return t.detach().contiguous().float()


def vae_loss(recon, original, mean, logvar, beta):
    """trainer.py:270-280, values only -> (total, recon_loss, kld_loss)."""
    recon_loss = ((_f(recon) - _f(original)) ** 2).sum()
    kld_loss = -0.5 * torch.sum(1 + _f(logvar) - _f(mean).pow(2) - _f(logvar).exp())
    return recon_loss + beta * kld_loss, recon_loss, kld_loss


def cosine_warm_restarts_lr(base_lr, epoch, T_0=30, T_mult=2, eta_min=0.0):
    """torch.optim.lr_scheduler.CosineAnnealingWarmRestarts(T_0, T_mult) after `epoch` scheduler.step() calls."""
    T_i, T_cur = T_0, epoch
    while T_cur >= T_i:
        T_cur -= T_i
        T_i *= T_mult
    return eta_min + (base_lr - eta_min) * (1 + math.cos(math.pi * T_cur / T_i)) / 2


_BN_KEYS = ("norm", "decoder.1", "decoder.5")


class SegmentVAETrainer:
    """optimizer = AdamW(model.parameters(), lr, weight_decay=1e-5) + the inner loop of train_vae (trainer.py:40-66):
    forward -> vae_loss -> backward -> clip_grad_norm_(1.0) -> step."""

    def __init__(self, model, lr=1e-3, beta=0.1, weight_decay=1e-5, max_norm=1.0, device="cuda:0", group=None):
        self.lib, self.model, self.group = _lib.load(), model, group
        self.device = torch.device(device)
        F, H, L = model._dims()
        self.H, self.L, self.p_drop = H, L, float(model.decoder[3].p)
        self.lr, self.beta, self.weight_decay, self.max_norm = float(lr), float(beta), float(weight_decay), float(max_norm)
        self.h = _lib.vp()
        _lib.check(self.lib.vaex_segvae_trainer_create(F, H, L, ctypes.byref(self.h)))
        self.n = int(self.lib.vaex_segvae_trainer_param_count(self.h))
        self.slices = {}
        for name, _ in model.named_parameters():
            o, c = _lib.i64(), _lib.i64()
            _lib.check(self.lib.vaex_segvae_trainer_param_offset(self.h, name.encode(), ctypes.byref(o), ctypes.byref(c)))
            self.slices[name] = (o.value, c.value)
        z = lambda: torch.zeros(self.n, device=self.device)   # noqa: E731
        self.params, self.grads, self.m, self.v = z(), z(), z(), z()
        self.bn_running = torch.zeros(3, 2, H, device=self.device)
        self.losses = torch.zeros(3, device=self.device)
        self.gnorm = torch.zeros(1, device=self.device)
        self.t = 0
        self.batches_tracked = 0
        self.import_params()

    def import_params(self):
        sd = self.model.state_dict()
        with torch.no_grad():
            for name, (o, c) in self.slices.items():
                assert c == sd[name].numel(), name
                self.params[o:o + c].copy_(sd[name].detach().reshape(-1).to(self.device, torch.float32))
            for i, k in enumerate(_BN_KEYS):
                self.bn_running[i, 0].copy_(sd[k + ".running_mean"])
                self.bn_running[i, 1].copy_(sd[k + ".running_var"])

    def export_params(self):
        """Writes parameters and BatchNorm running statistics back into the module."""
        sd = self.model.state_dict()
        with torch.no_grad():
            for name, (o, c) in self.slices.items():
                sd[name].copy_(self.params[o:o + c].reshape(sd[name].shape).to(sd[name].device))
            for i, k in enumerate(_BN_KEYS):
                sd[k + ".running_mean"].copy_(self.bn_running[i, 0])
                sd[k + ".running_var"].copy_(self.bn_running[i, 1])
                sd[k + ".num_batches_tracked"].add_(self.batches_tracked)
        self.batches_tracked = 0
        return self.model

    def forward_backward(self, segment_sizes, features, eps=None, drop_mask=None, drop_seed=0, want_outputs=False):
        """eps [P, L]: the randn of reparameterize (None: z = mean, logvar = 0).  drop_mask uint8 [P, H] (1 = keep) or
        drop_seed (Philox) or neither.  Fills self.grads, self.losses = [total, recon, kld], updates the running stats."""
        feats = _f(features)
        sizes = segment_sizes.detach().to(device=feats.device, dtype=torch.int32).contiguous()
        P = sizes.shape[0]
        off = torch.empty(P + 1, dtype=torch.int32, device=feats.device)
        st = _lib.stream_ptr()
        _lib.check(self.lib.vaex_segment_offsets(_lib.ptr(sizes), P, _lib.ptr(off), st))
        e = None if eps is None else _f(eps)
        mk = None if drop_mask is None else drop_mask.to(device=feats.device, dtype=torch.uint8).contiguous()
        mean = recon = None
        if want_outputs:
            mean, recon = torch.empty(P, self.L, device=feats.device), torch.empty(P, self.H, device=feats.device)
        _lib.check(self.lib.vaex_segvae_trainer_forward_backward(
            self.h, _lib.ptr(self.params), _lib.ptr(self.grads), _lib.ptr(feats), _lib.ptr(off), P, feats.shape[0], _lib.ptr(e),
            _lib.ptr(mk), int(drop_seed), self.p_drop, self.beta, _lib.ptr(self.bn_running), _lib.ptr(self.losses), _lib.ptr(mean),
            _lib.ptr(recon), st))
        self.batches_tracked += 1
        return (self.losses, mean, recon) if want_outputs else self.losses

    def optimizer_step(self, grad_scale=1.0, lr=None):
        """clip_grad_norm_(model.parameters(), max_norm) then AdamW.step()."""
        self.t += 1
        _lib.check(self.lib.vaex_adamw_step(_lib.ptr(self.params), _lib.ptr(self.grads), _lib.ptr(self.m), _lib.ptr(self.v), self.n,
                                            0, self.n, self.t, self.lr if lr is None else float(lr), 0.9, 0.999, 1e-8,
                                            self.weight_decay, self.max_norm, grad_scale, _lib.ptr(self.gnorm), 1, _lib.stream_ptr()))

    def step(self, segment_sizes, features, eps="auto", drop="auto", lr=None):
        P = segment_sizes.shape[0]
        if isinstance(eps, str):
            eps = torch.randn(P, self.L, device=self.device)
        seed, mask = 0, None
        if isinstance(drop, str):
            seed = int(torch.randint(1, 2 ** 62, (1,)).item()) if self.p_drop > 0 else 0
        else:
            mask = drop
        self.forward_backward(segment_sizes, features, eps, mask, seed)
        scale = shard.reduce_gradients(self.grads, self.group, getattr(self, "comm", None))
        self.optimizer_step(scale, lr)
        return self.losses

    def close(self):
        if self.h:
            self.lib.vaex_segvae_trainer_destroy(self.h)
            self.h = _lib.vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class VAE_Trainer:
    """scripts/vae_experiments/trainer.py:23-118: same constructor and train_vae signature / return tuple."""

    def __init__(self, train_loader, val_loader, input_dim=164, hidden_dim=256, latent_dim=128, dropout=0.1, device='cuda'):
        self.train_loader, self.val_loader = train_loader, val_loader
        self.input_dim, self.hidden_dim, self.latent_dim, self.dropout = input_dim, hidden_dim, latent_dim, dropout
        self.device = "cuda:0" if device == "cuda" else device

    vae_loss = staticmethod(vae_loss)

    def train_vae(self, epochs=200, lr=1e-3, beta=0.1, patience=30, verbose=False):
        from sklearn.metrics import r2_score
        model = SegmentVAE(self.input_dim, self.hidden_dim, self.latent_dim, self.dropout).to(self.device)
        trainer = SegmentVAETrainer(model, lr=lr, beta=beta, weight_decay=1e-5, device=self.device)
        best_recon_r2, best_kl_per_dim, patience_counter, best_state = -float('inf'), None, 0, None
        history = {'recon_r2': [], 'kl_per_dim': [], 'recon_loss': []}
        for epoch in range(epochs):
            cur_lr = cosine_warm_restarts_lr(lr, epoch)
            epoch_recon, n_batches = 0.0, 0
            for segment_sizes_batch, features_batch, _ in self.train_loader:
                losses = trainer.step(segment_sizes_batch.to(self.device), features_batch.to(self.device), lr=cur_lr)
                epoch_recon += float(losses[1]) / segment_sizes_batch.shape[0]
                n_batches += 1
            trainer.export_params()
            model.eval()
            val_recons, val_originals, val_kl, val_batches = [], [], 0.0, 0
            with torch.no_grad():
                for segment_sizes_batch, features_batch, _ in self.val_loader:
                    mean, logvar, z, recon, segsum = model(segment_sizes_batch.to(self.device), features_batch.to(self.device),
                                                           use_mean=True)
                    kld = -0.5 * torch.sum(1 + logvar - mean.pow(2) - logvar.exp())
                    val_kl += kld.item() / (mean.shape[0] * self.latent_dim)
                    val_batches += 1
                    val_recons.append(recon.cpu().numpy())
                    val_originals.append(segsum.cpu().numpy())
            val_kl /= max(val_batches, 1)
            recon_r2 = r2_score(np.concatenate(val_originals).flatten(), np.concatenate(val_recons).flatten())
            history['recon_r2'].append(recon_r2)
            history['kl_per_dim'].append(val_kl)
            history['recon_loss'].append(epoch_recon / max(n_batches, 1))
            if recon_r2 > best_recon_r2:
                best_recon_r2, best_kl_per_dim, patience_counter = recon_r2, val_kl, 0
                best_state = copy.deepcopy(model.state_dict())
            else:
                patience_counter += 1
            if verbose and (epoch + 1) % 50 == 0:
                print(f"  Epoch {epoch+1}: R2={recon_r2:.4f}, KL/dim={val_kl:.4f}")
            if patience_counter >= patience:
                break
        model.load_state_dict(best_state)
        trainer.close()
        return model, best_recon_r2, best_kl_per_dim, history


# ------------------------------------------------------------------ regression on the segment-family VAECostPredictor
REG_LOSS_NAMES = ("total_loss", "reg_loss", "pair_loss", "smooth_loss", "kld_loss")


class SegmentRegressionTrainer:
    """optimizer = AdamW([{encoder params, encoder_lr}, {predictor params, predictor_lr}], weight_decay=1e-5) + one
    iteration of train_phase's mini-epoch loop (scripts/vae_experiments/trainer.py:401-455): model.train() ->
    compute_total_loss (:538-556) -> backward -> clip_grad_norm_(1.0) -> step."""

    def __init__(self, model, encoder_lr, predictor_lr, weight_decay=1e-5, max_norm=1.0, device="cuda:0", group=None):
        self.lib, self.model, self.group = _lib.load(), model, group
        self.device = torch.device(device)
        F, H, L, PH, NL = model._dims()
        self.H, self.L, self.PH, self.NL = H, L, PH, NL
        drops = [m for m in model.cost_predictor if isinstance(m, torch.nn.Dropout)]
        self.p_drop = float(drops[0].p) if drops else 0.0
        self.lrs = {"encoder": float(encoder_lr), "predictor": float(predictor_lr)}
        self.weight_decay, self.max_norm = float(weight_decay), float(max_norm)
        self.h = _lib.vp()
        _lib.check(self.lib.vaex_segreg_trainer_create(F, H, L, PH, NL, ctypes.byref(self.h)))
        self.n = int(self.lib.vaex_segreg_trainer_param_count(self.h))
        self.slices = {name: self._offset(name) for name, _ in model.named_parameters()}
        self.groups = {"encoder": self._offset("@encoder_group"), "predictor": self._offset("@predictor_group")}
        self.bn_keys = ["norm"] + ["cost_predictor.%d" % (4 * i + 1) for i in range(NL)]
        z = lambda: torch.zeros(self.n, device=self.device)   # noqa: E731
        self.params, self.grads, self.m, self.v = z(), z(), z(), z()
        self.bn_running = torch.zeros(2 * H + NL * 2 * PH, device=self.device)
        self.losses = torch.zeros(5, device=self.device)
        self.gnorm = torch.zeros(1, device=self.device)
        self.t = 0
        self.steps_tracked = 0
        self.import_params()

    def _offset(self, name):
        o, c = _lib.i64(), _lib.i64()
        _lib.check(self.lib.vaex_segreg_trainer_param_offset(self.h, name.encode(), ctypes.byref(o), ctypes.byref(c)))
        return o.value, c.value

    def _bn_views(self):
        H, PH = self.H, self.PH
        views = [(self.bn_running[0:H], self.bn_running[H:2 * H])]
        for i in range(self.NL):
            b = 2 * H + i * 2 * PH
            views.append((self.bn_running[b:b + PH], self.bn_running[b + PH:b + 2 * PH]))
        return views

    def import_params(self):
        sd = self.model.state_dict()
        with torch.no_grad():
            for name, (o, c) in self.slices.items():
                assert c == sd[name].numel(), name
                self.params[o:o + c].copy_(sd[name].detach().reshape(-1).to(self.device, torch.float32))
            for k, (rm, rv) in zip(self.bn_keys, self._bn_views()):
                rm.copy_(sd[k + ".running_mean"])
                rv.copy_(sd[k + ".running_var"])

    def export_params(self):
        sd = self.model.state_dict()
        with torch.no_grad():
            for name, (o, c) in self.slices.items():
                sd[name].copy_(self.params[o:o + c].reshape(sd[name].shape).to(sd[name].device))
            for j, (k, (rm, rv)) in enumerate(zip(self.bn_keys, self._bn_views())):
                sd[k + ".running_mean"].copy_(rm)
                sd[k + ".running_var"].copy_(rv)
                sd[k + ".num_batches_tracked"].add_(self.steps_tracked * (1 if j == 0 else 3))   # three predictor passes per step
        self.steps_tracked = 0
        return self.model

    def forward_backward(self, segment_sizes, features, labels, config, noise, drop_masks=None, drop_seed=0, cost_out=None):
        """noise [P, L] = noise_std * randn; drop_masks uint8 [3, NL - 1, P, PH] (1 = keep) or drop_seed (Philox) or neither.
        Fills self.grads and self.losses = [total, reg, pair, smooth, kld]."""
        feats = _f(features)
        sizes = segment_sizes.detach().to(device=feats.device, dtype=torch.int32).contiguous()
        P = sizes.shape[0]
        off = torch.empty(P + 1, dtype=torch.int32, device=feats.device)
        st = _lib.stream_ptr()
        _lib.check(self.lib.vaex_segment_offsets(_lib.ptr(sizes), P, _lib.ptr(off), st))
        if config.get('loss_type', 'mse') != 'mse':
            raise NotImplementedError("loss_type 'mae' is not on the GPU path")
        cfg = (ctypes.c_float * 5)(config.get('lambda_pair', 0.1), config.get('margin', 0.1), config.get('gamma', 0.01),
                                   config.get('beta', 0.001), self.p_drop)
        mk = None if drop_masks is None else drop_masks.to(device=feats.device, dtype=torch.uint8).contiguous()
        _lib.check(self.lib.vaex_segreg_trainer_forward_backward(
            self.h, _lib.ptr(self.params), _lib.ptr(self.grads), _lib.ptr(feats), _lib.ptr(off), _lib.ptr(_f(labels).view(-1)), P,
            feats.shape[0], _lib.ptr(_f(noise)), _lib.ptr(mk), int(drop_seed), cfg, _lib.ptr(self.bn_running), _lib.ptr(self.losses),
            _lib.ptr(cost_out), st))
        self.steps_tracked += 1
        return self.losses

    def optimizer_step(self, grad_scale=1.0):
        self.t += 1
        first = True
        for name in ("encoder", "predictor"):
            o, c = self.groups[name]
            _lib.check(self.lib.vaex_adamw_step(_lib.ptr(self.params), _lib.ptr(self.grads), _lib.ptr(self.m), _lib.ptr(self.v),
                                                self.n, o, c, self.t, self.lrs[name], 0.9, 0.999, 1e-8, self.weight_decay,
                                                self.max_norm, grad_scale, _lib.ptr(self.gnorm), int(first), _lib.stream_ptr()))
            first = False

    def step(self, segment_sizes, features, labels, config, noise=None, drop="auto"):
        P = segment_sizes.shape[0]
        if noise is None:
            noise = config.get('noise_std', 0.1) * torch.randn(P, self.L, device=self.device)
        seed, masks = 0, None
        if isinstance(drop, str):
            seed = int(torch.randint(1, 2 ** 62, (1,)).item()) if self.p_drop > 0 else 0
        else:
            masks = drop
        self.forward_backward(segment_sizes, features, labels, config, noise, masks, seed)
        scale = shard.reduce_gradients(self.grads, self.group, getattr(self, "comm", None))
        self.optimizer_step(scale)
        return self.losses

    def components(self):
        return dict(zip(REG_LOSS_NAMES, self.losses.cpu().tolist()))

    def close(self):
        if self.h:
            self.lib.vaex_segreg_trainer_destroy(self.h)
            self.h = _lib.vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Regression_Trainer:
    """scripts/vae_experiments/trainer.py:296-483: same constructor (a dict with the pre-trained SegmentVAE and its dims),
    train_regression / train_phase signatures and return values."""

    def __init__(self, vae_model, device='cuda'):
        from .seg_models import VAECostPredictor
        self.vae_model = vae_model
        self.device = "cuda:0" if device == "cuda" else device
        self.model = VAECostPredictor(input_dim=vae_model['input_dim'], hidden_dim=vae_model['hidden_dim'],
                                      latent_dim=vae_model['latent_dim'], predictor_hidden=256, predictor_layers=3,
                                      dropout=0.1).to(self.device)
        self.model.load_pretrained_encoder(vae_model['model'].state_dict())

    def train_regression(self, train_feature_list, train_segment_sizes, train_labels, fea_norm_vec, config, phases=1):
        self.model, history, best_loss = self.train_phase(
            self.model, train_feature_list, train_segment_sizes, train_labels, fea_norm_vec, self.device, config,
            samples_per_phase=len(train_labels) // phases if phases > 0 else len(train_labels),
            mini_epochs=config.get('mini_epochs', 10), phases=phases)
        return history, best_loss

    def train_phase(self, model, train_feature_list, train_segment_sizes, train_labels, fea_norm_vec, device, config,
                    samples_per_phase=64, mini_epochs=10, phases=1):
        all_indices = np.arange(len(train_labels))
        best_loss, best_state, history = float('inf'), None, []
        trainer = SegmentRegressionTrainer(model, config['encoder_lr'], config['predictor_lr'], weight_decay=1e-5, device=device)
        for phase in range(phases):
            start = phase * samples_per_phase
            idx = all_indices[start:min(start + samples_per_phase, len(all_indices))].tolist()
            flat = np.concatenate([train_feature_list[i] for i in idx], axis=0).astype(np.float32)
            sizes = torch.tensor(np.asarray(train_segment_sizes)[idx], dtype=torch.int32).to(device)
            feats = torch.tensor(flat, dtype=torch.float32).to(device)
            labels = torch.tensor(np.asarray(train_labels)[idx], dtype=torch.float32).to(device)
            if fea_norm_vec is not None:
                feats = feats / fea_norm_vec.to(device)
            for epoch in range(mini_epochs):
                model.train()
                trainer.step(sizes, feats, labels, config)
                if (epoch + 1) % 5 == 0 or epoch == mini_epochs - 1:
                    comp = trainer.components()
                    history.append({'phase': phase + 1, 'epoch': epoch + 1, **comp})
                    if comp['total_loss'] < best_loss:
                        best_loss = comp['total_loss']
                        best_state = copy.deepcopy(trainer.export_params().state_dict())
        trainer.export_params()
        if best_state is not None:
            model.load_state_dict(best_state)
        trainer.close()
        return model, history, best_loss
