"""GPU training of the SegmentSumMLP cost model (reference: MLPModelInternal._fit_a_model /
_fine_tune_a_model / _validate, python/tvm/auto_scheduler/cost_model/mlp_model.py:512-661).

`Trainer` owns flat fp32 device buffers (parameters in state_dict order, gradients, Adam moments)
and drives the CUDA training kernels behind libvaex_b200.so: one call does forward + loss +
backward, one does clip + Adam.  Data-parallel training (one process per GPU) all-reduces the
flat gradient buffer over NCCL between the two calls -- 0.96 MB for the 239 873-parameter model.
The whole training set stays device-resident; shuffled mini-batches are assembled by a gather
kernel from the epoch's permutation (no per-batch host gather / H2D as in _fetch_indices)."""
import ctypes

import numpy as np
import torch
import torch.distributed as dist

from . import _lib, shard


class Trainer:
    def __init__(self, net, loss_type="lambdaRankLoss", device="cuda:0", group=None):
        self.lib = _lib.load()
        if loss_type not in _lib.LOSS_IDS:
            raise ValueError("Invalid loss type: " + loss_type)
        if not isinstance(net.norm, torch.nn.Identity):
            raise NotImplementedError("train-mode BatchNorm is not on the CUDA training path (MLPModelInternal never sets use_norm)")
        self.net, self.loss_id, self.group = net, _lib.LOSS_IDS[loss_type], group
        self.device = torch.device(device)
        self.F, self.H = net.segment_encoder[0].in_features, net.segment_encoder[0].out_features
        self.h = _lib.vp()
        _lib.check(self.lib.vaex_trainer_create(self.F, self.H, int(bool(net.add_sigmoid)), ctypes.byref(self.h)))
        self.n = int(self.lib.vaex_trainer_param_count(self.h))
        self.slices = {}
        for name, t in net.state_dict().items():
            o, c = _lib.i64(), _lib.i64()
            _lib.check(self.lib.vaex_trainer_param_offset(self.h, name.encode(), ctypes.byref(o), ctypes.byref(c)))
            assert c.value == t.numel(), name
            self.slices[name] = (o.value, c.value)
        self.params = torch.zeros(self.n, device=self.device)
        self.grads = torch.zeros_like(self.params)
        self.m = torch.zeros_like(self.params)
        self.v = torch.zeros_like(self.params)
        self.gnorm = torch.zeros(1, device=self.device)
        self.loss = torch.zeros(1, device=self.device)
        self.t = 0
        self.import_params()

    # -- parameters <-> module ---------------------------------------------------------------
    def import_params(self):
        with torch.no_grad():
            for name, t in self.net.state_dict().items():
                o, c = self.slices[name]
                self.params[o:o + c].copy_(t.detach().reshape(-1).to(self.device, torch.float32))

    def export_params(self):
        """Writes the trained flat buffer back into the module's parameters."""
        with torch.no_grad():
            sd = self.net.state_dict()
            for name, (o, c) in self.slices.items():
                sd[name].copy_(self.params[o:o + c].reshape(sd[name].shape).to(sd[name].device))
        return self.net

    def world(self):
        return dist.get_world_size(self.group) if dist.is_available() and dist.is_initialized() else 1

    # -- one optimisation step -----------------------------------------------------------------
    def forward_backward(self, offsets, features, labels, preds=None):
        """offsets int32 [n+1], features fp32 [rows, F], labels fp32 [n] (all CUDA).  Fills
        self.grads and self.loss (device scalars; nothing is read back)."""
        n, rows = labels.shape[0], features.shape[0]
        _lib.check(self.lib.vaex_trainer_forward_backward(
            self.h, _lib.ptr(self.params), _lib.ptr(self.grads), _lib.ptr(features), _lib.ptr(offsets), _lib.ptr(labels),
            n, rows, self.loss_id, _lib.ptr(self.loss), _lib.ptr(preds), _lib.stream_ptr()))
        return self.loss

    def all_reduce_grads(self):
        """Sum over ranks; the mean is folded into the optimizer kernel (grad_scale = 1 / world)."""
        return shard.reduce_gradients(self.grads, self.group, getattr(self, "comm", None))

    def adam_step(self, lr, weight_decay=0.0, grad_clip=0.0, betas=(0.9, 0.999), eps=1e-8, grad_scale=1.0):
        self.t += 1
        _lib.check(self.lib.vaex_adam_step(_lib.ptr(self.params), _lib.ptr(self.grads), _lib.ptr(self.m), _lib.ptr(self.v),
                                           self.n, self.t, lr, betas[0], betas[1], eps, weight_decay, grad_clip or 0.0,
                                           grad_scale, _lib.ptr(self.gnorm), _lib.stream_ptr()))

    def sgd_step(self, lr, weight_decay=0.0, grad_scale=1.0):
        _lib.check(self.lib.vaex_sgd_step(_lib.ptr(self.params), _lib.ptr(self.grads), self.n, lr, weight_decay, grad_scale,
                                          _lib.stream_ptr()))

    def step(self, offsets, features, labels, lr, weight_decay=0.0, grad_clip=0.0, optimizer="adam"):
        """loss.backward(); clip_grad_norm_; optimizer.step() of mlp_model.py:559-565 (Adam) / :636-638 (SGD)."""
        self.forward_backward(offsets, features, labels)
        scale = self.all_reduce_grads()
        if optimizer == "adam":
            self.adam_step(lr, weight_decay, grad_clip, grad_scale=scale)
        else:
            self.sgd_step(lr, weight_decay, grad_scale=scale)
        return self.loss

    def epoch(self, data, order, batch_size, lr, weight_decay=0.0, grad_clip=0.0, optimizer="adam", avg=None):
        """One pass over the resident set `data` in the order `order` (int64 CPU tensor) as ONE C-ABI call
        (vaex_trainer_epoch): gather, forward + loss + backward, all-reduce (self.comm, a shard.Comm, when set), clip +
        Adam / SGD and the moving-average loss for every batch, no Python between the launches.
        -> (moving-average loss as a device scalar, number of batches)."""
        order = order.to(torch.int64).contiguous()
        order_dev = order.to(self.device)
        first = avg is None
        if first:
            avg = torch.zeros(1, device=self.device)
        step, nb = _lib.i64(self.t), _lib.i64(0)
        comm = getattr(self, "comm", None)
        if comm is None and self.world() > 1:
            raise _lib.VaexError("data-parallel epochs run through the library's communicator: set trainer.comm = shard.Comm()")
        _lib.check(self.lib.vaex_trainer_epoch(
            self.h, comm.h if comm is not None else None, _lib.ptr(self.params), _lib.ptr(self.grads), _lib.ptr(self.m), _lib.ptr(self.v),
            _lib.ptr(data.features), _lib.ptr(data.offsets), data.sizes_i32.ctypes.data, _lib.ptr(data.labels),
            order.data_ptr(), _lib.ptr(order_dev), order.shape[0], int(batch_size), self.loss_id, 0 if optimizer == "adam" else 1,
            float(lr), float(weight_decay), float(grad_clip or 0.0), ctypes.byref(step), _lib.ptr(avg), int(first), ctypes.byref(nb),
            _lib.stream_ptr()))
        self.t = int(step.value)
        self._keep = (order, order_dev)              # alive until the stream has consumed them
        return avg, int(nb.value)

    def loss_value(self, preds, labels):
        """self.loss_func(preds, labels) on the GPU (<= 4096 predictions) -> device scalar."""
        out = torch.zeros(1, device=preds.device)
        _lib.check(self.lib.vaex_loss_eval(self.loss_id, _lib.ptr(preds.contiguous().float()), _lib.ptr(labels.contiguous().float()),
                                           preds.shape[0], _lib.ptr(out), None, _lib.stream_ptr()))
        return out

    def close(self):
        if self.h:
            self.lib.vaex_trainer_destroy(self.h)
            self.h = _lib.vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ResidentSet:
    """A training / validation set kept on the device: features [rows, F] (already normalised),
    int32 offsets [n+1], labels [n]; `batch(indices)` assembles a mini-batch with the gather kernel."""

    def __init__(self, loader, device):
        self.device = torch.device(device)
        self.sizes_host = loader.segment_sizes.numpy().astype(np.int64)
        self.sizes_i32 = np.ascontiguousarray(loader.segment_sizes.numpy().astype(np.int32))
        self.n = len(self.sizes_host)
        self.features = loader.features.to(self.device, torch.float32).contiguous()
        self.labels = loader.labels.to(self.device, torch.float32).contiguous()
        sizes = loader.segment_sizes.to(self.device, torch.int32).contiguous()
        self.offsets = torch.empty(self.n + 1, dtype=torch.int32, device=self.device)
        _lib.check(_lib.load().vaex_segment_offsets(_lib.ptr(sizes), self.n, _lib.ptr(self.offsets), _lib.stream_ptr()))
        self.F = self.features.shape[1]

    def batch(self, indices_host, indices_dev):
        """indices_*: the same int64 program indices on host (numpy) and device."""
        n = len(indices_host)
        rows = int(self.sizes_host[indices_host].sum())
        off = torch.empty(n + 1, dtype=torch.int32, device=self.device)
        feats = torch.empty(max(rows, 1), self.F, device=self.device)[:rows]
        labels = torch.empty(n, device=self.device)
        _lib.check(_lib.load().vaex_gather_segments(_lib.ptr(self.features), self.F, _lib.ptr(self.offsets), _lib.ptr(indices_dev), n,
                                                   _lib.ptr(off), _lib.ptr(feats), _lib.ptr(self.labels), _lib.ptr(labels),
                                                   _lib.stream_ptr()))
        return off, feats, labels
