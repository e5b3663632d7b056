"""Drop-in mirror of the TenSet MLP cost model
(reference: python/tvm/auto_scheduler/cost_model/mlp_model.py).

Same class names, constructor arguments, attribute names, state_dict keys and pickle tuple as
the reference; the bodies hand device pointers to the hand-written sm_100a kernels behind
libvaex_b200.so.  Nothing here computes on the CPU: without the library or a CUDA device the
calls raise.

  SegmentDataLoader   mlp_model.py:26-144   -> flatten_features() + on-device offsets
  SegmentSumMLPModule mlp_model.py:147-221  -> vaex_segsum_mlp_forward
  MLPModelInternal    mlp_model.py:340-804  -> predict / save / load; fit_base / fit_local / fine-tune on the
                                             CUDA training kernels (train.py)
  MLPModel            mlp_model.py:814-860  -> predict with the -inf rule (:842-844)
"""
import io
import logging
import pickle

import time

import ctypes

import numpy as np
import torch

from . import _lib
from .dataset import Dataset, LearningTask

logger = logging.getLogger("auto_scheduler")


def get_workload_embedding(workload_key):
    """xgb_model.py:79-87: keyword one-hot (9 wide) over the printed ComputeDAG of the
    workload.  Printing the DAG needs TVM; without it the tags are matched on the key itself
    (all-zero for hashed keys, which is what the synthetic configs use)."""
    tags = ['max', 'min', 'add', 'Conv2dOutput', 'conv2d_winograd', 'DepthwiseConv2d', 'dense', 'softmax', 'compute(b, i, j)']
    try:
        from tvm.auto_scheduler.compute_dag import ComputeDAG
        from tvm.auto_scheduler.workload_registry import workload_key_to_tensors
        dag_str = str(ComputeDAG(workload_key_to_tensors(workload_key)))
    except Exception:
        dag_str = str(workload_key)
    vec = [0] * len(tags)
    for i, t in enumerate(tags):
        if t in dag_str:
            vec[i] = 1
    return np.array(vec, dtype=np.float32)


def flatten_features(features, task_embedding=None):
    """SegmentDataLoader.__init__ (mlp_model.py:47-92) without the per-row Python loop:
    object array of [S_i, F] -> (float32 [sum S, F(+E)], int32 sizes[N])."""
    n = len(features)
    sizes = np.fromiter((len(r) for r in features), dtype=np.int32, count=n)
    if n == 0:
        return np.zeros((0, 0), np.float32), sizes
    flat = np.concatenate([np.asarray(r, dtype=np.float32) for r in features], axis=0)
    if task_embedding is not None:
        emb = np.broadcast_to(np.asarray(task_embedding, np.float32), (flat.shape[0], len(task_embedding)))
        flat = np.concatenate([flat, emb], axis=1)
    return np.ascontiguousarray(flat, dtype=np.float32), sizes


class SegmentDataLoader:
    """Inference-side mirror of mlp_model.py:26-144: same constructor, `features`,
    `segment_sizes`, `normalize`, iteration in batches -- but batches are views of one
    device-resident tensor (no per-batch gather / H2D)."""

    def __init__(self, dataset, batch_size, device, use_workload_embedding=True, use_target_embedding=False,
                 target_id_dict={}, fea_norm_vec=None, shuffle=False):
        self.device = device
        self.shuffle = shuffle
        self.batch_size = batch_size
        flats, sizes, labels = [], [], []
        for task in dataset.features:
            emb = None
            if use_workload_embedding or use_target_embedding:
                emb = np.zeros(10 if use_workload_embedding else 0, dtype=np.float32)
                if use_workload_embedding:
                    emb[:9] = get_workload_embedding(task.workload_key)
                if use_target_embedding:
                    tid = target_id_dict.get(str(task.target), np.random.randint(0, max(len(target_id_dict), 1)))
                    emb[9 + tid] = 1.0
            f, s = flatten_features(dataset.features[task], emb)
            flats.append(f)
            sizes.append(s)
            thr = dataset.throughputs[task]
            labels.append(np.zeros(len(s), np.float32) if thr is None else np.asarray(thr, np.float32))
        self.segment_sizes = torch.from_numpy(np.concatenate(sizes) if sizes else np.zeros(0, np.int32))
        self.labels = torch.from_numpy(np.concatenate(labels) if labels else np.zeros(0, np.float32))
        self.number = len(self.segment_sizes)
        self.features = torch.from_numpy(np.concatenate(flats, axis=0)) if flats else torch.zeros(0, 0)
        if fea_norm_vec is not None:
            self.normalize(fea_norm_vec)
        self.feature_offsets = (torch.cumsum(self.segment_sizes, 0, dtype=torch.int32) - self.segment_sizes).numpy()
        self.iter_order = self.pointer = None

    def normalize(self, norm_vector=None):
        if norm_vector is None:
            mx = self.features.max(dim=0).values
            norm_vector = torch.where(mx > 0, mx, torch.ones_like(mx))
        self.features /= norm_vector
        return norm_vector

    def __iter__(self):
        """mlp_model.py:107-114: a fresh permutation per epoch when shuffling (the same torch CPU stream as the
        reference, so a seeded run visits the programs in the reference's order)."""
        self.iter_order = torch.randperm(self.number) if self.shuffle else None
        self.pointer = 0
        return self

    def __next__(self):
        if self.pointer >= self.number:
            raise StopIteration
        s, e = self.pointer, min(self.pointer + self.batch_size, self.number)
        self.pointer += self.batch_size
        if self.iter_order is not None:
            return self._fetch_indices(self.iter_order[s:e])
        r0 = int(self.feature_offsets[s])
        r1 = int(self.feature_offsets[e - 1] + self.segment_sizes[e - 1])
        return (self.segment_sizes[s:e].to(self.device), self.features[r0:r1].to(self.device),
                self.labels[s:e].to(self.device))

    def _fetch_indices(self, indices):
        """mlp_model.py:129-141 without the per-program Python loop: the rows of programs `indices`, in that order."""
        indices = torch.as_tensor(indices, dtype=torch.int64)
        segment_sizes = self.segment_sizes[indices]
        sz = segment_sizes.numpy().astype(np.int64)
        starts = np.asarray(self.feature_offsets, dtype=np.int64)[indices.numpy()]
        ends_out = np.cumsum(sz)
        row_idx = np.arange(int(ends_out[-1]) if len(sz) else 0, dtype=np.int64) + np.repeat(starts - (ends_out - sz), sz)
        features = self.features[torch.from_numpy(row_idx)]
        return (segment_sizes.to(self.device), features.to(self.device), self.labels[indices].to(self.device))

    def __len__(self):
        return self.number


class RowsImage:
    """A pool's feature rows packed once for repeated scoring (vaex_rows_pack_image): fp16 [n_rows, 192] with the
    bias-one columns in place + the per-row non-zero bits.  SegmentSumMLPModule.forward / segment_sums take it in place
    of the fp32 feature tensor; the row encoder then loads its operand tiles with TMA.  Same scores bit for bit."""

    def __init__(self, features, in_dim=None):
        if not features.is_cuda:
            raise _lib.VaexError("RowsImage needs a CUDA tensor (no CPU fallback)")
        feats = _lib.check_device(features).detach().contiguous().float()
        self.n_rows, self.in_dim = int(feats.shape[0]), int(in_dim or feats.shape[1])
        lib = _lib.load()
        nbytes = int(lib.vaex_rows_image_bytes(self.n_rows))
        self.data = torch.empty(max(nbytes, 128), dtype=torch.uint8, device=feats.device)
        _lib.check(lib.vaex_rows_pack_image(_lib.ptr(feats), self.n_rows, self.in_dim, _lib.ptr(self.data), _lib.stream_ptr()))

    @property
    def device(self):
        return self.data.device


class SegmentSumMLPModule(torch.nn.Module):
    """mlp_model.py:147-221.  Parameters live in the same nn.Linear modules (identical
    state_dict keys, so reference pickles / checkpoints load); forward() runs the fused
    tcgen05 row-encoder + segment-sum kernel and the residual head on the GPU."""

    def __init__(self, in_dim, hidden_dim, out_dim, use_norm=False, add_sigmoid=False):
        super().__init__()
        assert out_dim == 1, "the cost model has a scalar output"
        self.segment_encoder = torch.nn.Sequential(
            torch.nn.Linear(in_dim, hidden_dim), torch.nn.ReLU(),
            torch.nn.Linear(hidden_dim, hidden_dim), torch.nn.ReLU())
        self.add_sigmoid = add_sigmoid
        self.norm = torch.nn.BatchNorm1d(hidden_dim) if use_norm else torch.nn.Identity()
        self.l0 = torch.nn.Sequential(torch.nn.Linear(hidden_dim, hidden_dim), torch.nn.ReLU())
        self.l1 = torch.nn.Sequential(torch.nn.Linear(hidden_dim, hidden_dim), torch.nn.ReLU())
        self.decoder = torch.nn.Linear(hidden_dim, out_dim)
        self._in_dim, self._hidden = in_dim, hidden_dim
        self._packed = {}     # version key -> _lib.Model

    def freeze_for_fine_tuning(self):
        for x in self.segment_encoder.parameters():
            x.requires_grad_(False)

    # -- packing ---------------------------------------------------------------------------
    def _version_key(self, norm_vec):
        """Cache key of the packed weights: storage address + autograd version per tensor (id() can be reused after garbage
        collection) and the norm vector's bytes.  No device read: the key must cost nothing on the scoring path and be legal
        under CUDA-graph capture.  Writes through .data do not bump _version: call invalidate_packed() after them."""
        key = tuple((p.data_ptr(), p._version)
                    for p in self.state_dict(keep_vars=True).values())
        nk = None
        if norm_vec is not None:
            nv = norm_vec.detach().cpu().numpy() if hasattr(norm_vec, "detach") else np.asarray(norm_vec)
            nk = hash(np.ascontiguousarray(nv, dtype=np.float32).tobytes())
        return key, nk

    def invalidate_packed(self):
        self._packed = {}

    def packed(self, fea_norm_vec=None, embedding=None):
        """Returns the packed kernel-side model (re-packed when weights / norm vector change).

        embedding: the constant per-task vector SegmentDataLoader tiles onto every row
        (mlp_model.py:77-79).  Its contribution W0[:, F:] @ (emb / norm[F:]) is the same for every row, so
        it is folded into the first bias and the packed model takes the raw F-wide rows: no host-side
        concat, 6 % fewer bytes over PCIe, and the all-zero-feature rule (mlp_model.py:842-844) sees exactly
        the columns the reference tests."""
        # older pickles made by the reference class do not carry the private attributes
        in_dim = self.segment_encoder[0].in_features
        hidden = self.segment_encoder[0].out_features
        ek = None if embedding is None else tuple(np.asarray(embedding, np.float32).tolist())
        key = self._version_key(fea_norm_vec) + (ek,)
        cache = getattr(self, "_packed", None)
        if not isinstance(cache, dict):
            cache = self._packed = {}
        if key in cache:
            return cache[key]
        use_norm = isinstance(self.norm, torch.nn.BatchNorm1d)
        sd, nv = self.state_dict(), fea_norm_vec
        if ek is not None and len(ek):
            E, F = len(ek), in_dim - len(ek)
            if F <= 0:
                raise ValueError("embedding is wider than the model's input")
            w0 = sd["segment_encoder.0.weight"].detach().cpu().double().numpy()
            b0 = sd["segment_encoder.0.bias"].detach().cpu().double().numpy()
            emb = np.asarray(ek, np.float64)
            if nv is not None:
                nv = np.asarray(nv.detach().cpu() if hasattr(nv, "detach") else nv, np.float64).reshape(-1)
                emb, nv = emb / nv[F:], nv[:F].astype(np.float32)
            sd = dict(sd)
            sd["segment_encoder.0.weight"] = torch.from_numpy(np.ascontiguousarray(w0[:, :F]).astype(np.float32))
            sd["segment_encoder.0.bias"] = torch.from_numpy((b0 + w0[:, F:] @ emb).astype(np.float32))
            in_dim = F
        m = _lib.Model(_lib.KIND_SEGSUM_MLP, [in_dim, hidden, int(use_norm), int(bool(self.add_sigmoid))])
        m.load_state(sd, nv)
        if len(cache) >= 8:                    # a handful of tasks alternate during a search round
            cache.pop(next(iter(cache)))
        cache[key] = m
        return m

    def __getstate__(self):
        d = self.__dict__.copy()
        d["_packed"] = {}
        return d

    # -- forward ---------------------------------------------------------------------------
    def forward(self, segment_sizes, features, params=None, flags=0, fea_norm_vec=None):
        if self.training and torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
            # the reference trains by autograd through this forward (mlp_model.py:548-565); here the forward, the loss and
            # the backward of a step are one C-ABI call on flat buffers, not an autograd graph
            raise _lib.VaexError("SegmentSumMLPModule.forward builds no autograd graph: train with "
                                 "MLPModelInternal.fit_base / fit_local or vaex_b200.train.Trainer "
                                 "(vaex_trainer_forward_backward), or call under torch.no_grad() / .eval() to score")
        if isinstance(features, RowsImage):
            feats, n_rows, flags = features.data, features.n_rows, flags | _lib.F_FEATS_IMAGE
        elif not features.is_cuda:
            raise _lib.VaexError("SegmentSumMLPModule.forward needs CUDA tensors (no CPU fallback)")
        elif features.dtype == torch.float16:      # rows already narrowed (the tcgen05 kernel rounds to fp16 first anyway)
            feats, flags = _lib.check_device(features).detach().contiguous(), flags | _lib.F_FEATS_F16
            n_rows = feats.shape[0]
        else:
            feats = _lib.check_device(features).detach().contiguous().float()
            n_rows = feats.shape[0]
        lib = _lib.load()
        m = self.packed(fea_norm_vec)
        sizes = segment_sizes.detach().to(device=feats.device, dtype=torch.int32).contiguous()
        n = sizes.shape[0]
        offsets = torch.empty(n + 1, dtype=torch.int32, device=feats.device)
        scores = torch.empty(n, dtype=torch.float32, device=feats.device)
        st = _lib.stream_ptr()
        _lib.check(lib.vaex_segment_offsets(_lib.ptr(sizes), n, _lib.ptr(offsets), st))
        _lib.check(lib.vaex_segsum_mlp_forward(m.h, _lib.ptr(feats), _lib.ptr(offsets), n, n_rows,
                                                _lib.ptr(scores), flags, st))
        return scores.squeeze()

    def segment_sums(self, segment_sizes, features, flags=0, fea_norm_vec=None):
        """[n_programs, hidden] per-program sums of the row encoder's output -- the tensor
        `scatter_add(...)` produces at mlp_model.py:200-208 (and SegmentVAE.segment_sum,
        scripts/vae_experiments/models/vae.py:68-85)."""
        if isinstance(features, RowsImage):
            feats, n_rows, flags = features.data, features.n_rows, flags | _lib.F_FEATS_IMAGE
        elif not features.is_cuda:
            raise _lib.VaexError("SegmentSumMLPModule.segment_sums needs CUDA tensors (no CPU fallback)")
        else:
            feats = _lib.check_device(features).detach().contiguous().float()
            n_rows = feats.shape[0]
        lib = _lib.load()
        m = self.packed(fea_norm_vec)
        sizes = segment_sizes.detach().to(device=feats.device, dtype=torch.int32).contiguous()
        n = sizes.shape[0]
        offsets = torch.empty(n + 1, dtype=torch.int32, device=feats.device)
        out = torch.empty(n, self.segment_encoder[0].out_features, dtype=torch.float32, device=feats.device)
        st = _lib.stream_ptr()
        _lib.check(lib.vaex_segment_offsets(_lib.ptr(sizes), n, _lib.ptr(offsets), st))
        _lib.check(lib.vaex_segsum_encode(m.h, _lib.ptr(feats), _lib.ptr(offsets), n, n_rows, _lib.ptr(out), flags, st))
        return out


def make_net(params):
    if params["type"] == "SegmentSumMLP":
        return SegmentSumMLPModule(params["in_dim"], params["hidden_dim"], params["out_dim"],
                                   add_sigmoid=params["add_sigmoid"])
    raise ValueError("Invalid type: " + params["type"] + " (only SegmentSumMLP is on the B200 path)")


class MLPModelInternal:
    """mlp_model.py:340-804: same hyper-parameter attributes and predict/save/load behaviour."""

    def __init__(self, device=None, few_shot_learning="base_only", use_workload_embedding=True,
                 use_target_embedding=False, loss_type='lambdaRankLoss'):
        if device is None:
            device = 'cuda:0'
        if not str(device).startswith('cuda'):
            raise _lib.VaexError("vaex_b200 runs on CUDA devices only (no CPU fallback)")
        self.net_params = {"type": "SegmentSumMLP", "in_dim": 164 + (10 if use_workload_embedding else 0),
                           "hidden_dim": 256, "out_dim": 1}
        self.target_id_dict = {}
        self.loss_type = loss_type
        self.n_epoch, self.lr = 100, 7e-4
        if loss_type == 'rmse':
            self.net_params['add_sigmoid'] = True
        elif loss_type == 'rankNetLoss':
            self.net_params['add_sigmoid'] = False
            self.n_epoch = 30
        elif loss_type == 'lambdaRankLoss':
            self.net_params['add_sigmoid'] = False
            self.n_epoch = 50
        elif loss_type == 'listNetLoss':
            self.net_params['add_sigmoid'] = False
            self.lr, self.n_epoch = 9e-4, 50
        else:
            raise ValueError("Invalid loss type: " + loss_type)
        self.grad_clip = 0.5
        self.few_shot_learning = few_shot_learning
        self.fea_norm_vec = None
        self.use_workload_embedding = use_workload_embedding
        self.use_target_embedding = use_target_embedding
        self.batch_size, self.infer_batch_size = 512, 4096
        self.wd = 1e-6
        self.device = device
        self.print_per_epoches = 5
        self.fine_tune_lr, self.fine_tune_batch_size = 4e-2, 512
        self.fine_tune_num_steps, self.fine_tune_wd = 10, 0
        self.base_model = None
        self.local_model = {}

    def register_new_task(self, task):
        target = str(task.target)
        if target not in self.target_id_dict:
            self.target_id_dict[target] = len(self.target_id_dict)

    # -- training (mlp_model.py:420-489,512-661) on the CUDA training kernels -----------------------
    def fit_base(self, train_set, valid_set=None, valid_train_set=None):
        if self.few_shot_learning == "local_only":
            self.base_model = None
        elif self.few_shot_learning == "MAML":
            raise NotImplementedError("MAML raises NotImplemented in the reference as well (mlp_model.py:424-428)")
        else:
            self.base_model = self._fit_a_model(train_set, valid_set, valid_train_set)

    def fit_local(self, train_set, valid_set=None):
        fsl = self.few_shot_learning
        if fsl == "base_only":
            return
        if fsl == "local_only_mix_task":
            local_model = self._fit_a_model(train_set, valid_set)
            for task in train_set.tasks():
                self.local_model[task] = local_model
        elif fsl == "local_only_per_task":
            for task in train_set.tasks():
                self.local_model[task] = self._fit_a_model(train_set.extract_subset([task]), valid_set)
        elif fsl in ("plus_mix_task", "plus_per_task"):
            if fsl == "plus_mix_task":
                self.net_params["hidden_dim"] = 128
                self.loss_type = 'rmse'
                self.net_params['add_sigmoid'] = True

            def diff_set(ds):
                base = self._predict_a_dataset(self.base_model, ds)
                out = Dataset()
                for task in ds.tasks():
                    out.load_task_data(task, ds.features[task], ds.throughputs[task] - base[task])
                return out
            if fsl == "plus_mix_task":
                diff_model = self._fit_a_model(diff_set(train_set), diff_set(valid_set) if valid_set else None)
                for task in train_set.tasks():
                    self.local_model[task] = diff_model
            else:
                for task in train_set.tasks():
                    self.local_model[task] = self._fit_a_model(diff_set(train_set.extract_subset([task])), valid_set)
        elif fsl == "fine_tune_mix_task":
            self.base_model = self._fine_tune_a_model(self.base_model, train_set, valid_set)
        else:
            raise ValueError("Invalid few-shot learning method: " + fsl)

    def _loaders(self, train_set, valid_set, batch_size, shuffle):
        for task in train_set.tasks():
            self.register_new_task(task)
        train_loader = SegmentDataLoader(train_set, batch_size, self.device, self.use_workload_embedding,
                                         self.use_target_embedding, self.target_id_dict, shuffle=shuffle)
        if self.fea_norm_vec is None:
            self.fea_norm_vec = train_loader.normalize()
        else:
            train_loader.normalize(self.fea_norm_vec)
        valid_loader = None
        if valid_set:
            for task in valid_set.tasks():
                self.register_new_task(task)
            valid_loader = SegmentDataLoader(valid_set, self.infer_batch_size, self.device, self.use_workload_embedding,
                                             self.use_target_embedding, self.target_id_dict, fea_norm_vec=self.fea_norm_vec)
        return train_loader, valid_loader

    def _epoch(self, trainer, data, batch_size, shuffle, lr, wd, clip, optimizer, avg):
        """One pass over the resident set in one C-ABI call (Trainer.epoch -> vaex_trainer_epoch); the running loss
        (moving_average, momentum 0.95) stays on the device and is read back once by the caller."""
        order = torch.randperm(data.n) if shuffle else torch.arange(data.n)       # same CPU stream as the reference
        return trainer.epoch(data, order, batch_size, lr, wd, clip, optimizer, avg)

    def _fit_a_model(self, train_set, valid_set=None, valid_train_set=None, n_epoch=None):
        from . import train as _train
        _lib.require_cuda("MLPModelInternal._fit_a_model")
        print("=" * 60 + "\nFit a net. Train size: %d" % len(train_set))
        train_loader, valid_loader = self._loaders(train_set, valid_set, self.batch_size, True)
        n_epoch = n_epoch or self.n_epoch
        early_stop = n_epoch // 6
        net = make_net(self.net_params).to(self.device)
        trainer = _train.Trainer(net, self.loss_type, self.device)
        data = _train.ResidentSet(train_loader, self.device)
        train_loss, best_epoch, best_train_loss = None, None, 1e10
        for epoch in range(n_epoch):
            tic = time.time()
            train_loss, batch = self._epoch(trainer, data, self.batch_size, True, self.lr, self.wd, self.grad_clip, "adam",
                                            train_loss)
            loss_now = float(train_loss.item())                       # the only device read of the epoch
            train_time = time.time() - tic
            if epoch % self.print_per_epoches == 0 or epoch == n_epoch - 1:
                valid_loss = self._validate(trainer.export_params(), valid_loader, trainer) if valid_loader else 0.0
                if self.loss_type == "rmse":
                    loss_msg = "Train RMSE: %.4f\tValid RMSE: %.4f" % (np.sqrt(loss_now), np.sqrt(valid_loss))
                else:
                    loss_msg = "Train Loss: %.4f\tValid Loss: %.4f" % (loss_now, valid_loss)
                print("Epoch: %d\tBatch: %d\t%s\tTrain Speed: %.0f" % (epoch, batch - 1, loss_msg, len(train_loader) / train_time))
            if loss_now < best_train_loss:
                best_train_loss, best_epoch = loss_now, epoch
            elif epoch - best_epoch >= early_stop:
                print("Early stop. Best epoch: %d" % best_epoch)
                break
        net = trainer.export_params()
        trainer.close()
        return net.eval()

    def _fine_tune_a_model(self, model, train_set, valid_set=None, verbose=1):
        from . import train as _train
        _lib.require_cuda("MLPModelInternal._fine_tune_a_model")
        if verbose >= 1:
            print("=" * 60 + "\nFine-tune a net. Train size: %d" % len(train_set))
        bs = self.fine_tune_batch_size or len(train_set)
        train_loader, valid_loader = self._loaders(train_set, valid_set, bs, False)
        trainer = _train.Trainer(model.to(self.device), self.loss_type, self.device)
        data = _train.ResidentSet(train_loader, self.device)
        tic = time.time()
        for step in range(self.fine_tune_num_steps):
            train_loss, _ = self._epoch(trainer, data, bs, False, self.fine_tune_lr, self.fine_tune_wd, 0.0, "sgd", None)
            if verbose >= 1:
                valid_loss = self._validate(trainer.export_params(), valid_loader, trainer) if valid_loader else 0
                loss_now = float(train_loss.item())
                if self.loss_type == "rmse":
                    loss_msg = "Train RMSE: %.4f\tValid RMSE: %.4f" % (np.sqrt(loss_now), np.sqrt(valid_loss))
                else:
                    loss_msg = "Train Loss: %.4f\tValid Loss: %.4f" % (loss_now, valid_loss)
                print("Fine-tune step: %d\t%s\tTime: %.1f" % (step, loss_msg, time.time() - tic))
        model = trainer.export_params()
        trainer.close()
        return model.eval()

    def _validate(self, model, valid_loader, trainer=None):
        """mlp_model.py:653-661: mean over infer_batch_size batches of loss_func(model(batch), labels);
        the forward is the fp32 check path so the validation loss is comparable with the training loss."""
        from . import train as _train
        own = trainer is None
        if own:
            trainer = _train.Trainer(model, self.loss_type, self.device)
        model.eval()
        losses = []
        for segment_sizes, features, labels in valid_loader:
            with torch.no_grad():
                preds = model(segment_sizes, features, flags=_lib.F_FORCE_SIMT)
            losses.append(trainer.loss_value(preds.reshape(-1), labels))
        if own:
            trainer.close()
        return float(torch.cat(losses).mean().item()) if losses else 0.0

    # -- inference ---------------------------------------------------------------------------
    def _task_embedding(self, task):
        if not (self.use_workload_embedding or self.use_target_embedding):
            return None
        emb = np.zeros(10 if self.use_workload_embedding else 0, dtype=np.float32)
        if self.use_workload_embedding:
            emb[:9] = get_workload_embedding(task.workload_key)
        if self.use_target_embedding:
            tid = self.target_id_dict.get(str(task.target), np.random.randint(0, max(len(self.target_id_dict), 1)))
            emb[9 + tid] = 1.0
        return emb

    def _predict_a_task(self, model, task, features, flags=0):
        """mlp_model.py:669-681.  One flatten on the host, then the C-ABI host call (pinned
        staging + chunked H2D overlapped with the kernels)."""
        if model is None:
            return np.zeros(len(features), dtype=np.float32)
        flat, sizes = flatten_features(features)
        return predict_flat(model, flat, sizes, self.fea_norm_vec, flags, self._task_embedding(task))

    def predict_blob(self, task, byte_arr, flags=_lib.F_MASK_INVALID):
        """Scores straight from the bytes of TVM's GetPerStoreFeaturesFromStates (feature.py:224-246):
        the blob is parsed in C++ into pinned staging and streamed to the GPU (vaex_segsum_mlp_predict_blob).
        -> np.float32[records]; -inf for records without features (mlp_model.py:842-844)."""
        model = self.base_model
        if self.few_shot_learning in ("local_only_mix_task", "local_only_per_task"):
            model = self.local_model[task]
        elif self.few_shot_learning in ("plus_mix_task", "plus_per_task"):
            raise NotImplementedError("predict_blob covers the single-model modes; use predict() for plus_*")
        from . import feature
        keep, addr, nbytes = feature._as_bytes(byte_arr)
        n = feature.scan(keep)[0]
        if model is None:
            return np.zeros(n, dtype=np.float32)
        m = model.packed(self.fea_norm_vec, self._task_embedding(task))
        out = np.empty(n, dtype=np.float32)
        _lib.check(_lib.load().vaex_segsum_mlp_predict_blob(m.h, addr, nbytes, out.ctypes.data, n, flags))
        return out

    def _predict_a_dataset(self, model, dataset, flags=0):
        return {task: self._predict_a_task(model, task, feats, flags) for task, feats in dataset.features.items()}

    def predict(self, dataset):
        fsl = self.few_shot_learning
        if fsl in ["base_only", "fine_tune_mix_task", "fine_tune_per_task", "MAML"]:
            return self._predict_a_dataset(self.base_model, dataset)
        if fsl in ["local_only_mix_task", "local_only_per_task"]:
            return {t: self._predict_a_task(self.local_model[t], t, dataset.features[t]) for t in dataset.tasks()}
        if fsl in ["plus_mix_task", "plus_per_task"]:
            base = self._predict_a_dataset(self.base_model, dataset)
            ret = {}
            for t in dataset.tasks():
                if t not in self.local_model and fsl == "plus_mix_task":
                    self.local_model[t] = list(self.local_model.values())[0]
                ret[t] = base[t] + self._predict_a_task(self.local_model[t], t, dataset.features[t])
            return ret
        raise ValueError("Invalid few show learing: " + fsl)

    # -- checkpoints (same pickle tuple as mlp_model.py:788-804) -------------------------------
    def load(self, filename):
        self.base_model, self.local_model, self.few_shot_learning, self.fea_norm_vec = \
            _RefUnpickler(open(filename, 'rb')).load()

    def save(self, filename):
        base = self.base_model.cpu() if self.base_model else None
        local = self.local_model.cpu() if self.local_model and hasattr(self.local_model, "cpu") else self.local_model
        pickle.dump((base, local, self.few_shot_learning, self.fea_norm_vec), open(filename, 'wb'))
        # mlp_model.py:803-804: the models go back to the device after the dump
        dev = getattr(self, "device", None)
        if dev is not None and torch.cuda.is_available():
            self.base_model = self.base_model.to(dev) if self.base_model else None
            if self.local_model and hasattr(self.local_model, "to"):
                self.local_model = self.local_model.to(dev)


class _RefUnpickler(pickle.Unpickler):
    """Loads pickles written by the reference (class path tvm.auto_scheduler.cost_model.
    mlp_model.*) onto this module's classes, storages mapped to CPU (mlp_model.py:806-811)."""

    def find_class(self, module, name):
        if module == 'torch.storage' and name == '_load_from_bytes':
            return lambda b: torch.load(io.BytesIO(b), map_location='cpu', weights_only=False)
        if module.endswith("cost_model.mlp_model") and name in globals():
            return globals()[name]
        return super().find_class(module, name)


def predict_flat(model, flat, sizes, fea_norm_vec=None, flags=0, embedding=None):
    """Host fp32 [rows, F] + int32 sizes -> np.float32[N] through vaex_segsum_mlp_predict_host."""
    lib = _lib.load()
    m = model.packed(fea_norm_vec, embedding)
    flat = np.ascontiguousarray(flat, dtype=np.float32)
    sizes = np.ascontiguousarray(sizes, dtype=np.int32)
    out = np.empty(len(sizes), dtype=np.float32)
    _lib.check(lib.vaex_segsum_mlp_predict_host(m.h, flat.ctypes.data, sizes.ctypes.data, len(sizes),
                                                 flat.shape[0], out.ctypes.data, flags))
    return out


def pack_features_f16(flat):
    """Host fp32 rows -> (np.float16 array of the same shape, lossy_zero flag): the rounding the tcgen05 row encoder
    applies, done ahead of the copy (vaex_pack_features_f16).  Feed the result to SegmentSumMLPModule.forward as a
    CUDA half tensor."""
    flat = np.ascontiguousarray(flat, dtype=np.float32)
    out = np.empty(flat.shape, dtype=np.float16)
    lossy = ctypes.c_int32(0)
    _lib.check(_lib.load().vaex_pack_features_f16(flat.ctypes.data, out.ctypes.data, flat.size, ctypes.addressof(lossy)))
    return out, bool(lossy.value)


try:        # with TVM importable the model is a PythonBasedModel, so task_scheduler.py:94-115 / SketchPolicy accept it
    from tvm.auto_scheduler.cost_model.cost_model import PythonBasedModel
except ImportError:
    class PythonBasedModel:
        """Stand-in base when TVM is absent: cost_model.py:115-130 registers update / predict / predict_stages
        as PackedFuncs with the C++ PythonBasedModelNode; without TVM there is nothing to register with."""

        def __init__(self):
            pass

        def predict_stages(self, task, states):
            raise NotImplementedError       # cost_model.py:166-201: MLPModel does not implement it either


class MLPModel(PythonBasedModel):
    """mlp_model.py:814-860.  `predict(task, states)` needs TVM's feature extractor
    (get_per_store_features_from_states); `predict_features(task, features)` is the same call
    from the point where the 164-d rows exist -- the boundary of this build."""

    def __init__(self, few_shot_learning="base_only", disable_update=False):
        super().__init__()
        self.disable_update = disable_update
        self.model = MLPModelInternal(few_shot_learning=few_shot_learning)
        self.dataset = Dataset()

    def update(self, inputs, results):
        """mlp_model.py:824-830: append the measured pairs to the data set and re-fit the base model on all of it
        (the CUDA training kernels, train.py)."""
        if self.disable_update or len(inputs) <= 0:
            return
        tic = time.time()
        self.dataset.update_from_measure_pairs(inputs, results)
        self.model.fit_base(self.dataset)
        logger.info("MLPModel Training time: %.2f s", time.time() - tic)

    def update_from_file(self, file_name, n_lines=None):
        """mlp_model.py:848-851.  Reading measure records is TVM's RecordReader."""
        try:
            from tvm.auto_scheduler.measure_record import RecordReader
        except ImportError as e:
            raise _lib.VaexError("update_from_file needs TVM's RecordReader: %s" % e)
        inputs, results = RecordReader(file_name).read_lines(n_lines)
        logger.info("MLPModel: Loaded %s measurement records from %s", len(inputs), file_name)
        self.update(inputs, results)

    def predict_features(self, task, features):
        learning_task = LearningTask(task.workload_key, str(task.target))
        if self.model is not None:
            ret = self.model._predict_a_task(self.model.base_model, learning_task, features,
                                             flags=_lib.F_MASK_INVALID if self.model.base_model is not None else 0)
            if self.model.base_model is None:
                for idx, feature in enumerate(features):
                    if feature.min() == feature.max() == 0:
                        ret[idx] = float('-inf')
        else:
            ret = np.random.uniform(0, 1, (len(features),))
        return ret

    def predict_blob(self, task, byte_arr):
        """predict() from the packed bytes of GetPerStoreFeaturesFromStates: no per-row Python unpacking."""
        learning_task = LearningTask(task.workload_key, str(task.target))
        if self.model is None:
            from . import feature
            return np.random.uniform(0, 1, (feature.scan(byte_arr)[0],))
        if self.model.base_model is not None or self.model.few_shot_learning.startswith("local_only"):
            return self.model.predict_blob(learning_task, byte_arr)
        from . import feature
        flat, sizes, _, _, _, _ = feature.unpack_feature_flat(byte_arr)
        ret = np.zeros(len(sizes), dtype=np.float32)
        if len(sizes):
            starts = np.concatenate([[0], np.cumsum(sizes[:-1], dtype=np.int64)])
            ret[np.maximum.reduceat(np.abs(flat).max(axis=1), starts) == 0] = float('-inf')
        return ret

    def predict_extents(self, task, features_of_sketch, extents, flags=_lib.F_MASK_INVALID):
        """predict(task, states) (mlp_model.py:832-846) for candidates of ONE sketch given by their tile extents: the rows
        get_per_store_features_from_states would extract through TVM come from the sketch's template on the GPU
        (sketch_features.SketchFeatures), and never leave it.  extents: int32 [n, slots] (CUDA tensor or array).
        -> np.float32[n] scores (random scores without a trained model, like predict).  flags: VAEX_F_* of the row encoder."""
        from . import sketch_features as SF
        if not isinstance(features_of_sketch, SF.SketchFeatures):
            features_of_sketch = SF.SketchFeatures(features_of_sketch)
        e = extents if isinstance(extents, torch.Tensor) else torch.as_tensor(np.asarray(extents, dtype=np.int32))
        e = e.to(device=features_of_sketch.program.device, dtype=torch.int32)
        n = e.shape[0]
        if self.model is None or n == 0:
            return np.random.uniform(0, 1, (n,))
        learning_task = LearningTask(task.workload_key, str(task.target))
        inner = self.model
        net = inner.local_model[learning_task] if inner.few_shot_learning.startswith("local_only") else inner.base_model
        if net is None:
            return np.zeros(n, dtype=np.float32)
        rows = features_of_sketch.rows(e)
        R = features_of_sketch.n_rows
        offsets = torch.arange(0, (n + 1) * R, R, dtype=torch.int32, device=rows.device)
        scores = torch.empty(n, device=rows.device)
        m = net.packed(inner.fea_norm_vec, inner._task_embedding(learning_task))
        _lib.check(_lib.load().vaex_segsum_mlp_forward(m.h, _lib.ptr(rows), _lib.ptr(offsets), n, rows.shape[0], _lib.ptr(scores),
                                                      flags, _lib.stream_ptr()))
        return scores.cpu().numpy()

    def predict(self, task, states):
        """mlp_model.py:832-846.  Needs TVM for the feature extraction itself; the packed bytes it returns go
        straight to the C++ parser (feature.py:224-246 without unpack_feature)."""
        from tvm.auto_scheduler import _ffi_api
        from tvm.auto_scheduler.feature import DEFAULT_MAX_N_BUFS
        from tvm.auto_scheduler.loop_state import State
        objs = [s.state_object for s in states] if isinstance(states[0], State) else states
        return self.predict_blob(task, _ffi_api.GetPerStoreFeaturesFromStates(objs, task, DEFAULT_MAX_N_BUFS))

    def save(self, file_name):
        self.model.save(file_name)

    def load(self, file_name):
        if self.model is None:
            self.model = MLPModelInternal()
        self.model.load(file_name)
        self.num_warmup_sample = -1
