"""GPU parity of the training row: CUDA losses, backward pass, clip + Adam / SGD against
the reference-generated goldens (tests/golden/train.npz) and the torch-autograd oracle."""
import numpy as np
import pytest
import torch

from conftest import load_golden, split_sd
from oracle import torch_oracle as O

import vaex_b200
from vaex_b200 import _lib, train

pytestmark = pytest.mark.gpu


def _offsets(sizes):
    return torch.from_numpy(np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)).cuda()


@pytest.mark.parametrize("n", [64, 257])
def test_losses_and_gradients_vs_reference(n):
    g = load_golden("train.npz")
    lib = _lib.load()
    preds, labels = g["loss_preds_%d" % n], g["loss_labels_%d" % n]
    for name, lid in _lib.LOSS_IDS.items():
        p = torch.from_numpy(preds).cuda()
        if name == "rmse":
            p = torch.sigmoid(p)
        loss, grad = torch.zeros(1, device="cuda"), torch.zeros(n, device="cuda")
        yd = torch.from_numpy(labels).cuda()
        _lib.check(lib.vaex_loss_eval(lid, _lib.ptr(p), _lib.ptr(yd), n, _lib.ptr(loss), _lib.ptr(grad), _lib.stream_ptr()))
        np.testing.assert_allclose(loss.item(), g["loss_%s_%d" % (name, n)], rtol=2e-5, err_msg=name)
        want = g["grad_%s_%d" % (name, n)]
        np.testing.assert_allclose(grad.cpu().numpy(), want, rtol=2e-4, atol=2e-6 * np.abs(want).max(), err_msg=name)


def test_loss_edge_cases():
    lib = _lib.load()
    for n in (1, 2, 1000, 4096):                       # single element, non-power-of-two, the 4096 limit
        rng = np.random.default_rng(n)
        p, y = rng.standard_normal(n).astype(np.float32), rng.random(n).astype(np.float32)
        for name, lid in _lib.LOSS_IDS.items():
            loss, grad = torch.zeros(1, device="cuda"), torch.zeros(n, device="cuda")
            pd, yd = torch.from_numpy(p).cuda(), torch.from_numpy(y).cuda()      # keep both alive across the launch
            _lib.check(lib.vaex_loss_eval(lid, _lib.ptr(pd), _lib.ptr(yd), n, _lib.ptr(loss), _lib.ptr(grad), _lib.stream_ptr()))
            wl, wg = O.loss_and_grad(name, p, y)
            np.testing.assert_allclose(loss.item(), float(wl), rtol=2e-4, atol=1e-6, err_msg="%s n=%d" % (name, n))
            np.testing.assert_allclose(grad.cpu().numpy(), wg.numpy(), rtol=1e-3, atol=1e-5 * max(1e-3, float(wg.abs().max())),
                                       err_msg="%s n=%d" % (name, n))
    loss = torch.zeros(1, device="cuda")
    with pytest.raises(_lib.VaexError):
        _lib.check(lib.vaex_loss_eval(2, _lib.ptr(loss), _lib.ptr(loss), 4097, _lib.ptr(loss), None, None))


@pytest.mark.parametrize("name", ["lambdaRankLoss", "rmse", "rankNetLoss", "listNetLoss"])
def test_train_trajectory_vs_reference(name):
    """six Adam steps (lr 7e-4, L2 1e-6, clip 0.5) on fixed batches: losses, pre-clip gradient norms and
    final parameters against the reference's own make_net / loss / torch.optim.Adam run."""
    g = load_golden("train.npz")
    sizes, feats, labels = g["train_sizes"], g["train_feats"], g["train_labels"]
    offs = np.concatenate([[0], np.cumsum(sizes)])
    net = vaex_b200.SegmentSumMLPModule(164, 64, 1, add_sigmoid=(name == "rmse"))
    net.load_state_dict(split_sd(g, "init."))
    tr = train.Trainer(net.cuda(), name)
    losses, norms = [], []
    for b in (0, 1, 2, 0, 1, 2):
        s, e = 48 * b, 48 * (b + 1)
        loss = tr.step(_offsets(sizes[s:e]), torch.from_numpy(feats[offs[s]:offs[e]]).cuda(), torch.from_numpy(labels[s:e]).cuda(),
                       7e-4, 1e-6, 0.5)
        losses.append(loss.item())
        norms.append(tr.gnorm.item())
    np.testing.assert_allclose(losses, g["traj_losses_%s" % name], rtol=2e-4)
    np.testing.assert_allclose(norms, g["traj_norms_%s" % name], rtol=2e-3)
    net = tr.export_params()
    for k, v in split_sd(g, "final_%s." % name).items():
        got = net.state_dict()[k].cpu().numpy()
        # Adam turns every gradient element, however small, into a step of ~lr: an element whose gradient is at
        # rounding-noise level may legitimately differ by a fraction of 6 * lr = 4.2e-3.  So: all updates within
        # a quarter of that budget, and 97 % of them within 2 % relative.
        if k == "decoder.bias" and name != "rmse":
            continue      # d loss / d bias = sum_i d loss / d pred_i is identically 0 for the ranking losses: pure rounding noise
        init = g["init." + k]
        du, dw = got - init, v.numpy() - init
        assert np.abs(du - dw).max() < 1e-3, k
        bad = int(np.sum(np.abs(du - dw) > 0.02 * np.abs(dw) + 2e-6))
        assert bad <= max(2, 0.03 * du.size), (k, bad)


def test_backward_h256_vs_autograd_and_sgd():
    """full-size model (F=164, H=256, 300 programs): every gradient tensor against torch autograd, then one
    SGD fine-tune step (mlp_model.py:626,636-638)."""
    torch.manual_seed(21)
    rng = np.random.default_rng(21)
    net = vaex_b200.SegmentSumMLPModule(164, 256, 1)
    P0 = {k: v.detach().clone() for k, v in net.state_dict().items()}
    sizes = rng.integers(1, 33, size=300).astype(np.int32)
    feats = (rng.random((int(sizes.sum()), 164), dtype=np.float32) * (rng.random(164) > 0.15)).astype(np.float32)
    labels = rng.random(300).astype(np.float32)
    tr = train.Trainer(net.cuda(), "lambdaRankLoss")
    tr.forward_backward(_offsets(sizes), torch.from_numpy(feats).cuda(), torch.from_numpy(labels).cuda())
    P = {k: v.clone().requires_grad_(True) for k, v in P0.items()}
    loss = O.loss_lambdarank(O.segsum_mlp_forward(P, sizes, torch.from_numpy(feats)), torch.from_numpy(labels))
    grads = torch.autograd.grad(loss, [P[k] for k in O.PARAM_ORDER])
    np.testing.assert_allclose(tr.loss.item(), float(loss), rtol=1e-4)
    gmax = max(float(gw.abs().max()) for gw in grads)
    for k, gw in zip(O.PARAM_ORDER, grads):
        o, c = tr.slices[k]
        got = tr.grads[o:o + c].cpu().numpy().reshape(gw.shape)
        # (the bias of the scalar decoder has an identically-zero gradient under a ranking loss: noise only)
        assert np.abs(got - gw.numpy()).max() <= 2e-4 * np.abs(gw.numpy()).max() + 2e-6 * gmax, k
    tr.sgd_step(4e-2, 0.0)
    net = tr.export_params()
    for k, gw in zip(O.PARAM_ORDER, grads):
        want = P0[k] - 4e-2 * gw
        np.testing.assert_allclose(net.state_dict()[k].cpu().numpy(), want.numpy(), rtol=1e-4, atol=4e-2 * 2e-4 * gmax + 2e-6, err_msg=k)
    # determinism: the same batch twice gives bit-identical gradients (two-pass reductions, no atomics)
    g1 = tr.grads.clone()
    tr.import_params()
    tr.forward_backward(_offsets(sizes), torch.from_numpy(feats).cuda(), torch.from_numpy(labels).cuda())
    g2 = tr.grads.clone()
    tr.forward_backward(_offsets(sizes), torch.from_numpy(feats).cuda(), torch.from_numpy(labels).cuda())
    assert torch.equal(g2, tr.grads) and not torch.equal(g1, g2)


def test_gather_segments_and_fit_base_learns():
    """the device mini-batch gather against numpy, then MLPModelInternal.fit_base (default LambdaRank loss, Adam,
    clip, shuffled batches of 512, early stopping) on a learnable synthetic task: held-out predictions must
    rank the programs; then the SGD fine-tuning path."""
    rng = np.random.default_rng(5)
    n, n_train = 14000, 12000
    sizes = rng.integers(1, 9, size=n)
    w = rng.standard_normal(164).astype(np.float32)
    feats, score = np.empty(n, dtype=object), np.empty(n, np.float32)
    for i in range(n):
        f = rng.random((sizes[i], 164), dtype=np.float32)
        feats[i] = f
        score[i] = f.mean(0) @ w
    labels = (np.argsort(np.argsort(score)) / float(n)).astype(np.float32)      # "throughput" = percentile of the hidden score
    task = vaex_b200.LearningTask("wk", "cuda")
    ds = vaex_b200.Dataset.create_one_task(task, feats[:n_train], labels[:n_train])
    test = vaex_b200.Dataset.create_one_task(task, feats[n_train:], labels[n_train:])
    m = vaex_b200.MLPModelInternal(device="cuda:0", use_workload_embedding=False)
    # gather kernel
    loader = vaex_b200.SegmentDataLoader(ds, 512, "cuda:0", False, False, {})
    rs = train.ResidentSet(loader, "cuda:0")
    idx = torch.randperm(n_train)[:77]
    off, bf, bl = rs.batch(idx.numpy(), idx.cuda())
    want = np.concatenate([feats[i] for i in idx.numpy()], 0)
    assert np.array_equal(bf.cpu().numpy(), want) and np.array_equal(bl.cpu().numpy(), labels[idx.numpy()])
    assert np.array_equal(off.cpu().numpy(), np.concatenate([[0], np.cumsum(sizes[idx.numpy()])]))
    # training
    m.n_epoch, m.print_per_epoches = 30, 10
    torch.manual_seed(0)
    m.fit_base(ds, valid_set=test)
    from scipy.stats import spearmanr
    rho = spearmanr(m.predict(test)[task], labels[n_train:]).statistic
    assert rho > 0.85, rho
    # fine-tuning path (SGD, mlp_model.py:609-651) runs and keeps the ranking
    m.few_shot_learning = "fine_tune_mix_task"
    m.fine_tune_num_steps, m.fine_tune_lr = 2, 1e-3
    m.fit_local(test)
    rho2 = spearmanr(m.predict(test)[task], labels[n_train:]).statistic
    assert rho2 > 0.85, rho2


def test_mlp_model_update_refits_and_predicts(monkeypatch):
    """MLPModel.update (mlp_model.py:824-830) on the GPU: measured pairs -> TVM's packed feature bytes (stub FFI) ->
    Dataset.update_from_measure_pairs -> fit_base on the CUDA training kernels -> predictions that rank a held-out
    round; the second update re-fits on both rounds."""
    import collections
    import sys
    import types
    from oracle import feature_blob as FB
    Task = collections.namedtuple("Task", ["workload_key", "target"])
    MeasureInput = collections.namedtuple("MeasureInput", ["task", "state"])
    MeasureResult = collections.namedtuple("MeasureResult", ["costs"])

    def get_pairs(inputs, results, skip, max_n_bufs):
        costs = np.array([np.mean(r.costs) for r in results])
        return FB.serialize_features([i.state for i in inputs], costs.min() / costs, np.zeros(len(inputs), np.int32), [costs.min()])
    tvm, autos, ffi = types.ModuleType("tvm"), types.ModuleType("tvm.auto_scheduler"), types.ModuleType("tvm.auto_scheduler._ffi_api")
    ffi.GetPerStoreFeaturesFromMeasurePairs = get_pairs
    for name, mod in (("tvm", tvm), ("tvm.auto_scheduler", autos), ("tvm.auto_scheduler._ffi_api", ffi)):
        monkeypatch.setitem(sys.modules, name, mod)

    rng = np.random.default_rng(9)
    w = rng.standard_normal(164).astype(np.float32)
    task = Task("wk", "cuda")

    def round_of(n):
        states = [rng.random((int(rng.integers(1, 9)), 164), dtype=np.float32) for _ in range(n)]
        score = np.array([s.mean(0) @ w for s in states])
        pct = np.argsort(np.argsort(score)) / float(n)          # a higher hidden score = a faster program:
        cost = 1.0 / (0.05 + pct)                               # normalised throughput = min cost / cost ~ U(0.05, 1)
        return [MeasureInput(task, s) for s in states], [MeasureResult([float(c)]) for c in cost], states, score

    model = vaex_b200.MLPModel()
    model.model.use_workload_embedding = False
    model.model.net_params["in_dim"] = 164
    model.model.n_epoch, model.model.print_per_epoches = 30, 100
    torch.manual_seed(0)
    i1, r1, _, _ = round_of(8000)
    model.update(i1, r1)
    assert len(model.dataset) == 8000 and model.model.base_model is not None
    i2, r2, s2, score2 = round_of(2000)
    feats = np.empty(len(s2), dtype=object)
    for k, s in enumerate(s2):
        feats[k] = s
    feats[5] = np.zeros((2, 164), np.float32)
    pred = model.predict_features(task, feats)
    assert np.isneginf(pred[5])
    from scipy.stats import spearmanr
    keep = np.arange(len(s2)) != 5
    rho1 = spearmanr(pred[keep], score2[keep]).statistic
    assert rho1 > 0.6, rho1
    model.update(i2, r2)
    assert len(model.dataset) == 10000
    pred2 = model.predict_features(task, feats)
    assert spearmanr(pred2[keep], score2[keep]).statistic > max(0.7, rho1 - 0.05)


def test_trainer_epoch_equals_the_step_loop():
    """vaex_trainer_epoch (the whole per-batch loop in one C-ABI call) == the same batches driven one by one through
    gather + Trainer.step, bit for bit: parameters, Adam moments, step counter and the moving-average loss; ragged last
    batch, Adam with clipping and the SGD fine-tuning form."""
    rng = np.random.default_rng(21)
    n = 1500
    feats = np.empty(n, dtype=object)
    for i in range(n):
        feats[i] = rng.random((int(rng.integers(1, 9)), 164), dtype=np.float32)
    labels = rng.random(n).astype(np.float32)
    ds = vaex_b200.Dataset.create_one_task(vaex_b200.LearningTask("wk", "cuda"), feats, labels)
    loader = vaex_b200.SegmentDataLoader(ds, 512, "cuda:0", False, False, {})
    data = train.ResidentSet(loader, "cuda:0")
    for optimizer, lr, wd, clip in (("adam", 7e-4, 1e-6, 0.5), ("sgd", 4e-2, 0.0, 0.0)):
        torch.manual_seed(3)
        net_a = vaex_b200.SegmentSumMLPModule(164, 256, 1).cuda()
        net_b = vaex_b200.SegmentSumMLPModule(164, 256, 1).cuda()
        net_b.load_state_dict(net_a.state_dict())
        ta, tb = train.Trainer(net_a, "lambdaRankLoss"), train.Trainer(net_b, "lambdaRankLoss")
        order = torch.randperm(n, generator=torch.Generator().manual_seed(5))
        avg_a, nb = None, 0
        for _ in range(2):                                         # two epochs: the average carries over
            avg_a, nb = ta.epoch(data, order, 512, lr, wd, clip, optimizer, avg_a)
        assert nb == 3
        avg_b = None
        for _ in range(2):
            for s in range(0, n, 512):
                idx = order[s:s + 512]
                off, bf, bl = data.batch(idx.numpy(), idx.cuda())
                loss = tb.step(off, bf, bl, lr, wd, clip, optimizer)
                avg_b = loss.clone() if avg_b is None else avg_b * 0.95 + loss * 0.05
        assert torch.equal(ta.params, tb.params) and ta.t == tb.t
        if optimizer == "adam":
            assert torch.equal(ta.m, tb.m) and torch.equal(ta.v, tb.v)
        np.testing.assert_allclose(avg_a.item(), avg_b.item(), rtol=1e-6)
