"""Segment-family VAECostPredictor training on the GPU against three steps of the reference's own
VAECostPredictor + Regression_Trainer.compute_total_loss + clip_grad_norm_ + two-group AdamW
(tests/golden/segreg_train.npz; the smoothness noise and the six dropout masks of every step are replayed)."""
import numpy as np
import pytest
import torch

from conftest import load_golden, split_sd
from oracle import torch_oracle as O

import vaex_b200
from vaex_b200 import seg_training
from vaex_b200.seg_models import VAECostPredictor

pytestmark = pytest.mark.gpu


def _setup():
    g = load_golden("segreg_train.npz")
    elr, plr, lp, gamma, beta, margin, nstd, p = [float(v) for v in g["hyper"]]
    cfg = {"lambda_pair": lp, "gamma": gamma, "beta": beta, "margin": margin, "noise_std": nstd}
    model = VAECostPredictor(164, 256, 64, 256, 3, p)
    model.load_state_dict(split_sd(g, "w."))
    tr = seg_training.SegmentRegressionTrainer(model.cuda(), elr, plr)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()   # noqa: E731
    return g, cfg, model, tr, dev(g["sizes"]), dev(g["feats"]), dev(g["labels"]), dev


def test_segreg_first_step_vs_reference():
    g, cfg, model, tr, sizes, feats, labels, dev = _setup()
    cost = torch.empty(sizes.shape[0], device="cuda")
    losses = tr.forward_backward(sizes, feats, labels, cfg, dev(g["noise"][0]), dev(g["mask"][0]), cost_out=cost)
    np.testing.assert_allclose(losses.cpu().numpy(), g["losses"][0], rtol=2e-4)
    enc, pred = O.segreg_param_order()
    gmax = max(np.abs(g["g." + k]).max() for k in enc + pred)
    for k in enc + pred:
        o, c = tr.slices[k]
        got, want = tr.grads[o:o + c].cpu().numpy().reshape(g["g." + k].shape), g["g." + k]
        assert np.abs(got - want).max() <= 3e-4 * np.abs(want).max() + 2e-6 * gmax, (k, np.abs(got - want).max(), np.abs(want).max())
    first = tr.grads.clone()
    tr.forward_backward(sizes, feats, labels, cfg, dev(g["noise"][0]), dev(g["mask"][0]))
    assert torch.equal(first, tr.grads)                             # fixed-order reductions


def test_segreg_three_steps_vs_reference():
    g, cfg, model, tr, sizes, feats, labels, dev = _setup()
    losses, norms = [], []
    for i in range(3):
        losses.append(tr.step(sizes, feats, labels, cfg, dev(g["noise"][i]), dev(g["mask"][i])).cpu().numpy().copy())
        norms.append(float(tr.gnorm.item()))
    np.testing.assert_allclose(np.array(losses), g["losses"], rtol=2e-3)
    np.testing.assert_allclose(norms, g["gnorm"], rtol=5e-3)
    net = tr.export_params()
    sd = net.state_dict()
    enc, pred = O.segreg_param_order()
    for k in enc + pred:
        lr = 3e-4 if k in enc else 1e-3
        du, dw = sd[k].cpu().numpy() - g["w." + k], g["p3." + k] - g["w." + k]
        assert np.abs(du - dw).max() < 0.75 * lr, k
        if k.endswith(".bias") and k.split(".")[1] in ("0", "4", "8"):
            continue       # Linear bias in front of a batch-statistics BatchNorm: exactly zero gradient, AdamW moves it on noise
        bad = int(np.sum(np.abs(du - dw) > 0.02 * np.abs(dw) + 2e-6))
        assert bad <= max(2, 0.03 * du.size), (k, bad)
    for j, b in enumerate(tr.bn_keys):
        np.testing.assert_allclose(sd[b + ".running_mean"].cpu().numpy(), g["p3." + b + ".running_mean"], rtol=1e-3, atol=3e-4)
        np.testing.assert_allclose(sd[b + ".running_var"].cpu().numpy(), g["p3." + b + ".running_var"], rtol=1e-3, atol=3e-4)
        assert int(sd[b + ".num_batches_tracked"]) == int(g["p3." + b + ".num_batches_tracked"]) == (3 if j == 0 else 9)


def test_regression_trainer_learns():
    """Regression_Trainer mirror end to end on a learnable synthetic task: the total loss falls and predictions rank the costs."""
    from scipy.stats import spearmanr
    rng = np.random.default_rng(5)
    torch.manual_seed(5)
    n = 256
    sizes = rng.integers(1, 9, size=n).astype(np.int32)
    base = rng.random((n, 164), dtype=np.float32)
    feature_list = [np.repeat(base[i:i + 1], s, axis=0) + 0.02 * rng.random((s, 164), dtype=np.float32) for i, s in enumerate(sizes)]
    w = rng.standard_normal(164).astype(np.float32)
    labels = np.tanh(base @ w / 4).astype(np.float32)
    vae = vaex_b200.SegmentVAE(164, 256, 64)
    rt = seg_training.Regression_Trainer({"input_dim": 164, "hidden_dim": 256, "latent_dim": 64, "model": vae}, device="cuda")
    cfg = {"encoder_lr": 1e-3, "predictor_lr": 2e-3, "lambda_pair": 0.3, "gamma": 0.01, "beta": 1e-4, "margin": 0.05,
           "noise_std": 0.01, "mini_epochs": 150}
    history, best = rt.train_regression(feature_list, sizes, labels, None, cfg, phases=1)
    assert history[-1]["total_loss"] < 0.3 * history[0]["total_loss"], (history[0], history[-1])
    model = rt.model.eval()
    flat = torch.from_numpy(np.concatenate(feature_list)).cuda()
    pred = model(torch.from_numpy(sizes).cuda(), flat, use_mean=True)[0].cpu().numpy()
    assert spearmanr(pred, labels).statistic > 0.8
