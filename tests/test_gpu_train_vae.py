"""GPU parity of the extent-family VAECostPredictor training row: pair / KL losses, the
forward + total loss + backward call and clip + AdamW (two groups) against goldens produced by the reference's own
pair_loss_fn / kld_loss_fn / compute_total_loss / torch.optim.AdamW (tests/golden/reg_train.npz)."""
import numpy as np
import pytest
import torch

from conftest import load_golden, split_sd
from oracle import torch_oracle as O

import vaex_b200
from vaex_b200 import _lib, training

pytestmark = pytest.mark.gpu


def _cfg(g):
    c = g["cfg"]
    return {"lambda_reg": float(c[0]), "lambda_pair": float(c[1]), "margin": float(c[2]), "gamma": float(c[3]), "beta": float(c[4]),
            "noise_std": 0.5}


def _model(g):
    m = vaex_b200.VAECostPredictor(48, hidden_dim=64, latent_dim=16, predictor_hidden=64, dropout=float(g["cfg"][5]))
    m.load_state_dict(split_sd(g, "init."))
    return m.cuda()


def test_pair_and_kld_losses_vs_reference():
    g = load_golden("reg_train.npz")
    p, t = torch.from_numpy(g["pair_pred"]).cuda(), torch.from_numpy(g["pair_target"]).cuda()
    loss, grad = training.pair_loss_fn(p, t, margin=0.3, return_grad=True)
    np.testing.assert_allclose(loss.item(), float(g["pair_loss"]), rtol=2e-5)
    np.testing.assert_allclose(grad.cpu().numpy(), g["pair_grad"], rtol=1e-4, atol=1e-7)
    mu, lv = torch.from_numpy(g["kld_mu"]).cuda(), torch.from_numpy(g["kld_lv"]).cuda()
    loss, dmu, dlv = training.kld_loss_fn(mu, lv, return_grad=True)
    np.testing.assert_allclose(loss.item(), float(g["kld_loss"]), rtol=2e-5)
    np.testing.assert_allclose(dmu.cpu().numpy(), g["kld_dmu"], rtol=1e-5, atol=1e-9)
    np.testing.assert_allclose(dlv.cpu().numpy(), g["kld_dlv"], rtol=1e-4, atol=1e-8)
    # edge cases: a single prediction has no pairs; two equal targets contribute exactly the margin; 4096 is the limit
    one = torch.ones(1, device="cuda")
    assert training.pair_loss_fn(one, one).item() == 0.0
    two = torch.tensor([0.3, -0.2], device="cuda")
    assert abs(training.pair_loss_fn(two, torch.zeros(2, device="cuda"), margin=0.25).item() - 0.25) < 1e-7
    rng = np.random.default_rng(1)
    pb, tb = rng.standard_normal(4096).astype(np.float32), rng.random(4096).astype(np.float32)
    np.testing.assert_allclose(training.pair_loss_fn(torch.from_numpy(pb).cuda(), torch.from_numpy(tb).cuda(), 0.1).item(),
                               float(O.pair_loss(pb, tb, 0.1)), rtol=1e-4)
    with pytest.raises(_lib.VaexError):
        training.pair_loss_fn(torch.zeros(4097, device="cuda"), torch.zeros(4097, device="cuda"))


def test_forward_backward_vs_reference_gradients():
    g = load_golden("reg_train.npz")
    tr = training.RegressionTrainer(_model(g), 3e-4, 1e-3)
    cost = torch.empty(96, device="cuda")
    tr.forward_backward(torch.from_numpy(g["x"][0]).cuda(), torch.from_numpy(g["y"][0]).cuda(), _cfg(g),
                        drop_mask=torch.from_numpy(g["mask"][0]).cuda(), noise=torch.from_numpy(g["noise"][0]).cuda(), cost_out=cost)
    np.testing.assert_allclose(tr.losses.cpu().numpy(), g["losses"][0], rtol=5e-5)
    np.testing.assert_allclose(cost.cpu().numpy(), g["cost0"], rtol=1e-4, atol=1e-5)
    gmax = max(np.abs(g["grad0." + k]).max() for k in O.REG_PARAM_ORDER)
    for k in O.REG_PARAM_ORDER:
        o, c = tr.slices[k]
        got, want = tr.grads[o:o + c].cpu().numpy().reshape(g["grad0." + k].shape), g["grad0." + k]
        assert np.abs(got - want).max() <= 2e-4 * np.abs(want).max() + 2e-6 * gmax, k
    # deterministic: the same call again gives bit-identical gradients
    g1 = tr.grads.clone()
    tr.forward_backward(torch.from_numpy(g["x"][0]).cuda(), torch.from_numpy(g["y"][0]).cuda(), _cfg(g),
                        drop_mask=torch.from_numpy(g["mask"][0]).cuda(), noise=torch.from_numpy(g["noise"][0]).cuda())
    assert torch.equal(g1, tr.grads)
    # values-only mirror of compute_total_loss (eval-mode model: no dropout -> compare with the oracle)
    m = _model(g).eval()
    x, y = torch.from_numpy(g["x"][1]).cuda(), torch.from_numpy(g["y"][1]).cuda()
    m.set_tensor_cores(False)
    cost_pred, mean, logvar, z = m(x, use_mean=True)
    _, logvar = m.encode(x)
    total, comp = training.compute_total_loss(m, cost_pred, mean, logvar, z, y, None, _cfg(g), noise=torch.from_numpy(g["noise"][1]).cuda())
    wt, wc, _ = O.reg_total_loss(split_sd(g, "init."), g["x"][1], g["y"][1], _cfg(g), None, g["noise"][1])
    np.testing.assert_allclose(total.item(), float(wt), rtol=1e-4)
    for a, b in (("reg_loss", "reg"), ("pair_loss", "pair"), ("smooth_loss", "smooth"), ("kld_loss", "kld")):
        np.testing.assert_allclose(comp[a], float(wc[b]), rtol=2e-4)


@pytest.mark.parametrize("freeze", [False, True])
def test_adamw_trajectory_vs_reference(freeze):
    """three steps of zero_grad / backward / clip_grad_norm_(1.0) / AdamW(two groups, wd 1e-5): losses, pre-clip norms and the
    parameter updates; with the encoder frozen (vae_extent_search.py:1386-1391) against the oracle restatement."""
    g = load_golden("reg_train.npz")
    tr = training.RegressionTrainer(_model(g), 3e-4, 1e-3, encoder_freeze=freeze)
    losses, norms = [], []
    for i in range(3):
        tr.forward_backward(torch.from_numpy(g["x"][i]).cuda(), torch.from_numpy(g["y"][i]).cuda(), _cfg(g),
                            drop_mask=torch.from_numpy(g["mask"][i]).cuda(), noise=torch.from_numpy(g["noise"][i]).cuda())
        tr.optimizer_step()
        losses.append(tr.losses.cpu().numpy().copy())
        norms.append(tr.gnorm.item())
    if freeze:
        batches = [(g["x"][i], g["y"][i], g["mask"][i].astype(np.float32), g["noise"][i]) for i in range(3)]
        P, wl, wn, _ = O.reg_train_steps(split_sd(g, "init."), batches, _cfg(g), 3e-4, 1e-3, p_drop=float(g["cfg"][5]), encoder_freeze=True)
        final = {k: v.numpy() for k, v in P.items()}
    else:
        wl, wn, final = g["losses"], g["norms"], {k: g["final." + k] for k in O.REG_PARAM_ORDER}
    np.testing.assert_allclose(np.array(losses), np.array(wl), rtol=2e-4)
    np.testing.assert_allclose(norms, wn, rtol=1e-3)
    net = tr.export_params()
    for k in O.REG_PARAM_ORDER:
        got, init, want = net.state_dict()[k].cpu().numpy(), g["init." + k], final[k]
        du, dw = got - init, want - init
        if freeze and k in O.REG_ENCODER_KEYS:
            assert np.array_equal(got, init), k
            continue
        # AdamW turns a gradient element of any size into a step of ~lr: elements at rounding-noise level may move by a
        # fraction of 3 * lr; all updates within a quarter of that, 97 % within 2 % relative
        lr = 3e-4 if k in O.REG_ENCODER_KEYS else 1e-3
        assert np.abs(du - dw).max() < 0.75 * lr, k
        bad = int(np.sum(np.abs(du - dw) > 0.02 * np.abs(dw) + 2e-6))
        assert bad <= max(2, 0.03 * du.size), (k, bad)


def test_train_regression_learns():
    """train_regression mirror on a learnable synthetic task: predictions must rank the held-out costs."""
    from scipy.stats import spearmanr
    rng = np.random.default_rng(7)
    torch.manual_seed(7)
    x = rng.standard_normal((1536, 48)).astype(np.float32)
    w = rng.standard_normal(48).astype(np.float32)
    y = np.tanh(x @ w / 6).astype(np.float32)
    model = vaex_b200.VAECostPredictor(48, hidden_dim=256, latent_dim=64).cuda()
    tr = training.RegressionTrainer(model, 1e-3, 1e-3)
    xt, yt = torch.from_numpy(x[:1024]).cuda(), torch.from_numpy(y[:1024]).cuda()
    loader = [(xt[i:i + 256], yt[i:i + 256].unsqueeze(-1)) for i in range(0, 1024, 256)]
    cfg = {"lambda_reg": 1.0, "lambda_pair": 1.0, "margin": 0.05, "gamma": 0.01, "beta": 0.001, "noise_std": 0.01, "epochs": 60}
    model = training.train_regression(model, tr, loader, cfg, warmup_epochs=10).eval()
    pred = model(torch.from_numpy(x[1024:]).cuda(), use_mean=True)[0].cpu().numpy()
    rho = spearmanr(pred, y[1024:]).statistic
    assert rho > 0.9, rho


def test_vae_pretraining_vs_reference():
    """vae_train's step (forward(use_mean=False) + vae_feature_loss + Adam) against the reference's VAE_feature_head,
    vae_feature_loss and torch.optim.Adam (tests/golden/vae_train.npz): first-step gradients, three losses, parameters."""
    g = load_golden("vae_train.npz")
    vae = vaex_b200.VAE_feature_head(48, None, 16, 64)
    vae.load_state_dict(split_sd(g, "init."))
    tr = training.VAETrainer(vae.cuda(), lr=1e-3, alpha_recon=1.0, beta=0.01)
    tr.forward_backward(torch.from_numpy(g["x"][0]).cuda(), torch.from_numpy(g["eps"][0]).cuda())
    np.testing.assert_allclose(tr.losses.cpu().numpy(), g["losses"][0], rtol=5e-5)
    gmax = max(np.abs(g["grad0." + k]).max() for k in O.VAE_PARAM_ORDER)
    for k in O.VAE_PARAM_ORDER:
        o, c = tr.slices[k]
        got, want = tr.grads[o:o + c].cpu().numpy().reshape(g["grad0." + k].shape), g["grad0." + k]
        assert np.abs(got - want).max() <= 2e-4 * np.abs(want).max() + 2e-6 * gmax, k
    losses = []
    for i in range(3):
        losses.append(tr.step(torch.from_numpy(g["x"][i]).cuda(), torch.from_numpy(g["eps"][i]).cuda()).cpu().numpy().copy())
    np.testing.assert_allclose(np.array(losses), g["losses"], rtol=2e-4)
    net = tr.export_params()
    for k in O.VAE_PARAM_ORDER:
        du, dw = net.state_dict()[k].cpu().numpy() - g["init." + k], g["final." + k] - g["init." + k]
        assert np.abs(du - dw).max() < 0.75e-3, k
        bad = int(np.sum(np.abs(du - dw) > 0.02 * np.abs(dw) + 2e-6))
        assert bad <= max(2, 0.03 * du.size), (k, bad)
    # use_mean path (eps = None) against the oracle, and the values-only loss mirror
    tr.import_params()
    tr.forward_backward(torch.from_numpy(g["x"][1]).cuda(), None)
    P = {k: v.detach().cpu() for k, v in net.state_dict().items()}
    wt, wr, wk = O.vae_loss(P, g["x"][1], None)
    np.testing.assert_allclose(tr.losses.cpu().numpy(), [float(wt), float(wr), float(wk)], rtol=1e-4)
    net = net.eval()
    net.set_tensor_cores(False)
    x = torch.from_numpy(g["x"][1]).cuda()
    xr, mu, lv, z, fp = net(x, use_mean=True)
    _, lv = net.encode(x)
    loss, recon, kl, _ = training.vae_feature_loss(xr, x, mu, lv, None, None, alpha_recon=1.0, beta=0.01)
    np.testing.assert_allclose([loss.item(), recon.item(), kl.item()], [float(wt), float(wr), float(wk)], rtol=1e-4)
