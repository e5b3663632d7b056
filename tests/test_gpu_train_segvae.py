"""SegmentVAE training on the GPU (segment family) against three steps of the reference's own
SegmentVAE + VAE_Trainer.vae_loss + clip_grad_norm_ + torch.optim.AdamW (tests/golden/segvae_train.npz; the noise and
the dropout mask of every step were captured from the reference run and are replayed here)."""
import numpy as np
import pytest
import torch

from conftest import load_golden, split_sd
from oracle import torch_oracle as O

import vaex_b200
from vaex_b200 import seg_training

pytestmark = pytest.mark.gpu


def _setup():
    g = load_golden("segvae_train.npz")
    beta, lr, wd, clip, p = [float(v) for v in g["hyper"]]
    model = vaex_b200.SegmentVAE(164, 256, 64, dropout=p)
    model.load_state_dict(split_sd(g, "w."))
    tr = seg_training.SegmentVAETrainer(model.cuda(), lr=lr, beta=beta, weight_decay=wd, max_norm=clip)
    return g, model, tr, torch.from_numpy(g["sizes"]).cuda(), torch.from_numpy(g["feats"]).cuda()


def test_segvae_first_step_vs_reference():
    g, model, tr, sizes, feats = _setup()
    losses, mean, recon = tr.forward_backward(sizes, feats, torch.from_numpy(g["eps"][0]).cuda(), torch.from_numpy(g["mask"][0]).cuda(),
                                              want_outputs=True)
    np.testing.assert_allclose(losses.cpu().numpy(), g["losses"][0], rtol=1e-4)
    assert np.abs(mean.cpu().numpy() - g["mean0"]).max() < 2e-4 * np.abs(g["mean0"]).max()
    assert np.abs(recon.cpu().numpy() - g["recon0"]).max() < 2e-4 * np.abs(g["recon0"]).max()
    gmax = max(np.abs(g["g." + k]).max() for k in O.SEGVAE_PARAM_ORDER)
    for k in O.SEGVAE_PARAM_ORDER:
        o, c = tr.slices[k]
        got, want = tr.grads[o:o + c].cpu().numpy().reshape(g["g." + k].shape), g["g." + k]
        assert np.abs(got - want).max() <= 3e-4 * np.abs(want).max() + 2e-6 * gmax, (k, np.abs(got - want).max(), np.abs(want).max())
    # same batch again: bit-identical gradients (fixed-order reductions); running statistics advanced twice
    first = tr.grads.clone()
    tr.forward_backward(sizes, feats, torch.from_numpy(g["eps"][0]).cuda(), torch.from_numpy(g["mask"][0]).cuda())
    assert torch.equal(first, tr.grads)


def test_segvae_three_steps_vs_reference():
    g, model, tr, sizes, feats = _setup()
    losses, norms = [], []
    for i in range(3):
        losses.append(tr.step(sizes, feats, torch.from_numpy(g["eps"][i]).cuda(), torch.from_numpy(g["mask"][i]).cuda()).cpu().numpy().copy())
        norms.append(float(tr.gnorm.item()))
    np.testing.assert_allclose(np.array(losses), g["losses"], rtol=5e-4)
    np.testing.assert_allclose(norms, g["gnorm"], rtol=2e-3)
    net = tr.export_params()
    sd = net.state_dict()
    for k in O.SEGVAE_PARAM_ORDER:
        du, dw = sd[k].cpu().numpy() - g["w." + k], g["p3." + k] - g["w." + k]
        assert np.abs(du - dw).max() < 0.75e-3, k                      # AdamW: any element may move by ~lr per step
        if k in ("decoder.0.bias", "decoder.4.bias"):
            continue   # a bias in front of a batch-statistics BatchNorm has an exactly zero gradient: what AdamW
                       # normalises there is rounding noise, in the reference as much as here
        bad = int(np.sum(np.abs(du - dw) > 0.02 * np.abs(dw) + 2e-6))
        assert bad <= max(2, 0.03 * du.size), (k, bad)
    for b in O.SEGVAE_BN:
        np.testing.assert_allclose(sd[b + ".running_mean"].cpu().numpy(), g["p3." + b + ".running_mean"], rtol=2e-4, atol=2e-5)
        np.testing.assert_allclose(sd[b + ".running_var"].cpu().numpy(), g["p3." + b + ".running_var"], rtol=2e-4, atol=2e-5)
        assert int(sd[b + ".num_batches_tracked"]) == int(g["p3." + b + ".num_batches_tracked"]) == 3
    # the exported module scores like the reference's trained module in eval mode (inference kernels, BN folded)
    net.eval()
    P = {k[3:]: torch.from_numpy(np.ascontiguousarray(v)) for k, v in g.items() if k.startswith("p3.")}
    wm, wl, wS = O.segvae_encode(P, g["sizes"], g["feats"])
    mean, logvar, S = net.encode_segments(sizes, feats)
    assert np.abs(mean.cpu().numpy() - wm.numpy()).max() < 5e-3 * np.abs(wm.numpy()).max() + 5e-3


def test_segvae_use_mean_and_philox_dropout():
    g, model, tr, sizes, feats = _setup()
    # eps = None: z = mean, logvar = 0 -> fc_logvar receives no gradient, kld = 0.5 * sum(mean^2)
    losses, mean, recon = tr.forward_backward(sizes, feats, None, None, want_outputs=True)
    o, c = tr.slices["fc_logvar.weight"]
    assert not tr.grads[o:o + c].any()
    np.testing.assert_allclose(float(losses[2]), 0.5 * float((mean ** 2).sum()), rtol=1e-4)
    # Philox dropout: same seed -> same gradients, different seed -> different; the kept fraction is ~ 1 - p
    tr.forward_backward(sizes, feats, torch.from_numpy(g["eps"][0]).cuda(), None, drop_seed=11)
    a = tr.grads.clone()
    tr.forward_backward(sizes, feats, torch.from_numpy(g["eps"][0]).cuda(), None, drop_seed=11)
    assert torch.equal(a, tr.grads)
    tr.forward_backward(sizes, feats, torch.from_numpy(g["eps"][0]).cuda(), None, drop_seed=12)
    assert not torch.equal(a, tr.grads)
    with pytest.raises(vaex_b200._lib.VaexError):
        tr.forward_backward(sizes[:1], feats[: int(g["sizes"][0])], None, None)     # BatchNorm needs 2 programs


def test_vae_trainer_learns():
    """VAE_Trainer.train_vae mirror end to end: the epoch loop, schedule, validation and early-stopping bookkeeping run,
    and the training objective falls.  (The reference's objective does not detach its target -- the row encoder is free
    to shrink the segment sums it is asked to reconstruct -- so held-out R^2 is not a property to assert on synthetic rows.)"""
    rng = np.random.default_rng(3)
    torch.manual_seed(3)

    def batch(n):
        sizes = rng.integers(1, 9, size=n).astype(np.int32)
        base = rng.random((n, 164), dtype=np.float32)
        feats = np.concatenate([np.repeat(base[i:i + 1], s, axis=0) + 0.05 * rng.random((s, 164), dtype=np.float32)
                                for i, s in enumerate(sizes)]).astype(np.float32)
        return torch.from_numpy(sizes), torch.from_numpy(feats), torch.zeros(n)

    train, val = [batch(128) for _ in range(6)], [batch(128)]
    vt = seg_training.VAE_Trainer(train, val, input_dim=164, hidden_dim=256, latent_dim=64, dropout=0.1, device="cuda")
    model, r2, kl, hist = vt.train_vae(epochs=40, lr=2e-3, beta=1e-4, patience=40)
    assert len(hist["recon_r2"]) == len(hist["recon_loss"]) == 40 and np.isfinite(hist["recon_loss"]).all() and np.isfinite(r2)
    assert hist["recon_loss"][-1] < 0.2 * hist["recon_loss"][0], (hist["recon_loss"][0], hist["recon_loss"][-1])
    assert r2 == max(hist["recon_r2"]) and not model.training and all(int(model.state_dict()[b + ".num_batches_tracked"]) > 0
                                                                     for b in O.SEGVAE_BN)
