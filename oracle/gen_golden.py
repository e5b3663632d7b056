"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz from the UNMODIFIED reference.

Run in the build container (needs /root/reference):  python oracle/gen_golden.py
Every array written here is an input to, a weight of, or an output of the reference's own
modules (imported by oracle/ref_import.py).  Fixtures are a few MB in total (dominated by the fp32 weights) and committed.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.ref_import import load_reference  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def sd_np(module, prefix="w."):
    return {prefix + k: v.detach().cpu().numpy() for k, v in module.state_dict().items()}


def synth_programs(rng, n, F=164, smax=32, zero_frac=0.02):
    progs = []
    for _ in range(n):
        s = int(rng.integers(1, smax + 1))
        p = (rng.random((s, F), dtype=np.float32) * rng.integers(0, 25, size=(1, F))).astype(np.float32)
        p[:, rng.random(F) < 0.15] = 0.0
        if rng.random() < zero_frac:
            p[:] = 0.0
        progs.append(p)
    return progs


def gen_segsum_mlp(ns):
    """a1-a3: SegmentDataLoader + SegmentSumMLPModule + MLPModelInternal.predict."""
    rng = np.random.default_rng(1264)
    for tag, F, emb in (("f164", 164, False), ("f174", 174, True)):
        torch.manual_seed(7)
        m = ns.mlp_model.MLPModelInternal(device="cpu", use_workload_embedding=emb)
        m.base_model = ns.mlp_model.make_net(dict(m.net_params))
        m.base_model.eval()
        progs = synth_programs(rng, 97)
        progs[3][:] = 0.0  # invalid program -> -inf at the MLPModel.predict level
        feats = np.empty(len(progs), dtype=object)
        for i, p in enumerate(progs):
            feats[i] = p
        task = ns.dataset.LearningTask("wk", "cuda")
        ds = ns.dataset.Dataset.create_one_task(task, feats, np.zeros(len(progs)))
        loader = ns.mlp_model.SegmentDataLoader(ds, 4096, "cpu", emb, False, {}, None)
        norm = loader.normalize()
        m.fea_norm_vec = norm
        flat_norm = loader.features.numpy().copy()
        sizes = loader.segment_sizes.numpy().copy()
        with torch.no_grad():
            direct = m.base_model(loader.segment_sizes, loader.features).numpy()
            m.infer_batch_size = 40  # exercise batching (ragged last batch)
            pred = m.predict(ds)[task]
        # MLPModel.predict's invalid-state rule (mlp_model.py:842-844), applied by hand
        masked = pred.copy()
        for i, p in enumerate(progs):
            if p.min() == p.max() == 0:
                masked[i] = -np.inf
        np.savez_compressed(
            os.path.join(OUT, "segsum_mlp_%s.npz" % tag),
            raw=np.concatenate(progs, 0), sizes=sizes, norm=norm.numpy(), flat_norm_head=flat_norm[:64],
            flat_norm_sum=flat_norm.astype(np.float64).sum(axis=0),
            direct=direct, pred=pred, masked=masked, **sd_np(m.base_model))


def gen_extent_vae(ns):
    """a5, a6, a12, a13: VAE_feature_head / VAECostPredictor (extent family)."""
    D, H, L = 48, 256, 64
    torch.manual_seed(2023)
    vae = ns.vae_model.VAE_feature_head(D, None, L, H).eval()
    reg = ns.vae_model.VAECostPredictor(D, hidden_dim=H, latent_dim=L).eval()
    reg.load_pretrained_encoder(vae.state_dict())
    g = torch.Generator().manual_seed(2023)
    x = torch.randn(300, D, generator=g)
    z = torch.randn(300, L, generator=g)
    with torch.no_grad():
        mu, logvar = vae.encode(x)
        dec = vae.decode(z)
        recon = vae(x, use_mean=True)[0]
        cost = reg.predict_cost(z)
        cost_x, mean_x, logvar_x, z_x = reg(x, use_mean=True)
    zz = z.clone().requires_grad_(True)
    c = reg.predict_cost(zz)
    grad = torch.autograd.grad(c.sum(), zz)[0]
    # mc_predict (model.py:241-273) with explicit Bernoulli(0.9) masks in place of nn.Dropout's
    # private Philox stream; same graph, same 1/(1-p) scaling, mean + unbiased var over T.
    T = 5
    gm = torch.Generator().manual_seed(99)
    mk = (torch.rand(T, 300, H, generator=gm) >= 0.1).float()
    with torch.no_grad():
        preds = []
        for t in range(T):
            h = reg.cost_predictor[1](reg.cost_predictor[0](mean_x))
            h = h * mk[t] / 0.9
            h = reg.cost_predictor[4](reg.cost_predictor[3](h))
            preds.append(reg.cost_predictor[6](h).squeeze(-1))
        preds = torch.stack(preds, 0)
        mc_mean, mc_var = preds.mean(0), preds.var(0)
    np.savez_compressed(
        os.path.join(OUT, "extent_vae.npz"),
        x=x.numpy(), z=z.numpy(), mu=mu.numpy(), logvar=logvar.numpy(), dec=dec.numpy(),
        recon=recon.numpy(), cost=cost.numpy(), cost_x=cost_x.numpy(), mean_x=mean_x.numpy(),
        grad=grad.numpy(), masks=mk.numpy().astype(np.uint8), mc_mean=mc_mean.numpy(),
        mc_var=mc_var.numpy(), **sd_np(vae, "vae."), **sd_np(reg, "reg."))


def gen_snap(ns):
    """a9: postprocess (vae_sch_extent.ipynb:1374-1383) + clamp(min=1) (:1457), restated
    literally with the same torch calls (the notebook cell is not importable)."""
    g = torch.Generator().manual_seed(5)
    x = torch.randn(200, 48, generator=g) * 1.5
    mean = torch.rand(48, generator=g) * 2.5 + 0.5
    std = torch.rand(48, generator=g) * 1.2 + 0.3
    x[0, :8] = torch.tensor([-50.0, 50.0, 0.0, 1e-3, -1e-3, 3.0, -3.0, 0.5])
    v = x * (std + 1e-8) + mean
    ext = torch.clamp(torch.round(torch.exp(v)).int(), min=1)
    v2 = x * std + mean
    ext_log1p = torch.clamp(torch.round(torch.expm1(v2)).int(), min=1)
    np.savez_compressed(os.path.join(OUT, "snap.npz"), x=x.numpy(), mean=mean.numpy(),
                        std=std.numpy(), ext=ext.numpy(), ext_log1p=ext_log1p.numpy())


def gen_select(ns):
    """a11/a12/a14: util_select_topk, select_topk_z_grad and select_programs on a seeded pool."""
    D, H, L = 48, 256, 64
    torch.manual_seed(11)
    reg = ns.vae_model.VAECostPredictor(D, hidden_dim=H, latent_dim=L).eval()
    g = torch.Generator().manual_seed(12)
    x = torch.randn(1024, D, generator=g)
    with torch.no_grad():
        z, _ = reg.encode(x)
        cost = reg.predict_cost(z).numpy()
    remaining = set(range(1024)) - set(range(0, 1024, 7))
    sel, rem_after = ns.select.util_select_topk(cost, set(remaining), 48)
    used = set(range(0, 1024, 7))
    sel_prog, rem_prog = ns.select.select_programs(
        reg, x.numpy(), set(used), set(remaining), num_select=64, T_mc=4, uncertainty_topk=10 ** 9,
        w_cost=1.0, w_unc=0.0, w_div=0.0, grad_num=2, rand_num=0,
        rng=np.random.default_rng(42), device=torch.device("cpu"), topk_factor=5)
    np.savez_compressed(
        os.path.join(OUT, "select.npz"), x=x.numpy(), cost=cost, z=z.numpy(),
        remaining=np.array(sorted(remaining), np.int64), topk48=np.sort(sel),
        used=np.array(sorted(used), np.int64), sel_prog=sel_prog,
        rem_prog=np.array(sorted(rem_prog), np.int64), **sd_np(reg, "reg."))


def gen_segment_family(ns):
    """a7/a8 with the reference's *shipped trained* checkpoints (real BatchNorm running stats):
    SegmentVAE = model_final/checkpoints/vae_small/vae_rank10_h256_l32_*.pt, segment
    VAECostPredictor = for_inference/vae_latent_augmented_best.pth (carries its fea_norm_vec)."""
    ref = os.environ.get("VAEX_REFERENCE_ROOT", "/root/reference")
    rng = np.random.default_rng(2024)
    progs = synth_programs(rng, 67, zero_frac=0.0)
    sizes = np.array([len(p) for p in progs], np.int32)
    raw = np.concatenate(progs, 0)
    # --- SegmentVAE
    ck = torch.load(os.path.join(ref, "scripts/pre_experiments/model_final/checkpoints/vae_small/"
                                      "vae_rank10_h256_l32_b5e-04_lr5e-04.pt"), map_location="cpu", weights_only=False)
    cfg = ck["config"]
    vae = ns.seg_vae.SegmentVAE(164, hidden_dim=cfg["hidden_dim"], latent_dim=cfg["latent_dim"])
    vae.load_state_dict(ck["model_state_dict"])
    vae.eval()
    feats = torch.from_numpy(raw / 25.0)
    with torch.no_grad():
        h = vae.segment_encoder(feats)
        S = vae.segment_sum(torch.from_numpy(sizes), h)
        mean, logvar = vae.encode(S)
        recon = vae.decode(mean)
        torch.manual_seed(5)
        fm, fl, fz, frecon, fS = vae(torch.from_numpy(sizes), feats, use_mean=True)   # z = mean + eps (logvar forced to 0)
    np.savez_compressed(os.path.join(OUT, "segment_vae.npz"), feats=feats.numpy(), sizes=sizes, segsum=S.numpy(),
                        mean=mean.numpy(), logvar=logvar.numpy(), recon=recon.numpy(),
                        fwd_eps=(fz - fm).numpy(), fwd_logvar=fl.numpy(), fwd_recon=frecon.numpy(),
                        hidden=cfg["hidden_dim"], latent=cfg["latent_dim"], **sd_np(vae))
    # --- segment VAECostPredictor
    ck = torch.load(os.path.join(ref, "scripts/pre_experiments/for_inference/vae_latent_augmented_best.pth"),
                    map_location="cpu", weights_only=False)
    cfg = ck["config"]
    reg = ns.seg_reg.VAECostPredictor(cfg["input_dim"], cfg["hidden_dim"], cfg["latent_dim"], cfg["predictor_hidden"],
                                      cfg["predictor_layers"], cfg["dropout"])
    reg.load_state_dict(ck["model_state_dict"])
    reg.eval()
    norm = np.asarray(ck["fea_norm_vec"], np.float32).reshape(-1)
    norm = np.where(norm == 0, 1.0, norm).astype(np.float32)
    featn = torch.from_numpy(raw / norm)
    with torch.no_grad():
        cost, mean, logvar, z = reg(torch.from_numpy(sizes), featn, use_mean=True)
        zt = torch.randn(33, cfg["latent_dim"], generator=torch.Generator().manual_seed(6))
        cost_z = reg.predict_cost(zt)
    np.savez_compressed(os.path.join(OUT, "segment_cost.npz"), raw=raw, norm=norm, sizes=sizes, cost=cost.numpy(),
                        mean=mean.numpy(), logvar=logvar.numpy(), z_test=zt.numpy(), cost_z=cost_z.numpy(),
                        dims=np.array([cfg["input_dim"], cfg["hidden_dim"], cfg["latent_dim"], cfg["predictor_hidden"],
                                       cfg["predictor_layers"]]), **sd_np(reg))


def gen_train(ns):
    """a4: the reference's loss classes (value + autograd gradient) and short optimisation
    trajectories with its own make_net / Adam / clip_grad_norm_ (mlp_model.py:541-565)."""
    M = ns.mlp_model
    rng = np.random.default_rng(77)
    out = {}
    ref_losses = {"rmse": torch.nn.MSELoss(), "rankNetLoss": M.RankNetLoss(), "lambdaRankLoss": M.LambdaRankLoss(),
                  "listNetLoss": M.ListNetLoss()}
    for n in (64, 257):
        preds = torch.from_numpy(rng.standard_normal(n).astype(np.float32))
        labels = torch.from_numpy(rng.random(n).astype(np.float32))
        labels[::7] = 0.0                                  # failed measurements: throughput 0, ties in the labels
        out["loss_preds_%d" % n], out["loss_labels_%d" % n] = preds.numpy(), labels.numpy()
        for name, fn in ref_losses.items():
            p = (torch.sigmoid(preds) if name == "rmse" else preds).clone().requires_grad_(True)
            kw = {"device": "cpu"} if name == "lambdaRankLoss" else {}
            loss = fn(p, labels, **kw)
            (g,) = torch.autograd.grad(loss, p)
            out["loss_%s_%d" % (name, n)], out["grad_%s_%d" % (name, n)] = np.float32(loss.item()), g.numpy()
    # optimisation trajectories (hidden_dim 64 keeps the fixture small)
    progs = synth_programs(rng, 3 * 48, smax=8, zero_frac=0.0)
    sizes = np.array([len(p) for p in progs], np.int32)
    feats = np.concatenate(progs, 0) / 25.0
    labels = rng.random(len(progs)).astype(np.float32)
    offs = np.concatenate([[0], np.cumsum(sizes)])
    out.update(train_feats=feats.astype(np.float32), train_sizes=sizes, train_labels=labels)
    for name in ("lambdaRankLoss", "rmse", "rankNetLoss", "listNetLoss"):
        torch.manual_seed(11)
        net = M.make_net({"type": "SegmentSumMLP", "in_dim": 164, "hidden_dim": 64, "out_dim": 1, "add_sigmoid": name == "rmse"})
        net.train()
        for k, v in net.state_dict().items():
            out["init.%s" % k] = v.detach().numpy().copy()
        opt = torch.optim.Adam(net.parameters(), lr=7e-4, weight_decay=1e-6)
        losses, norms = [], []
        for step in range(6):
            b = step % 3
            s, e = 48 * b, 48 * (b + 1)
            opt.zero_grad()
            kw = {"device": "cpu"} if name == "lambdaRankLoss" else {}
            loss = ref_losses[name](net(torch.from_numpy(sizes[s:e]), torch.from_numpy(feats[offs[s]:offs[e]].astype(np.float32))),
                                    torch.from_numpy(labels[s:e]), **kw)
            loss.backward()
            norms.append(float(torch.nn.utils.clip_grad_norm_(net.parameters(), 0.5)))
            opt.step()
            losses.append(float(loss))
        out["traj_losses_%s" % name], out["traj_norms_%s" % name] = np.array(losses, np.float64), np.array(norms, np.float64)
        for k, v in net.state_dict().items():
            out["final_%s.%s" % (name, k)] = v.detach().numpy().copy()
    np.savez_compressed(os.path.join(OUT, "train.npz"), **out)


def gen_reg_train(ns):
    """a15: the reference's pair / kld / total loss (utils/training.py) and three train_regression steps
    (vae_extent_search.py:1389-1440: AdamW two groups, clip 1.0) on a small seeded VAECostPredictor.
    nn.Dropout is replaced on the instance by a fixed-mask twin (same h * mask / (1 - p) in train mode) and the
    smooth-loss noise is reproduced by seeding right before compute_total_loss, so both are recorded."""
    from oracle.ref_import import load_reference_training
    T = load_reference_training()
    D, H, L, PH, n, p = 48, 64, 16, 64, 96, 0.1
    torch.manual_seed(31)
    reg = ns.vae_model.VAECostPredictor(D, hidden_dim=H, latent_dim=L, predictor_hidden=PH, predictor_layers=2, dropout=p)

    class FixedDrop(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.mask = None

        def forward(self, h):
            return h * self.mask / (1.0 - p) if self.training else h

    drop = FixedDrop()
    reg.cost_predictor[2] = drop
    out = {k: v.copy() for k, v in sd_np(reg, "init.").items()}      # copies: the optimizer updates the tensors in place
    g = torch.Generator().manual_seed(32)
    cfg = {"lambda_reg": 0.5, "lambda_pair": 6.0, "margin": 0.3, "gamma": 2.0, "beta": 0.01, "noise_std": 0.5, "loss_type": "mse"}
    out["cfg"] = np.array([cfg["lambda_reg"], cfg["lambda_pair"], cfg["margin"], cfg["gamma"], cfg["beta"], p], np.float32)
    # stand-alone losses
    pred, target = torch.randn(257, generator=g), torch.rand(257, generator=g)
    target[5] = target[9]                                        # an equal-target pair (label 0 -> contributes the margin)
    pr = pred.clone().requires_grad_(True)
    lp = T.pair_loss_fn(pr, target, margin=0.3)
    lp.backward()
    mu, lv = torch.randn(40, 16, generator=g).requires_grad_(True), (0.3 * torch.randn(40, 16, generator=g)).requires_grad_(True)
    lk = T.kld_loss_fn(mu, lv)
    lk.backward()
    out.update({"pair_pred": pred.numpy(), "pair_target": target.numpy(), "pair_loss": lp.detach().numpy(), "pair_grad": pr.grad.numpy(),
                "kld_mu": mu.detach().numpy(), "kld_lv": lv.detach().numpy(), "kld_loss": lk.detach().numpy(),
                "kld_dmu": mu.grad.numpy(), "kld_dlv": lv.grad.numpy()})
    # three training steps
    opt = torch.optim.AdamW([{"params": reg.get_encoder_params(), "lr": 3e-4},
                             {"params": reg.get_cost_predictor_params(), "lr": 1e-3}], weight_decay=1e-5)
    xs, ys, masks, noises, losses, norms = [], [], [], [], [], []
    for step in range(3):
        x, y = torch.randn(n, D, generator=g), 4.0 * torch.rand(n, generator=g)
        drop.mask = (torch.rand(n, PH, generator=g) > p).float()
        reg.train()
        cost_pred, mean, logvar, z = reg(x, use_mean=True)
        torch.manual_seed(100 + step)
        noise = cfg["noise_std"] * torch.randn(n, L)
        torch.manual_seed(100 + step)
        total, comp = T.compute_total_loss(reg, cost_pred, mean, logvar, z, y, None, cfg)
        opt.zero_grad()
        total.backward()
        if step == 0:
            out.update({"grad0." + k: v.grad.detach().numpy().copy() for k, v in reg.named_parameters()})
            out["cost0"] = cost_pred.detach().numpy()
        norms.append(float(torch.nn.utils.clip_grad_norm_(reg.parameters(), max_norm=1.0)))
        opt.step()
        losses.append([float(total), comp["reg_loss"], comp["pair_loss"], comp["smooth_loss"], comp["kld_loss"]])
        xs.append(x.numpy()); ys.append(y.numpy()); masks.append(drop.mask.numpy().astype(np.uint8)); noises.append(noise.numpy())
    out.update({"x": np.stack(xs), "y": np.stack(ys), "mask": np.stack(masks), "noise": np.stack(noises),
                "losses": np.array(losses, np.float64), "norms": np.array(norms, np.float64)})
    out.update({"final." + k: v.detach().numpy() for k, v in reg.state_dict().items()})
    np.savez_compressed(os.path.join(OUT, "reg_train.npz"), **out)


def gen_vae_train(ns):
    """a15: three vae_train steps (utils/training.py:65-88: forward(use_mean=False), vae_feature_loss, Adam) of the
    reference's VAE_feature_head; the reparameterisation noise is reproduced by seeding right before the forward."""
    from oracle.ref_import import load_reference_training
    T = load_reference_training()
    D, H, L, n = 48, 64, 16, 80
    torch.manual_seed(41)
    vae = ns.vae_model.VAE_feature_head(input_dim=D, latent_dim=L, hidden_dim=H)
    out = {k: v.copy() for k, v in sd_np(vae, "init.").items()}
    opt = torch.optim.Adam(vae.parameters(), lr=1e-3)
    g = torch.Generator().manual_seed(42)
    xs, epss, losses = [], [], []
    for step in range(3):
        x = torch.randn(n, D, generator=g)
        torch.manual_seed(200 + step)
        eps = torch.randn(n, L)
        torch.manual_seed(200 + step)
        vae.train()
        x_recon, mu, logvar, z, fp = vae(x, use_mean=False)
        loss, recon, kl, _ = T.vae_feature_loss(x_recon, x, mu, logvar, fp, None, alpha_recon=1.0, alpha_feature=0.0, beta=0.01)
        opt.zero_grad()
        loss.backward()
        if step == 0:
            out.update({"grad0." + k: v.grad.detach().numpy().copy() for k, v in vae.named_parameters()})
        opt.step()
        losses.append([float(loss), float(recon), float(kl)])
        xs.append(x.numpy()); epss.append(eps.numpy())
    out.update({"x": np.stack(xs), "eps": np.stack(epss), "losses": np.array(losses, np.float64)})
    out.update({"final." + k: v.detach().numpy() for k, v in vae.state_dict().items()})
    np.savez_compressed(os.path.join(OUT, "vae_train.npz"), **out)


def gen_kcenter(ns):
    """n2: select_topk_latent_diversity (select.py:185-270) on seeded latents, and select_programs with a
    diversity budget.  The candidate order the reference iterates (list(set)) is stored with the result."""
    g = torch.Generator().manual_seed(21)
    z = torch.randn(2000, 64, generator=g) * torch.rand(2000, 1, generator=g) * 3
    out = {"z": z.numpy()}
    rng = np.random.default_rng(22)
    perm = rng.permutation(2000)
    cases = {"a": (perm[:320], perm[320:620], 24), "b": (perm[700:800], np.zeros(0, np.int64), 10),
             "c": (perm[900:916], perm[1000:1003], 16)}
    for tag, (cand, used, k) in cases.items():
        cset = set(cand.tolist())
        order = np.array(list(cset), dtype=np.int64)
        sel, rest = ns.select.select_topk_latent_diversity(z, set(cset), set(used.tolist()), k)
        out.update({"cand_" + tag: order, "used_" + tag: np.array(list(set(used.tolist())), np.int64), "k_" + tag: np.int64(k),
                    "sel_" + tag: sel, "rest_" + tag: np.array(sorted(rest), np.int64)})
    sel_ref = np.load(os.path.join(OUT, "select.npz"))
    torch.manual_seed(11)
    reg = ns.vae_model.VAECostPredictor(48, hidden_dim=256, latent_dim=64).eval()
    reg.load_state_dict({k[4:]: torch.from_numpy(sel_ref[k]) for k in sel_ref.files if k.startswith("reg.")})
    sel_prog, rem_prog = ns.select.select_programs(
        reg, sel_ref["x"], set(sel_ref["used"].tolist()), set(sel_ref["remaining"].tolist()), num_select=64, T_mc=4,
        uncertainty_topk=10 ** 9, w_cost=0.5, w_unc=0.0, w_div=0.5, grad_num=2, rand_num=0,
        rng=np.random.default_rng(42), device=torch.device("cpu"), topk_factor=5)
    out.update({"sel_prog_div": sel_prog, "rem_prog_div": np.array(sorted(rem_prog), np.int64)})
    np.savez_compressed(os.path.join(OUT, "kcenter.npz"), **out)


def gen_acquisition(ns):
    """select_features (for_inference/tuning_improved.py:11-175) run unmodified on the shipped segment VAECostPredictor
    (for_inference/vae_latent_augmented_best.pth).  The two model methods it calls are absent from the reference's model
    class, so they are bound here from the reference's own layers: get_latent_embedding = encode()[0]; mc_predict = T
    passes through reg.cost_predictor with every Dropout module replaced by a recorded keep-mask."""
    from oracle.ref_import import load_reference_tuning
    tun = load_reference_tuning()
    ref = os.environ.get("VAEX_REFERENCE_ROOT", "/root/reference")
    ck = torch.load(os.path.join(ref, "scripts/pre_experiments/for_inference/vae_latent_augmented_best.pth"),
                    map_location="cpu", weights_only=False)
    cfg = ck["config"]
    reg = ns.seg_reg.VAECostPredictor(cfg["input_dim"], cfg["hidden_dim"], cfg["latent_dim"], cfg["predictor_hidden"],
                                      cfg["predictor_layers"], cfg["dropout"])
    reg.load_state_dict(ck["model_state_dict"])
    reg.eval()
    norm = np.asarray(ck["fea_norm_vec"], np.float32).reshape(-1)
    norm = np.where(norm == 0, 1.0, norm).astype(np.float32)
    rng = np.random.default_rng(4242)
    n_all, n_used, T_mc, num_select = 120, 24, 6, 8
    progs = synth_programs(rng, n_all, zero_frac=0.0)
    used = np.sort(rng.choice(n_all, n_used, replace=False)).astype(np.int64)
    p = float(cfg["dropout"])
    n_u, PH, ND = n_all - n_used, cfg["predictor_hidden"], cfg["predictor_layers"] - 1
    masks = (rng.random((T_mc, ND, n_u, PH)) >= p).astype(np.uint8)
    rec = {}

    def get_latent_embedding(segment_sizes, features, use_mean=True):
        with torch.no_grad():
            return reg.encode(segment_sizes, features)[0]

    def mc_predict(segment_sizes, features, T=20, use_mean=True):
        with torch.no_grad():
            z = reg.encode(segment_sizes, features)[0]
            outs = []
            for t in range(T):
                x, j = z, 0
                for mod in reg.cost_predictor:
                    if isinstance(mod, torch.nn.Dropout):
                        x = x * torch.from_numpy(masks[t, j]).float() / (1.0 - p)
                        j += 1
                    else:
                        x = mod(x)
                outs.append(x.squeeze(-1))
            o = torch.stack(outs)
            rec["z_mc"], rec["mu"], rec["var"] = z.numpy(), o.mean(0).numpy(), o.var(0).numpy()
            return o.mean(0), o.var(0)

    reg.get_latent_embedding, reg.mc_predict = get_latent_embedding, mc_predict
    sel = tun.select_features(reg, progs, set(used.tolist()), torch.from_numpy(norm), num_select=num_select, T_mc=T_mc,
                              w_cost=0.5, w_unc=0.3, w_dist=0.2, topk_factor=5)
    unl = [i for i in range(n_all) if i not in set(used.tolist())]
    with torch.no_grad():
        sz = torch.tensor([len(progs[i]) for i in unl])
        z_u = reg.encode(sz, torch.from_numpy(np.vstack([progs[i] / norm for i in unl])))[0].numpy()
        z_l = reg.encode(torch.tensor([len(progs[i]) for i in used]), torch.from_numpy(np.vstack([progs[i] / norm for i in used])))[0].numpy()
    np.savez_compressed(os.path.join(OUT, "acquisition.npz"), raw=np.concatenate(progs, 0),
                        sizes=np.array([len(q) for q in progs], np.int32), used=used, norm=norm,
                        masks=np.packbits(masks), masks_shape=np.array(masks.shape), p=np.float32(p),
                        num_select=np.int64(num_select), T_mc=np.int64(T_mc), sel=np.array(sel, np.int64),
                        z_u=z_u, z_l=z_l, z_mc=rec["z_mc"], mu=rec["mu"], var=rec["var"])


def gen_feature_blob(ns):
    """A packed per-store feature blob (layout of SerializeFeatures, feature.cc:1562-1622, written by
    oracle/feature_blob.serialize_features) through the reference's own unpack_feature (feature.py:49-158)."""
    from oracle import feature_blob as FB
    from oracle.ref_import import load_reference_feature
    ref = load_reference_feature()
    rng = np.random.default_rng(77)
    feats = []
    for i in range(41):
        if i in (0, 7, 40):
            feats.append(None)                              # failed lowering -> size 0
        elif i == 13:
            feats.append(np.zeros((3, 164), np.float32))    # lowered, but all-zero features
        else:
            f = (rng.random((int(rng.integers(1, 33)), 164)) * 24).astype(np.float32)
            f[:, rng.random(164) < 0.3] = 0
            feats.append(f)
    thr = rng.random(41).astype(np.float32)
    task_ids = rng.integers(0, 5, 41).astype(np.int32)
    min_costs = rng.random(5).astype(np.float32)
    blob = FB.serialize_features(feats, thr, task_ids, min_costs)
    f_ref, thr_ref, ids_ref, min_ref = ref.unpack_feature(bytearray(blob))
    out = {"blob": np.frombuffer(blob, np.uint8), "thr": np.asarray(thr_ref), "task_ids": np.asarray(ids_ref),
           "min_costs": np.asarray(min_ref), "sizes": np.array([len(f) for f in f_ref], np.int32),
           "flat": np.concatenate([np.asarray(f, np.float64) for f in f_ref], 0)}
    np.savez_compressed(os.path.join(OUT, "feature_blob.npz"), **out)


README_STATE = """Placeholder: placeholder, placeholder, placeholder
parallel ax0.0@ax1.0@ax2.0@ (0,4)
  for i1 (0,57)
    for i2 ((floormod(ax0.outer.outer.ax1.outer.outer.fused.ax2.outer.outer.fused, 4)*14),15)
      for i3 (0,64)
        PaddedInput = ...
  for ax3.0 (0,2)
    for ax2.1 (0,7)
      for ax3.1 (0,8)
        Conv2dOutput auto_unroll: 16
        for rx.0 (0,3)
          for rc.0 (0,4)
            for ry.1 (0,3)
              for rc.1 (0,16)
                for yy.3 (0,28)
                  vectorize ff.3 (0,8)
                    Conv2dOutput = ...
        for ax1.2 (0,28)
          vectorize ax3.2 (0,8)
            T_relu = ...
"""   # the printed program of the reference README (:86-108), dedented


def synth_states(rng, n, ragged):
    """Printed Ansor states in PrintState's format (src/auto_scheduler/loop_state.cc:354-440): one conv-like sketch
    with random tile extents; `ragged` adds the odd cases (no / several auto_unroll lines, trivial loops, non-zero
    minimums, undefined ranges, blanks after the comma, very long digit strings, leading / trailing blanks)."""
    ann = ["for", "parallel", "vectorize", "unroll", "blockIdx.x", "threadIdx.x", "vthread"]
    div = [1, 2, 3, 4, 7, 8, 14, 16, 28, 32, 56, 57, 64, 128, 512]
    out = []
    for i in range(n):
        lines = ["Placeholder: data, kernel" if i % 3 else "Placeholder: placeholder, placeholder, placeholder"]
        ind = 0
        depth = 14 if not ragged else int(rng.integers(3, 20))
        for j in range(depth):
            name = "%s.%d" % ("ijkxyc"[j % 6], j // 6)
            e = int(rng.choice(div))
            if ragged:
                r = rng.random()
                if r < 0.08:
                    lines.append(" " * ind + "for %s (%d,%d)" % (name, int(rng.integers(1, 9)), e)); ind += 2; continue
                if r < 0.12:
                    lines.append(" " * ind + "for %s (None)" % name); ind += 2; continue
                if r < 0.18:
                    lines.append(" " * ind + "for %s (0, %d)" % (name, e)); ind += 2; continue
                if r < 0.21:
                    lines.append(" " * ind + "for %s (0,%d)" % (name, int(rng.integers(10 ** 9, 10 ** 18)))); ind += 2; continue
                if r < 0.24:
                    lines.append(" " * ind + "for %s (0,\t %d)" % (name, e)); ind += 2; continue
                if r < 0.27:
                    lines.append(""); continue
                if r < 0.30 and ind >= 2:
                    lines.append(" " * ind + "A.local = ..."); ind -= 2; continue
            if j in (2, 5) or (ragged and rng.random() < 0.1):
                e = 1                                             # trivial loops: j = 2, 5 are common to all states
            lines.append(" " * ind + "%s %s (0,%d)" % (ann[int(rng.integers(0, len(ann)))] if j not in (2, 5) else "for", name, e))
            ind += 2
            if j == 8 and (not ragged or rng.random() < 0.6):
                lines.append(" " * ind + "conv auto_unroll: %d" % int(rng.choice([16, 64, 512, 1024])))
            if ragged and rng.random() < 0.05:
                lines.append(" " * ind + "pad auto_unroll:  %d" % int(rng.choice([0, 16, 64])))
        lines.append(" " * ind + "conv = ...")
        txt = "\n".join(lines) + "\n"
        if ragged and rng.random() < 0.3:
            txt = "  \n" + txt + " \t"
        out.append(txt)
    return out


def _pack_text(strs):
    enc = [s.encode() for s in strs]
    return (np.frombuffer(b"".join(enc), dtype=np.uint8).copy(),
            np.concatenate([[0], np.cumsum([len(b) for b in enc])]).astype(np.int64))


def gen_extent_text(ns):
    """the reference's own state_to_records (utils/extent.py:94-136) on synthetic printed states and on the
    README's program; sklearn StandardScaler (the routine the reference calls, vae_extent_search.py:394-397) on the
    fixed-sketch pool."""
    import contextlib
    import io
    from sklearn.preprocessing import StandardScaler
    from oracle.ref_import import load_reference_extent
    ext = load_reference_extent()
    rng = np.random.default_rng(2023)
    out = {}
    for tag, strs in (("rag", synth_states(rng, 400, True) + [README_STATE, "no placeholder here (0,7) auto_unroll: 3", ""]),
                      ("pool", synth_states(rng, 3000, False))):
        with contextlib.redirect_stdout(io.StringIO()):
            rec = ext.state_to_records(strs)          # list in, so the order is kept (the reference passes a set)
        out[tag + "_text"], out[tag + "_offsets"] = _pack_text(strs)
        out[tag + "_len"] = np.array([len(a) for a in rec["all"]], dtype=np.int32)
        out[tag + "_n_ext"] = np.array([len(a) for a in rec["extents"]], dtype=np.int32)
        out[tag + "_all"] = np.concatenate(rec["all"]).astype(np.float32)
        out[tag + "_clean_text"], out[tag + "_clean_offsets"] = _pack_text(rec["cleaned_schedules"])
    raw = np.array(rec["all"], dtype=np.float32)
    raw = np.concatenate([raw, np.full((raw.shape[0], 1), 7.0, np.float32)], axis=1)   # + a constant column
    x = np.log1p(raw)
    sc = StandardScaler()
    xs = sc.fit_transform(x)
    out.update(scaler_raw=raw, scaler_mean=sc.mean_, scaler_var=sc.var_, scaler_scale=sc.scale_, scaler_out=xs.astype(np.float32))
    out["readme_extents"] = np.array([4, 57, 64, 2, 7, 8, 3, 4, 3, 16, 28, 8, 28, 8, 16], dtype=np.float32)
    np.savez_compressed(os.path.join(OUT, "extent_text.npz"), **out)


def gen_segvae_train(ns):
    """SURVEY 8a a15, segment family: three training steps of the reference's SegmentVAE (models/vae.py) exactly as
    VAE_Trainer.train_vae runs them (trainer.py:57-66): train-mode forward, VAE_Trainer.vae_loss, backward,
    clip_grad_norm_(1.0), AdamW(lr, weight_decay=1e-5).  The reparameterisation noise and the dropout keep-mask of each
    step are captured from the run so that the CUDA path can replay them."""
    from oracle.ref_import import load_reference_seg_trainer
    ref_tr = load_reference_seg_trainer()
    torch.manual_seed(7)
    rng = np.random.default_rng(7)
    F, H, L, P = 164, 256, 64, 40
    model = ns.seg_vae.SegmentVAE(F, H, L, dropout=0.1)
    with torch.no_grad():
        for bn in (model.norm, model.decoder[1], model.decoder[5]):
            bn.weight.copy_(1 + 0.2 * torch.randn(H))
            bn.bias.copy_(0.1 * torch.randn(H))
            bn.running_mean.copy_(0.05 * torch.randn(H))
            bn.running_var.copy_(1 + 0.1 * torch.rand(H))
    out = {"w." + k: v.detach().clone().numpy() for k, v in model.state_dict().items()}
    sizes = rng.integers(1, 13, size=P).astype(np.int64)
    feats = (rng.random((int(sizes.sum()), F), dtype=np.float32) * 1.5).astype(np.float32)
    feats[:, rng.random(F) < 0.15] = 0.0
    out["sizes"], out["feats"] = sizes.astype(np.int32), feats
    beta, lr = 0.1, 1e-3
    out["hyper"] = np.array([beta, lr, 1e-5, 1.0, 0.1], dtype=np.float64)       # beta, lr, weight_decay, max_norm, p_drop
    opt = torch.optim.AdamW(model.parameters(), lr=lr, weight_decay=1e-5)
    captured = {}
    real_randn_like = torch.randn_like

    def randn_like(t, *a, **k):
        captured["eps"] = real_randn_like(t, *a, **k)
        return captured["eps"]

    def hook(mod, inp, outp):
        captured["mask"] = ((outp != 0) | (inp[0] == 0)).to(torch.uint8)

    hd = model.decoder[3].register_forward_hook(hook)
    model.train()
    ts, tf = torch.from_numpy(sizes), torch.from_numpy(feats)
    eps_l, mask_l, loss_l, gn_l = [], [], [], []
    for step in range(3):
        torch.randn_like = randn_like
        try:
            mean, logvar, z, recon, segsum = model(ts, tf)
        finally:
            torch.randn_like = real_randn_like
        loss, recon_loss, kld = ref_tr.VAE_Trainer.vae_loss(None, recon, segsum, mean, logvar, beta)
        opt.zero_grad()
        loss.backward()
        if step == 0:
            out.update({"g." + k: p.grad.detach().clone().numpy() for k, p in model.named_parameters()})
            out["mean0"], out["recon0"], out["segsum0"] = mean.detach().numpy(), recon.detach().numpy(), segsum.detach().numpy()
        gn_l.append(float(torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm=1.0)))
        opt.step()
        eps_l.append(captured["eps"].numpy().copy())
        mask_l.append(captured["mask"].numpy().copy())
        loss_l.append([loss.item(), recon_loss.item(), kld.item()])
    hd.remove()
    out["eps"], out["mask"] = np.stack(eps_l), np.stack(mask_l)
    out["losses"], out["gnorm"] = np.array(loss_l, dtype=np.float64), np.array(gn_l, dtype=np.float64)
    out.update({"p3." + k: v.detach().clone().numpy() for k, v in model.state_dict().items()})
    np.savez_compressed(os.path.join(OUT, "segvae_train.npz"), **out)


def gen_segreg_train(ns):
    """SURVEY 8a a15, segment family, regression: three steps of the reference's segment VAECostPredictor
    (models/regression.py) exactly as Regression_Trainer.train_phase runs them (trainer.py:446-455): model.train(),
    Regression_Trainer.compute_total_loss (three train-mode predictor passes), backward, clip_grad_norm_(1.0), AdamW with the
    encoder / predictor groups.  The smoothness noise and the six dropout keep-masks of each step (3 passes x 2 Dropout
    layers, in call order) are captured from the run."""
    from oracle.ref_import import load_reference_seg_trainer
    ref_tr = load_reference_seg_trainer()
    torch.manual_seed(11)
    rng = np.random.default_rng(11)
    F, H, L, PH, P = 164, 256, 64, 256, 36
    model = ns.seg_reg.VAECostPredictor(F, H, L, PH, 3, 0.1)
    bns = [model.norm, model.cost_predictor[1], model.cost_predictor[5], model.cost_predictor[9]]
    with torch.no_grad():
        for bn in bns:
            bn.weight.copy_(1 + 0.2 * torch.randn(bn.num_features))
            bn.bias.copy_(0.1 * torch.randn(bn.num_features))
            bn.running_mean.copy_(0.05 * torch.randn(bn.num_features))
            bn.running_var.copy_(1 + 0.1 * torch.rand(bn.num_features))
    out = {"w." + k: v.detach().clone().numpy() for k, v in model.state_dict().items()}
    sizes = rng.integers(1, 13, size=P).astype(np.int64)
    feats = (rng.random((int(sizes.sum()), F), dtype=np.float32) * 1.5).astype(np.float32)
    feats[:, rng.random(F) < 0.15] = 0.0
    labels = rng.random(P).astype(np.float32)
    labels[5] = labels[9]                                       # a tie: sign(0) = 0 pairs contribute `margin`
    out["sizes"], out["feats"], out["labels"] = sizes.astype(np.int32), feats, labels
    config = {"encoder_lr": 3e-4, "predictor_lr": 1e-3, "lambda_pair": 0.3, "gamma": 0.05, "beta": 0.01, "margin": 0.1, "noise_std": 0.1}
    out["hyper"] = np.array([config[k] for k in ("encoder_lr", "predictor_lr", "lambda_pair", "gamma", "beta", "margin", "noise_std")] + [0.1],
                            dtype=np.float64)
    opt = torch.optim.AdamW([{"params": model.get_encoder_params(), "lr": config["encoder_lr"]},
                             {"params": model.get_predictor_params(), "lr": config["predictor_lr"]}], weight_decay=1e-5)
    trainer = object.__new__(ref_tr.Regression_Trainer)         # the loss methods use no instance state
    cap = {"masks": []}
    real_randn_like = torch.randn_like

    def randn_like(t, *a, **k):
        cap["eps"] = real_randn_like(t, *a, **k)
        return cap["eps"]

    def hook(mod, inp, outp):
        cap["masks"].append(((outp != 0) | (inp[0] == 0)).to(torch.uint8).numpy().copy())

    hooks = [model.cost_predictor[3].register_forward_hook(hook), model.cost_predictor[7].register_forward_hook(hook)]
    ts, tf, tl = torch.from_numpy(sizes), torch.from_numpy(feats), torch.from_numpy(labels)
    noise_l, mask_l, loss_l, gn_l = [], [], [], []
    for step in range(3):
        model.train()
        opt.zero_grad()
        cap["masks"] = []
        torch.randn_like = randn_like
        try:
            total, comp = trainer.compute_total_loss(model, ts, tf, tl, config, return_components=True)
        finally:
            torch.randn_like = real_randn_like
        total.backward()
        if step == 0:
            out.update({"g." + k: p.grad.detach().clone().numpy() for k, p in model.named_parameters()})
        gn_l.append(float(torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm=1.0)))
        opt.step()
        assert len(cap["masks"]) == 6                            # call order: pass0 (d0, d1), pass1 (d0, d1), pass2 (d0, d1)
        noise_l.append((config["noise_std"] * cap["eps"]).numpy().copy())
        mask_l.append(np.stack(cap["masks"]).reshape(3, 2, P, PH))
        loss_l.append([comp["total_loss"], comp["reg_loss"], comp["pair_loss"], comp["smooth_loss"], comp["kld_loss"]])
    for h in hooks:
        h.remove()
    out["noise"], out["mask"] = np.stack(noise_l), np.stack(mask_l)
    out["losses"], out["gnorm"] = np.array(loss_l, dtype=np.float64), np.array(gn_l, dtype=np.float64)
    out.update({"p3." + k: v.detach().clone().numpy() for k, v in model.state_dict().items()})
    np.savez_compressed(os.path.join(OUT, "segreg_train.npz"), **out)


def gen_select_unc(ns):
    """a13: select_topk_uncertainty and select_programs with an uncertainty budget (w_unc > 0), run by the reference
    itself (utils/select.py:151-182,277-376 -> mc_predict, utils/model.py:241-273).  The predictor's nn.Dropout is
    replaced by a subclass that takes its keep-masks from a recorded Bernoulli(0.9) draw instead of torch's private
    stream (one mask per mc_predict pass, same x * mask / (1 - p) arithmetic), so the CUDA path can replay it."""
    sel_ref = np.load(os.path.join(OUT, "select.npz"))
    x = sel_ref["x"]
    N, PH, T = x.shape[0], 256, 6

    class RecordedDropout(torch.nn.Dropout):
        def __init__(self, p, masks):
            super().__init__(p)
            self.masks, self.calls = masks, 0

        def forward(self, inp):
            if not self.training:
                return inp
            m = self.masks[self.calls % len(self.masks)]
            self.calls += 1
            return inp * m / (1.0 - self.p)

    def model_with(masks):
        torch.manual_seed(11)
        reg = ns.vae_model.VAECostPredictor(48, hidden_dim=256, latent_dim=64).eval()
        reg.load_state_dict({k[4:]: torch.from_numpy(sel_ref[k]) for k in sel_ref.files if k.startswith("reg.")})
        assert isinstance(reg.cost_predictor[2], torch.nn.Dropout)
        drop = RecordedDropout(reg.cost_predictor[2].p, masks)
        reg.cost_predictor[2] = drop
        return reg.eval(), drop

    gm = torch.Generator().manual_seed(313)
    masks = (torch.rand(T, N, PH, generator=gm) >= 0.1).float()
    out = {"masks": masks.numpy().astype(np.uint8), "T": np.int64(T)}
    used, remaining = set(sel_ref["used"].tolist()), set(sel_ref["remaining"].tolist())
    # the function on its own
    reg, drop = model_with(masks)
    with torch.no_grad():
        mc_mean, mc_var = reg.mc_predict(torch.from_numpy(x), T=T)
    assert drop.calls == T
    out.update({"mc_mean": mc_mean.numpy(), "mc_var": mc_var.numpy()})
    reg, drop = model_with(masks)
    sel, rem = ns.select.select_topk_uncertainty(reg, torch.from_numpy(x), set(remaining), 24, T_mc=T)
    assert not reg.training and drop.calls == T
    out.update({"unc24": sel, "unc24_rem": np.array(sorted(rem), np.int64)})
    # select_programs, both branches of the uncertainty stage (select.py:350-357): whole remaining pool when few
    # programs have been measured, the top-5k candidate pool afterwards
    for tag, unc_topk in (("pool", 10 ** 9), ("cand", 10)):
        reg, drop = model_with(masks)
        sel_prog, rem_prog = ns.select.select_programs(
            reg, x, set(used), set(remaining), num_select=64, T_mc=T, uncertainty_topk=unc_topk,
            w_cost=0.4, w_unc=0.4, w_div=0.2, grad_num=2, rand_num=0,
            rng=np.random.default_rng(42), device=torch.device("cpu"), topk_factor=5)
        assert drop.calls == T
        out.update({"sel_prog_" + tag: sel_prog, "rem_prog_" + tag: np.array(sorted(rem_prog), np.int64)})
    np.savez_compressed(os.path.join(OUT, "select_unc.npz"), **out)


def main():
    os.makedirs(OUT, exist_ok=True)
    ns = load_reference()
    gen_segsum_mlp(ns)
    gen_extent_vae(ns)
    gen_snap(ns)
    gen_select(ns)
    gen_segment_family(ns)
    gen_train(ns)
    gen_feature_blob(ns)
    gen_kcenter(ns)
    gen_reg_train(ns)
    gen_vae_train(ns)
    gen_extent_text(ns)
    gen_segvae_train(ns)
    gen_segreg_train(ns)
    gen_select_unc(ns)
    gen_acquisition(ns)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    if len(sys.argv) > 1:        # python oracle/gen_golden.py gen_select_unc  -> only that fixture
        _ns = load_reference()
        for _name in sys.argv[1:]:
            globals()[_name](_ns)
    else:
        main()
