"""TEST INFRASTRUCTURE ONLY -- CPU restatement (oracle) of the reference hot path.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this file.  The product package never does.

The reference implements this path as PyTorch eager fp32 modules; this file restates the
same arithmetic as plain functions over ``state_dict``-style dicts (same key names as the
reference modules, so shipped checkpoints and golden fixtures plug in directly) using torch
CPU fp32 ops.  Each function cites the reference lines it follows (relative to
/root/reference).  Parity pin: ``tests/test_oracle_golden.py`` checks every function here
against ``tests/golden/*.npz`` -- outputs of the *unmodified* reference modules generated
by ``oracle/gen_golden.py`` in the build container (the reference's own tests hold no
vectors for this path).
"""
import numpy as np
import torch

F32 = torch.float32


def _t(x, dtype=F32):
    if isinstance(x, torch.Tensor):
        return x if x.requires_grad else x.detach().to("cpu", dtype)
    return torch.as_tensor(np.asarray(x), dtype=dtype)


def _lin(x, W, b):
    return torch.addmm(_t(b), x, _t(W).t())


# --------------------------------------------------------------------------------------
# a1  SegmentDataLoader: flatten / embed / normalise / offsets
#     python/tvm/auto_scheduler/cost_model/mlp_model.py:47-92 (flatten), :95-105 (normalize)
# --------------------------------------------------------------------------------------
def flatten_programs(programs, task_embedding=None):
    """programs: sequence of [S_i, F] float arrays -> (features [sum S, F(+E)], sizes int32[N])."""
    sizes = np.array([len(p) for p in programs], dtype=np.int32)
    rows = []
    for p in programs:
        p = np.asarray(p, dtype=np.float32)
        if task_embedding is not None:
            emb = np.tile(np.asarray(task_embedding, np.float32), (len(p), 1))
            p = np.concatenate([p, emb], axis=1)
        rows.append(p)
    feats = np.concatenate(rows, axis=0).astype(np.float32) if rows else np.zeros((0, 0), np.float32)
    return feats, sizes


def fea_norm_vec(features):
    """mlp_model.py:95-103: per-column max where max > 0 else 1."""
    f = _t(features)
    mx = f.max(dim=0).values
    return torch.where(mx > 0, mx, torch.ones_like(mx))


def segment_offsets(sizes):
    """mlp_model.py:91-92: exclusive scan of segment sizes."""
    s = np.asarray(sizes, dtype=np.int64)
    return np.concatenate([[0], np.cumsum(s)]).astype(np.int64)


# --------------------------------------------------------------------------------------
# a2  SegmentSumMLPModule.forward   mlp_model.py:179-221 (ctor :148-173)
# --------------------------------------------------------------------------------------
def segment_sum(rows, sizes):
    """mlp_model.py:195-208 (repeat_interleave + scatter_add) == per-program row sum."""
    sizes = _t(sizes, torch.int64)
    n = sizes.shape[0]
    idx = torch.repeat_interleave(torch.arange(n), sizes)
    out = torch.zeros((n, rows.shape[1]), dtype=rows.dtype)
    return out.index_add_(0, idx, rows)


def segsum_mlp_encode_rows(P, feats):
    h = torch.relu(_lin(_t(feats), P["segment_encoder.0.weight"], P["segment_encoder.0.bias"]))
    return torch.relu(_lin(h, P["segment_encoder.2.weight"], P["segment_encoder.2.bias"]))


def _bn_eval(x, P, pre):
    """BatchNorm1d in eval mode (mlp_model.py:160-161, vae.py:40, regression.py:42)."""
    w, b = _t(P[pre + ".weight"]), _t(P[pre + ".bias"])
    mu, var = _t(P[pre + ".running_mean"]), _t(P[pre + ".running_var"])
    return (x - mu) / torch.sqrt(var + 1e-5) * w + b


def segsum_mlp_head(P, s):
    if "norm.weight" in P:
        s = _bn_eval(s, P, "norm")
    s = torch.relu(_lin(s, P["l0.0.weight"], P["l0.0.bias"])) + s
    s = torch.relu(_lin(s, P["l1.0.weight"], P["l1.0.bias"])) + s
    return s


def segsum_mlp_forward(P, sizes, feats, add_sigmoid=False):
    h = segsum_mlp_encode_rows(P, feats)
    s = segsum_mlp_head(P, segment_sum(h, sizes))
    out = _lin(s, P["decoder.weight"], P["decoder.bias"]).squeeze(-1)
    return torch.sigmoid(out) if add_sigmoid else out


# --------------------------------------------------------------------------------------
# a7/a8  SegmentVAE (scripts/vae_experiments/models/vae.py:87-137) and the segment VAECostPredictor
#        (models/regression.py:82-135), eval mode (BatchNorm running stats, Dropout = identity)
# --------------------------------------------------------------------------------------
def segvae_encode(P, sizes, feats):
    """-> (mean, logvar, segment_sum_vec): vae.py:118-128 / regression.py:82-101."""
    S = segment_sum(segsum_mlp_encode_rows(P, feats), sizes)
    h = segsum_mlp_head(P, S)
    return _lin(h, P["fc_mean.weight"], P["fc_mean.bias"]), _lin(h, P["fc_logvar.weight"], P["fc_logvar.bias"]), S


def segvae_decode(P, z):
    """vae.py:57-66,109-111: Linear, BN, ReLU, Dropout, Linear, BN, ReLU, Linear."""
    h = torch.relu(_bn_eval(_lin(_t(z), P["decoder.0.weight"], P["decoder.0.bias"]), P, "decoder.1"))
    h = torch.relu(_bn_eval(_lin(h, P["decoder.4.weight"], P["decoder.4.bias"]), P, "decoder.5"))
    return _lin(h, P["decoder.7.weight"], P["decoder.7.bias"])


def segcost_predict(P, z, n_layers=3):
    """regression.py:47-61,109-111: n x [Linear, BN, ReLU, Dropout] + Linear -> squeeze."""
    h = _t(z)
    for i in range(n_layers):
        pre = "cost_predictor.%d" % (4 * i)
        h = torch.relu(_bn_eval(_lin(h, P[pre + ".weight"], P[pre + ".bias"]), P, "cost_predictor.%d" % (4 * i + 1)))
    pre = "cost_predictor.%d" % (4 * n_layers)
    return _lin(h, P[pre + ".weight"], P[pre + ".bias"]).squeeze(-1)


# --------------------------------------------------------------------------------------
# a4  losses (mlp_model.py:863-944) and the optimisation step of _fit_a_model (:541-565)
#     restated with torch autograd on the CPU
# --------------------------------------------------------------------------------------
def loss_mse(preds, labels):
    return torch.mean((preds - labels) ** 2)


def loss_ranknet(preds, labels):
    """RankNetLoss, :863-876."""
    s_ij = preds - preds.unsqueeze(1)
    p_ij = torch.triu(1 / (torch.exp(s_ij) + 1), diagonal=1)
    y_ij = torch.triu((labels.unsqueeze(1) > labels).float(), diagonal=1)
    return torch.nn.functional.binary_cross_entropy(p_ij, y_ij)


def loss_lambdarank(preds, labels, eps=1e-10, sigma=1.0):
    """LambdaRankLoss.forward with k=None, :887-925."""
    y_pred, y_true = preds[None, :], labels[None, :]
    y_pred_sorted, idx = y_pred.sort(descending=True, dim=-1)
    y_true_sorted, _ = y_true.sort(descending=True, dim=-1)
    t = torch.gather(y_true, 1, idx)
    diffs = t[:, :, None] - t[:, None, :]
    mask = torch.isfinite(diffs) & (diffs > 0)
    t = t.clamp(min=0.0)
    y_true_sorted = y_true_sorted.clamp(min=0.0)
    D = torch.log2(1.0 + torch.arange(1, y_pred.shape[1] + 1).float())[None, :]
    max_dcg = torch.sum((torch.pow(2, y_true_sorted) - 1) / D, dim=-1).clamp(min=eps)
    G = (torch.pow(2, t) - 1) / max_dcg[:, None]
    w = torch.abs(torch.pow(D[:, :, None], -1.0) - torch.pow(D[:, None, :], -1.0)) * torch.abs(G[:, :, None] - G[:, None, :])
    sd = (y_pred_sorted[:, :, None] - y_pred_sorted[:, None, :]).clamp(min=-1e8, max=1e8)
    sd = torch.where(torch.isnan(sd), torch.zeros_like(sd), sd)
    wp = (torch.sigmoid(sigma * sd).clamp(min=eps) ** w).clamp(min=eps)
    return -torch.sum(torch.log2(wp)[mask])


def loss_listnet(preds, labels, eps=1e-10):
    """ListNetLoss, :928-944."""
    q = torch.softmax(preds[None, :], dim=1) + eps
    return torch.mean(-torch.sum(torch.softmax(labels[None, :], dim=1) * torch.log(q), dim=1))


LOSSES = {"rmse": loss_mse, "rankNetLoss": loss_ranknet, "lambdaRankLoss": loss_lambdarank, "listNetLoss": loss_listnet}
PARAM_ORDER = ["segment_encoder.0.weight", "segment_encoder.0.bias", "segment_encoder.2.weight", "segment_encoder.2.bias",
               "l0.0.weight", "l0.0.bias", "l1.0.weight", "l1.0.bias", "decoder.weight", "decoder.bias"]


def loss_and_grad(loss_type, preds, labels):
    p = _t(preds).clone().requires_grad_(True)
    loss = LOSSES[loss_type](p, _t(labels))
    (g,) = torch.autograd.grad(loss, p)
    return loss.detach(), g


def train_steps(P0, batches, loss_type, lr=7e-4, wd=1e-6, grad_clip=0.5, optimizer="adam"):
    """batches: list of (sizes, feats, labels).  Runs optimizer.zero_grad / loss.backward /
    clip_grad_norm_ / optimizer.step per batch (mlp_model.py:559-565; SGD variant :636-638).
    -> (final params dict, [loss per step], [pre-clip grad norm per step])"""
    P = {k: _t(P0[k]).clone().requires_grad_(True) for k in PARAM_ORDER}
    params = [P[k] for k in PARAM_ORDER]
    opt = torch.optim.Adam(params, lr=lr, weight_decay=wd) if optimizer == "adam" else torch.optim.SGD(params, lr=lr, weight_decay=wd)
    losses, norms = [], []
    for sizes, feats, labels in batches:
        opt.zero_grad()
        out = segsum_mlp_forward(P, sizes, _t(feats), add_sigmoid=(loss_type == "rmse"))
        loss = LOSSES[loss_type](out, _t(labels))
        loss.backward()
        if grad_clip:
            norms.append(float(torch.nn.utils.clip_grad_norm_(params, grad_clip)))
        opt.step()
        losses.append(float(loss))
    return {k: v.detach() for k, v in P.items()}, losses, norms


# --------------------------------------------------------------------------------------
# a3  MLPModelInternal._predict_a_task (mlp_model.py:669-681) + MLPModel.predict (:832-846)
# --------------------------------------------------------------------------------------
def predict_programs(P, programs, norm_vec=None, task_embedding=None, batch=4096,
                     add_sigmoid=False, mask_invalid=True):
    """Returns np.float32[N]; all-zero programs -> -inf (mlp_model.py:842-844)."""
    if P is None:
        return np.zeros(len(programs), dtype=np.float32)
    feats, sizes = flatten_programs(programs, task_embedding)
    feats = _t(feats)
    if norm_vec is not None:
        feats = feats / _t(norm_vec)
    offs = segment_offsets(sizes)
    preds = []
    for s in range(0, len(sizes), batch):
        e = min(s + batch, len(sizes))
        preds.append(segsum_mlp_forward(P, sizes[s:e], feats[offs[s]:offs[e]], add_sigmoid))
    out = torch.cat(preds).numpy().astype(np.float32) if preds else np.zeros(0, np.float32)
    if mask_invalid:
        for i, p in enumerate(programs):
            p = np.asarray(p)
            if p.min() == p.max() == 0:
                out[i] = -np.inf
    return out


# --------------------------------------------------------------------------------------
# a5  VAE_feature_head   scripts/pre_experiments/model_myself/utils/model.py:6-81
# a6  VAECostPredictor   same file :86-273
# --------------------------------------------------------------------------------------
def _mlp_relu(x, P, pre, idxs, last_act=True):
    for j, i in enumerate(idxs):
        x = _lin(x, P["%s.%d.weight" % (pre, i)], P["%s.%d.bias" % (pre, i)])
        if last_act or j + 1 < len(idxs):
            x = torch.relu(x)
    return x


def vae_encode(P, x):
    """model.py:52-56 / :145-160 -> (mu, logvar)."""
    h = _mlp_relu(_t(x), P, "encoder", (0, 2, 4))
    return _lin(h, P["fc_mu.weight"], P["fc_mu.bias"]), _lin(h, P["fc_logvar.weight"], P["fc_logvar.bias"])


def reparameterize(mu, logvar, eps):
    """model.py:58-61 with eps supplied (torch Philox stream is not reproducible elsewhere)."""
    return _t(mu) + _t(eps) * torch.exp(0.5 * _t(logvar))


def vae_decode(P, z):
    """model.py:28-38, :63-64: 3x(Linear+ReLU) + Linear."""
    return _mlp_relu(_t(z), P, "decoder", (0, 2, 4, 6), last_act=False)


def predict_cost(P, z, drop_mask=None, p_drop=0.1):
    """model.py:119-131, :168-170.  cost_predictor.{0,3,6} (2 layers) or the 3-layer segment
    variant regression.py:48-61 is handled by predict_cost_bn.  drop_mask [N,256] of 0/1
    reproduces train-mode Dropout(0.1) after the first ReLU (scaled by 1/(1-p))."""
    h = torch.relu(_lin(_t(z), P["cost_predictor.0.weight"], P["cost_predictor.0.bias"]))
    if drop_mask is not None:
        h = h * _t(drop_mask) / (1.0 - p_drop)
    h = torch.relu(_lin(h, P["cost_predictor.3.weight"], P["cost_predictor.3.bias"]))
    return _lin(h, P["cost_predictor.6.weight"], P["cost_predictor.6.bias"]).squeeze(-1)


def mc_predict(P, x, masks, p_drop=0.1):
    """model.py:241-273: T stochastic passes; mean and *unbiased* variance over T."""
    mu, _ = vae_encode(P, x)
    preds = torch.stack([predict_cost(P, mu, m, p_drop) for m in masks], dim=0)
    return preds.mean(dim=0), preds.var(dim=0)


def cost_grad_z(P, z):
    """utils/select.py:128-135: d(sum cost)/dz through the predictor (eval mode)."""
    z = _t(z).clone().requires_grad_(True)
    c = predict_cost(P, z)
    (g,) = torch.autograd.grad(c.sum(), z)
    return c.detach(), g


# --------------------------------------------------------------------------------------
# a9  postprocess + clamp   vae_sch_extent.ipynb:1374-1383,1457 ; inverse of
#     vae_extent_search.py:394-397 (log1p + StandardScaler)
# --------------------------------------------------------------------------------------
def snap_round_clamp(x, mean, scale, log1p=False):
    x = _t(x) * (_t(scale) + (0.0 if log1p else 1e-8)) + _t(mean)
    e = torch.expm1(x) if log1p else torch.exp(x)
    return torch.clamp(torch.round(e).to(torch.int32), min=1)


# --------------------------------------------------------------------------------------
# a10 tile-size legality: src/auto_scheduler/search_policy/utils.cc:446-463 (GetFactors),
#     sketch_policy_rules.cc:493-529, max_innermost_split_factor = 64 (search_policy.py:191)
# --------------------------------------------------------------------------------------
def get_factors(n):
    """Sorted divisors of n (utils.cc:446-463: loop to sqrt, push both, sort)."""
    res = []
    step = 2 if n % 2 == 1 else 1
    i = 1
    while i * i <= n:
        if n % i == 0:
            res.append(i)
            if n // i != i:
                res.append(n // i)
        i += step
    return sorted(res)


def nearest_divisor(n, v, cap=None):
    """Divisor of n nearest to v (ties -> smaller); optionally capped (innermost <= 64)."""
    best = 1
    for d in get_factors(int(n)):
        if cap is not None and d > cap:
            break
        if abs(d - v) < abs(best - v):
            best = d
    return best


def legalise_divisor(raw, slot_extent, slot_group_last, max_innermost=64):
    """raw int32 [N,D] (already round+clamp); per slot j: slot_extent[j] = loop extent the
    split belongs to (0 = free slot e.g. auto_unroll, kept as is); slots of one split are
    consecutive, outer->inner; slot_group_last[j]=1 marks the innermost length of a split.
    Each length is snapped to the nearest divisor of the *remaining* extent."""
    raw = np.asarray(raw, dtype=np.int64)
    out = raw.copy()
    N, D = raw.shape
    for i in range(N):
        rem = None
        for j in range(D):
            ext = int(slot_extent[j])
            if ext <= 0:
                rem = None
                continue
            if rem is None:
                rem = ext
            cap = max_innermost if slot_group_last[j] else None
            d = nearest_divisor(rem, int(raw[i, j]), cap)
            out[i, j] = d
            rem //= d
            if slot_group_last[j]:
                rem = None
    return out.astype(np.int32)


def legalise_divisor_fast(raw, slot_extent, slot_group_last, max_innermost=64):
    """legalise_divisor vectorised over the rows (numpy; same results, pinned against the loop version by
    tests/test_oracle_golden.py): per slot the rows are grouped by their remaining extent and snapped to the nearest
    entry of that extent's sorted divisor list (ties -> smaller) by one searchsorted.  This is what bench.py times as
    the CPU leg of the legalise stage -- the per-element Python loop above would flatter the GPU."""
    raw = np.asarray(raw, dtype=np.int64)
    out = raw.copy()
    N, D = raw.shape
    rem = None
    for j in range(D):
        ext = int(slot_extent[j])
        if ext <= 0:
            rem = None
            continue
        if rem is None:
            rem = np.full(N, ext, np.int64)
        last = bool(slot_group_last[j])
        v = raw[:, j]
        res = np.empty(N, np.int64)
        for r in np.unique(rem):
            divs = np.array([d for d in get_factors(int(r)) if not last or d <= max_innermost], dtype=np.int64)
            sel = rem == r
            vv = v[sel]
            pos = np.searchsorted(divs, vv)                              # first divisor >= v
            lo = divs[np.clip(pos - 1, 0, len(divs) - 1)]
            hi = divs[np.clip(pos, 0, len(divs) - 1)]
            res[sel] = np.where(pos == 0, hi, np.where(pos == len(divs), lo, np.where(vv - lo <= hi - vv, lo, hi)))
        out[:, j] = res
        rem = rem // res
        if last:
            rem = None
    return out.astype(np.int32)


# --------------------------------------------------------------------------------------
# a11 top-k   utils/select.py:54-81 (np.argsort(pred)[-k:], ascending order of score)
# --------------------------------------------------------------------------------------
def topk_desc(scores, k):
    """k largest, ordered (score desc, index asc) -- the kernel's deterministic tie-break.
    The reference's argsort is unstable so only the *set* is comparable when gaps are 0."""
    s = np.asarray(scores, dtype=np.float32)
    k = min(k, len(s))
    order = np.lexsort((np.arange(len(s)), -s.astype(np.float64)))
    idx = order[:k]
    return s[idx], idx.astype(np.int64)


# --------------------------------------------------------------------------------------
# config 3: multi-step latent ascent (NEW surface; single-step precedent select.py:128-135)
# --------------------------------------------------------------------------------------
def latent_ascent(P, z0, steps, lr):
    z = _t(z0).clone()
    for _ in range(steps):
        _, g = cost_grad_z(P, z)
        z = z + lr * g
    return z, predict_cost(P, z).detach()


def decoder_head_score(Pv, Pm, z):
    """The north-star ascent graph (BASELINE.json north_star (3)): the decoder's three hidden layers
    (utils/model.py:31-39) feeding the SegmentSumMLP residual head (mlp_model.py:210-221) on the 256-d vector."""
    a = _mlp_relu(z if isinstance(z, torch.Tensor) else _t(z), Pv, "decoder", (0, 2, 4))
    return _lin(segsum_mlp_head(Pm, a), Pm["decoder.weight"], Pm["decoder.bias"]).squeeze(-1)


def latent_ascent_decoder_head(Pv, Pm, z0, steps, lr):
    """plain torch.autograd loop around the composed modules"""
    z = _t(z0).clone()
    for _ in range(steps):
        zz = z.clone().requires_grad_(True)
        g = torch.autograd.grad(decoder_head_score(Pv, Pm, zz).sum(), zz)[0]
        z = z + lr * g
    return z, decoder_head_score(Pv, Pm, z).detach()


def kcenter_select(z, cand, used, k, eps=1e-12):
    """select_topk_latent_diversity (utils/select.py:185-270) on an explicit candidate order: k-center greedy
    over L2-normalised rows, centres initialised with `used`; returns the chosen entries of cand in order."""
    z = _t(z)
    zn = z / (z.norm(dim=1, keepdim=True) + eps)
    cand = np.asarray(cand, dtype=np.int64)
    zc = zn[torch.from_numpy(cand)]
    if len(used):
        min_d = torch.cdist(zc, zn[torch.from_numpy(np.asarray(used, dtype=np.int64))], p=2).min(dim=1).values
    else:
        min_d = torch.full((len(cand),), float("inf"))
    sel = []
    for _ in range(int(k)):
        j = int(torch.argmax(min_d))
        sel.append(cand[j])
        min_d[j] = -float("inf")
        min_d = torch.minimum(min_d, torch.cdist(zc, zc[j:j + 1], p=2).squeeze(1))
    return np.array(sel, dtype=np.int64)


def fps_select(z, cand, first_pos, k):
    """Farthest-point sampling on raw rows from a given first pick: the greedy loops of select_features
    (for_inference/tuning_improved.py:138-170) and select_init_latent_diversity (utils/select.py:413-428); lowest
    candidate position on ties (the strict > of :161, torch.argmax of :424).  Returns the chosen entries of cand."""
    z = np.asarray(z, dtype=np.float32)
    cand = np.arange(len(z), dtype=np.int64) if cand is None else np.asarray(cand, dtype=np.int64)
    zc = z[cand]
    dist = lambda j: np.sqrt(((zc - zc[j]) ** 2).sum(axis=1, dtype=np.float32), dtype=np.float32)
    sel = [int(first_pos)]
    min_d = dist(first_pos)
    min_d[first_pos] = -np.inf
    for _ in range(1, int(k)):
        j = int(np.argmax(min_d))
        sel.append(j)
        min_d = np.minimum(min_d, dist(j))
        min_d[j] = -np.inf
    return cand[np.array(sel, dtype=np.int64)]


def acquisition_select(z_u, z_l, mu, var, num_select, w_cost=0.5, w_unc=0.3, w_dist=0.2, topk_factor=5):
    """The scoring and selection arithmetic of select_features (tuning_improved.py:95-172) on given unlabeled / labeled
    latents and MC mean / variance: positions into the unlabeled pool, in pick order, and the scores."""
    z_u = np.asarray(z_u, np.float32)
    mu, var = np.asarray(mu, np.float32), np.asarray(var, np.float32)
    if z_l is not None and len(z_l):
        z_l = np.asarray(z_l, np.float32)
        d2 = (z_u ** 2).sum(1, keepdims=True) + (z_l ** 2).sum(1, keepdims=True).T - 2 * z_u @ z_l.T      # :103-106
        d_min = np.sqrt(np.maximum(d2, 0).min(axis=1))
    else:
        d_min = np.ones(len(z_u))
    norm01 = lambda x: (x - x.min()) / (x.max() - x.min() + 1e-8)
    scores = w_cost * norm01(-mu) + w_unc * norm01(var) + w_dist * norm01(d_min)
    K = min(num_select * topk_factor, len(z_u))
    top = np.argsort(scores)[::-1][:K]
    return fps_select(z_u, top, 0, min(num_select, K)), scores


def segcost_mc_predict(P, z, masks, p_drop):
    """T dropout passes of the segment cost predictor (regression.py:49-60: Linear, BatchNorm (eval), ReLU, Dropout after
    every block but the last, then Linear -> 1) with recorded keep-masks [T, blocks - 1, n, PH]: (mean, unbiased var)."""
    z = _t(z)
    masks = torch.as_tensor(np.asarray(masks)).float()
    nb = masks.shape[1] + 1
    outs = []
    for t in range(masks.shape[0]):
        x = z
        for i in range(nb):
            w, b = _t(P["cost_predictor.%d.weight" % (4 * i)]), _t(P["cost_predictor.%d.bias" % (4 * i)])
            bn = "cost_predictor.%d." % (4 * i + 1)
            x = x @ w.T + b
            x = (x - _t(P[bn + "running_mean"])) / torch.sqrt(_t(P[bn + "running_var"]) + 1e-5) * _t(P[bn + "weight"]) + _t(P[bn + "bias"])
            x = torch.relu(x)
            if i + 1 < nb:
                x = x * masks[t, i] / (1.0 - p_drop)
        outs.append((x @ _t(P["cost_predictor.%d.weight" % (4 * nb)]).T + _t(P["cost_predictor.%d.bias" % (4 * nb)])).squeeze(-1))
    o = torch.stack(outs)
    return o.mean(0), (o.var(0) if o.shape[0] > 1 else torch.zeros_like(o[0]))


# --------------------------------------------------------------------------------------
# a15  regression losses and training step of the extent-family VAECostPredictor
#      (utils/training.py:133-226; train_regression, vae_extent_search.py:1413-1440)
# --------------------------------------------------------------------------------------
REG_PARAM_ORDER = ["encoder.0.weight", "encoder.0.bias", "encoder.2.weight", "encoder.2.bias", "encoder.4.weight",
                   "encoder.4.bias", "fc_mu.weight", "fc_mu.bias", "fc_logvar.weight", "fc_logvar.bias",
                   "cost_predictor.0.weight", "cost_predictor.0.bias", "cost_predictor.3.weight", "cost_predictor.3.bias",
                   "cost_predictor.6.weight", "cost_predictor.6.bias"]
REG_ENCODER_KEYS = REG_PARAM_ORDER[:10]


def pair_loss(pred, target, margin=0.1):
    """pair_loss_fn, training.py:144-166: margin ranking over all pairs i < j."""
    pred, target = _t(pred).view(-1), _t(target).view(-1)
    n = pred.shape[0]
    if n < 2:
        return torch.tensor(0.0)
    i, j = torch.triu_indices(n, n, offset=1)
    y = torch.sign(target[j] - target[i])
    return torch.clamp(-y * (pred[j] - pred[i]) + margin, min=0).mean()


def kld_loss(mu, logvar):
    """kld_loss_fn, training.py:181-186."""
    mu, logvar = _t(mu), _t(logvar)
    return -0.5 * torch.mean(1 + logvar - mu.pow(2) - logvar.exp())


def reg_total_loss(P, x, labels, cfg, drop_mask=None, noise=None, p_drop=0.1):
    """forward(x, use_mean=True) + compute_total_loss (training.py:200-226).  drop_mask [n, PH] = the Bernoulli
    keep mask of the train-mode forward (None: eval); noise [n, L] = noise_std * randn of smooth_loss_fn, whose
    two predictor passes run in eval mode and whose noisy input is detached.
    -> (total, {reg, pair, smooth, kld}, cost_pred)"""
    mu, logvar = vae_encode(P, x)
    cost = predict_cost(P, mu, drop_mask, p_drop)
    y = _t(labels).view(-1)
    reg = torch.mean((cost - y) ** 2)
    pair = pair_loss(cost, y, cfg.get("margin", 0.1))
    z_noisy = (mu + (_t(noise) if noise is not None else 0.0)).detach()
    smooth = torch.mean((predict_cost(P, mu) - predict_cost(P, z_noisy)) ** 2)
    kld = kld_loss(mu, logvar)
    total = cfg["lambda_reg"] * reg + cfg["lambda_pair"] * pair + cfg["gamma"] * smooth + cfg["beta"] * kld
    return total, {"reg": reg, "pair": pair, "smooth": smooth, "kld": kld}, cost


def reg_train_steps(P0, batches, cfg, encoder_lr, predictor_lr, weight_decay=1e-5, clip=1.0, p_drop=0.1, encoder_freeze=False):
    """batches: list of (x, labels, drop_mask or None, noise or None).  zero_grad / backward / clip_grad_norm_(1.0) /
    AdamW with the encoder and predictor groups (vae_extent_search.py:1395-1398, :1435-1438).
    -> (final params, [[total, reg, pair, smooth, kld] per step], [pre-clip grad norm per step], first-step grads)"""
    P = {k: _t(P0[k]).clone().requires_grad_(True) for k in REG_PARAM_ORDER}
    enc = [P[k] for k in REG_ENCODER_KEYS]
    pred = [P[k] for k in REG_PARAM_ORDER[10:]]
    if encoder_freeze:          # vae_extent_search.py:1386-1391: requires_grad = False, predictor group only
        for t in enc:
            t.requires_grad_(False)
        opt = torch.optim.AdamW([{"params": pred, "lr": predictor_lr}], weight_decay=weight_decay)
    else:
        opt = torch.optim.AdamW([{"params": enc, "lr": encoder_lr}, {"params": pred, "lr": predictor_lr}], weight_decay=weight_decay)
    losses, norms, first = [], [], None
    for x, labels, mask, noise in batches:
        opt.zero_grad()
        total, comp, _ = reg_total_loss(P, x, labels, cfg, mask, noise, p_drop)
        total.backward()
        if first is None:
            first = {k: P[k].grad.detach().clone() for k in REG_PARAM_ORDER if P[k].grad is not None}
        norms.append(float(torch.nn.utils.clip_grad_norm_(pred if encoder_freeze else enc + pred, clip)))
        opt.step()
        losses.append([float(total)] + [float(comp[k]) for k in ("reg", "pair", "smooth", "kld")])
    return {k: v.detach() for k, v in P.items()}, losses, norms, first


VAE_PARAM_ORDER = ["encoder.0.weight", "encoder.0.bias", "encoder.2.weight", "encoder.2.bias", "encoder.4.weight", "encoder.4.bias",
                   "fc_mu.weight", "fc_mu.bias", "fc_logvar.weight", "fc_logvar.bias", "decoder.0.weight", "decoder.0.bias",
                   "decoder.2.weight", "decoder.2.bias", "decoder.4.weight", "decoder.4.bias", "decoder.6.weight", "decoder.6.bias"]


def vae_loss(P, x, eps, alpha_recon=1.0, beta=0.01):
    """forward(x, use_mean=False) with the given reparameterisation noise + vae_feature_loss (training.py:14-32,
    no feature head) -> (total, recon, kl)."""
    mu, logvar = vae_encode(P, x)
    z = mu if eps is None else mu + _t(eps) * torch.exp(0.5 * logvar)
    recon = torch.mean((vae_decode(P, z) - _t(x)) ** 2)
    kl = kld_loss(mu, logvar)
    return alpha_recon * recon + beta * kl, recon, kl


def vae_train_steps(P0, batches, alpha_recon=1.0, beta=0.01, lr=1e-3):
    """batches: list of (x, eps).  zero_grad / backward / Adam.step (training.py:65,84-88).
    -> (final params, [[total, recon, kl] per step], first-step grads)"""
    P = {k: _t(P0[k]).clone().requires_grad_(True) for k in VAE_PARAM_ORDER}
    params = [P[k] for k in VAE_PARAM_ORDER]
    opt = torch.optim.Adam(params, lr=lr)
    losses, first = [], None
    for x, eps in batches:
        opt.zero_grad()
        total, recon, kl = vae_loss(P, x, eps, alpha_recon, beta)
        total.backward()
        if first is None:
            first = {k: P[k].grad.detach().clone() for k in VAE_PARAM_ORDER}
        opt.step()
        losses.append([float(total), float(recon), float(kl)])
    return {k: v.detach() for k, v in P.items()}, losses, first


# --------------------------------------------------------------------------------------
# a15 (segment family)  SegmentVAE training step: train-mode forward (vae.py:113-137), VAE_Trainer.vae_loss
#      (scripts/vae_experiments/trainer.py:270-280), clip_grad_norm_(1.0) + AdamW(weight_decay=1e-5) (:40,57-66)
# --------------------------------------------------------------------------------------
SEGVAE_PARAM_ORDER = ["segment_encoder.0.weight", "segment_encoder.0.bias", "segment_encoder.2.weight", "segment_encoder.2.bias",
                      "norm.weight", "norm.bias", "l0.0.weight", "l0.0.bias", "l1.0.weight", "l1.0.bias", "fc_mean.weight",
                      "fc_mean.bias", "fc_logvar.weight", "fc_logvar.bias", "decoder.0.weight", "decoder.0.bias", "decoder.1.weight",
                      "decoder.1.bias", "decoder.4.weight", "decoder.4.bias", "decoder.5.weight", "decoder.5.bias", "decoder.7.weight",
                      "decoder.7.bias"]
SEGVAE_BN = ["norm", "decoder.1", "decoder.5"]


def _bn_train(x, P, pre, running=None, momentum=0.1):
    """BatchNorm1d with batch statistics: biased variance normalises, the unbiased one feeds running_var."""
    n = x.shape[0]
    mu = x.mean(dim=0)
    var = ((x - mu) ** 2).mean(dim=0)
    if running is not None:
        running[pre + ".running_mean"] = (1 - momentum) * running[pre + ".running_mean"] + momentum * mu.detach()
        running[pre + ".running_var"] = (1 - momentum) * running[pre + ".running_var"] + momentum * var.detach() * n / (n - 1)
    return (x - mu) / torch.sqrt(var + 1e-5) * P[pre + ".weight"] + P[pre + ".bias"]


def segvae_train_loss(P, sizes, feats, eps, drop_mask, beta, p_drop=0.1, running=None):
    """-> (total, recon_sum, kld_sum, mean, recon, segment_sum_vec).  eps: reparameterisation noise [P, L];
    drop_mask: keep mask [P, H] of decoder.3 (None = no dropout)."""
    S = segment_sum(segsum_mlp_encode_rows(P, feats), sizes)
    h = _bn_train(S, P, "norm", running)
    h = torch.relu(_lin(h, P["l0.0.weight"], P["l0.0.bias"])) + h
    h = torch.relu(_lin(h, P["l1.0.weight"], P["l1.0.bias"])) + h
    mean, logvar = _lin(h, P["fc_mean.weight"], P["fc_mean.bias"]), _lin(h, P["fc_logvar.weight"], P["fc_logvar.bias"])
    z = mean + _t(eps) * torch.exp(0.5 * logvar)
    d = torch.relu(_bn_train(_lin(z, P["decoder.0.weight"], P["decoder.0.bias"]), P, "decoder.1", running))
    if drop_mask is not None:
        d = d * _t(drop_mask) / (1.0 - p_drop)
    d = torch.relu(_bn_train(_lin(d, P["decoder.4.weight"], P["decoder.4.bias"]), P, "decoder.5", running))
    recon = _lin(d, P["decoder.7.weight"], P["decoder.7.bias"])
    recon_loss = ((recon - S) ** 2).sum()                          # the target keeps its graph (no detach in the reference)
    kld = -0.5 * torch.sum(1 + logvar - mean ** 2 - torch.exp(logvar))
    return recon_loss + beta * kld, recon_loss, kld, mean, recon, S


def segvae_train_steps(P0, sizes, feats, eps_list, mask_list, beta, lr, weight_decay=1e-5, clip=1.0, p_drop=0.1):
    """-> (final params, running stats, [[total, recon, kld]], first-step grads, grad norms)"""
    P = {k: _t(P0[k]).clone().requires_grad_(True) for k in SEGVAE_PARAM_ORDER}
    running = {b + s: _t(P0[b + s]).clone() for b in SEGVAE_BN for s in (".running_mean", ".running_var")}
    params = [P[k] for k in SEGVAE_PARAM_ORDER]
    opt = torch.optim.AdamW(params, lr=lr, weight_decay=weight_decay)
    losses, norms, first = [], [], None
    for eps, mask in zip(eps_list, mask_list):
        opt.zero_grad()
        total, recon, kld = segvae_train_loss(P, sizes, feats, eps, mask, beta, p_drop, running)[:3]
        total.backward()
        if first is None:
            first = {k: P[k].grad.detach().clone() for k in SEGVAE_PARAM_ORDER}
        norms.append(float(torch.nn.utils.clip_grad_norm_(params, clip)))
        opt.step()
        losses.append([float(total), float(recon), float(kld)])
    return {k: v.detach() for k, v in P.items()}, running, losses, first, norms


# --------------------------------------------------------------------------------------
# a15 (segment family, regression)  Regression_Trainer.compute_total_loss (scripts/vae_experiments/trainer.py:485-556) on the
#      segment VAECostPredictor in train mode (models/regression.py:82-135), clip + AdamW over two groups (trainer.py:401-455)
# --------------------------------------------------------------------------------------
def segreg_param_order(n_layers=3):
    enc = SEGVAE_PARAM_ORDER[:14]
    pred = []
    for i in range(n_layers):
        pred += ["cost_predictor.%d.weight" % (4 * i), "cost_predictor.%d.bias" % (4 * i),
                 "cost_predictor.%d.weight" % (4 * i + 1), "cost_predictor.%d.bias" % (4 * i + 1)]
    pred += ["cost_predictor.%d.weight" % (4 * n_layers), "cost_predictor.%d.bias" % (4 * n_layers)]
    return enc, pred


def _segreg_predict_train(P, z, masks, p_drop, running, n_layers=3):
    """One train-mode predict_cost pass: n x [Linear, BatchNorm (batch stats), ReLU, Dropout (all but the last)] + Linear."""
    h = z
    for i in range(n_layers):
        pre = "cost_predictor.%d" % (4 * i)
        h = torch.relu(_bn_train(_lin(h, P[pre + ".weight"], P[pre + ".bias"]), P, "cost_predictor.%d" % (4 * i + 1), running))
        if i < n_layers - 1 and masks is not None:
            h = h * _t(masks[i]) / (1.0 - p_drop)
    pre = "cost_predictor.%d" % (4 * n_layers)
    return _lin(h, P[pre + ".weight"], P[pre + ".bias"]).squeeze(-1)


def segreg_total_loss(P, sizes, feats, labels, cfg, noise, masks, p_drop=0.1, running=None, n_layers=3):
    """-> (total, reg, pair, smooth, kld).  noise [P, L] = noise_std * randn; masks [3 passes][n_layers - 1][P, PH] or None."""
    S = segment_sum(segsum_mlp_encode_rows(P, feats), sizes)
    h = _bn_train(S, P, "norm", running)
    h = torch.relu(_lin(h, P["l0.0.weight"], P["l0.0.bias"])) + h
    h = torch.relu(_lin(h, P["l1.0.weight"], P["l1.0.bias"])) + h
    mean, logvar = _lin(h, P["fc_mean.weight"], P["fc_mean.bias"]), _lin(h, P["fc_logvar.weight"], P["fc_logvar.bias"])
    z = mean                                                       # forward(use_mean=True)
    m = (lambda q: None if masks is None else masks[q])
    cost = _segreg_predict_train(P, z, m(0), p_drop, running, n_layers)
    y = _t(labels)
    reg = torch.mean((cost - y) ** 2)
    pair = pair_loss(cost, y, cfg.get("margin", 0.1))
    c1 = _segreg_predict_train(P, z, m(1), p_drop, running, n_layers)
    c2 = _segreg_predict_train(P, z + _t(noise), m(2), p_drop, running, n_layers)
    smooth = torch.mean((c1 - c2) ** 2)
    kld = kld_loss(mean, logvar)
    total = reg + cfg.get("lambda_pair", 0.1) * pair + cfg.get("gamma", 0.01) * smooth + cfg.get("beta", 0.001) * kld
    return total, reg, pair, smooth, kld


def segreg_train_steps(P0, sizes, feats, labels, cfg, noise_list, mask_list, encoder_lr, predictor_lr, weight_decay=1e-5, clip=1.0,
                       p_drop=0.1, n_layers=3):
    enc, pred = segreg_param_order(n_layers)
    P = {k: _t(P0[k]).clone().requires_grad_(True) for k in enc + pred}
    bns = ["norm"] + ["cost_predictor.%d" % (4 * i + 1) for i in range(n_layers)]
    running = {b + s: _t(P0[b + s]).clone() for b in bns for s in (".running_mean", ".running_var")}
    opt = torch.optim.AdamW([{"params": [P[k] for k in enc], "lr": encoder_lr}, {"params": [P[k] for k in pred], "lr": predictor_lr}],
                            weight_decay=weight_decay)
    losses, norms, first = [], [], None
    for noise, masks in zip(noise_list, mask_list):
        opt.zero_grad()
        vals = segreg_total_loss(P, sizes, feats, labels, cfg, noise, masks, p_drop, running, n_layers)
        vals[0].backward()
        if first is None:
            first = {k: P[k].grad.detach().clone() for k in enc + pred}
        norms.append(float(torch.nn.utils.clip_grad_norm_([P[k] for k in enc + pred], clip)))
        opt.step()
        losses.append([float(v) for v in vals])
    return {k: v.detach() for k, v in P.items()}, running, losses, first, norms
