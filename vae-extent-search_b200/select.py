"""Candidate selection of the active-learning loop -- mirror of
scripts/pre_experiments/model_myself/utils/select.py (same function names / signatures /
return values).  Scores stay on the GPU; every top-k is the library's deterministic radix
select (score desc, index asc) restricted by a device mask instead of
np.argsort(pred[list(set)])[-k:] on the host (select.py:54-81)."""
import numpy as np
import torch

from . import _lib, vae_model


def _mask_from(indices, n, device):
    m = torch.zeros(n, dtype=torch.uint8, device=device)
    if len(indices):
        m[torch.as_tensor(np.fromiter(indices, dtype=np.int64, count=len(indices)), device=device)] = 1
    return m


def random_select_indices(remaining_indices, select_size, rng=np.random.default_rng(42)):
    """select.py:31-40."""
    if select_size == 0:
        return np.array([], dtype=np.int64), remaining_indices
    random_indices = rng.choice(list(remaining_indices), size=select_size, replace=False)
    remaining_indices.difference_update(set(random_indices.tolist()))
    return random_indices, remaining_indices


def util_select_topk(predictions, remaining_indices, num_select):
    """select.py:54-81: indices of the num_select largest predictions among remaining_indices
    (returned in ascending score order like argsort()[-k:]); updates the set."""
    pred = predictions if isinstance(predictions, torch.Tensor) else torch.as_tensor(np.asarray(predictions, np.float32))
    if not pred.is_cuda:
        pred = pred.cuda()
    k = min(num_select, len(remaining_indices))
    _, idx = vae_model.topk(pred.float(), k, _mask_from(remaining_indices, pred.shape[0], pred.device))
    selected = idx.flip(0).cpu().numpy().astype(np.int64)
    remaining_indices.difference_update(selected.tolist())
    return selected, remaining_indices


def select_topk_cost(cost_pred, remaining_indices, num_select):
    """select.py:88-108."""
    if num_select == 0:
        return np.array([], dtype=np.int64), remaining_indices
    return util_select_topk(cost_pred.detach() if isinstance(cost_pred, torch.Tensor) else cost_pred,
                            remaining_indices, num_select)


def select_topk_z_grad(z, cost_pred, remaining_indices, num_select):
    """select.py:111-148: top-k of ||d cost.sum()/dz||_2.  cost_pred must come from
    model.predict_cost(z) with z.requires_grad (its backward is the fused CUDA gradient)."""
    if num_select == 0:
        return np.array([], dtype=np.int64), remaining_indices
    z_grad = torch.autograd.grad(outputs=cost_pred.sum(), inputs=z, retain_graph=False, create_graph=False)[0]
    norm = torch.norm(z_grad, dim=1)
    remaining_indices = set(remaining_indices)
    return util_select_topk(norm, remaining_indices, num_select)


def select_topk_uncertainty(model, input_tensor, remaining_indices, num_select, T_mc=10, mc_masks=None, mc_seed=None):
    """select.py:151-182: top-k of the MC-dropout variance.  mc_masks (uint8 [T_mc, N, PH], 1 = keep) replays a given
    Bernoulli draw instead of the kernel's Philox stream (parity tests); mc_seed fixes that stream."""
    if num_select == 0:
        return np.array([], dtype=np.int64), remaining_indices
    with torch.no_grad():
        _, mc_var = model.mc_predict(input_tensor, T=T_mc, masks=mc_masks, seed=mc_seed)
    return util_select_topk(mc_var, remaining_indices, num_select)


def select_topk_latent_diversity(z, candidate_indices, used_indices, select_n_div, chunk_size=1024, eps=1e-12):
    """select.py:185-270: k-center greedy on L2-normalised latents, centres = used_indices.
    One call: normalise, min distance to the centres, then all k greedy steps in a single CTA (vaex_kcenter_select)."""
    if select_n_div == 0 or len(candidate_indices) == 0:
        return np.array([], dtype=np.int64), candidate_indices
    if not z.is_cuda:
        raise _lib.VaexError("vaex_b200 needs CUDA tensors (no CPU fallback)")
    zc = z.detach().contiguous().float()
    cand = np.array(list(candidate_indices), dtype=np.int64)
    k = min(select_n_div, len(cand))
    used = np.array(list(used_indices), dtype=np.int64)
    cand_d = torch.from_numpy(cand).to(z.device)
    used_d = torch.from_numpy(used).to(z.device) if len(used) else None
    sel = torch.empty(k, dtype=torch.int64, device=z.device)
    _lib.check(_lib.load().vaex_kcenter_select(_lib.ptr(zc), zc.shape[0], zc.shape[1], _lib.ptr(cand_d), len(cand),
                                              _lib.ptr(used_d), len(used), k, float(eps), _lib.ptr(sel), _lib.stream_ptr()))
    diverse = sel.cpu().numpy()
    candidate_indices = set(candidate_indices)
    candidate_indices.difference_update(diverse.tolist())
    return diverse, candidate_indices


def select_programs(model, input_data_scaled, used_indices, remaining_indices, num_select=64, T_mc=10,
                    uncertainty_topk=128, w_cost=0.5, w_unc=0.3, w_div=0.2, grad_num=2, rand_num=0,
                    rng=np.random.default_rng(42), device=torch.device("cuda"), topk_factor=5, mc_masks=None, mc_seed=None):
    """select.py:277-376, same budget arithmetic and stage order (mc_masks / mc_seed: see select_topk_uncertainty)."""
    total = num_select
    budget = total - grad_num - rand_num
    if num_select == 0 and rand_num > 0:
        return random_select_indices(remaining_indices, rand_num, rng=rng)
    n_cost, n_unc, n_div = int(budget * w_cost), int(budget * w_unc), int(budget * w_div)
    s = n_cost + n_unc + n_div
    if s < budget:
        n_cost += budget - s
    x = torch.as_tensor(input_data_scaled, dtype=torch.float32, device=device)
    model.eval()
    with torch.no_grad():
        z, _ = model.encode(x)
    z = z.detach().requires_grad_(True)
    cost_pred = model.predict_cost(z).view(-1)
    k_pref = min(len(remaining_indices), total * topk_factor)
    _, pool = vae_model.topk(cost_pred.detach(), k_pref, _mask_from(remaining_indices, x.shape[0], x.device))
    candidate = set(pool.cpu().numpy().tolist())
    currently_used = set()
    sel, candidate = select_topk_cost(cost_pred, candidate, n_cost)
    currently_used.update(sel.tolist())
    sel, candidate = select_topk_z_grad(z, cost_pred, candidate, grad_num)
    currently_used.update(sel.tolist())
    if len(used_indices) >= uncertainty_topk:
        unc, candidate = select_topk_uncertainty(model, x, candidate, n_unc, T_mc=T_mc, mc_masks=mc_masks, mc_seed=mc_seed)
    else:
        pool_unc = set(remaining_indices)
        pool_unc.difference_update(currently_used)
        unc, _ = select_topk_uncertainty(model, x, pool_unc, n_unc, T_mc=T_mc, mc_masks=mc_masks, mc_seed=mc_seed)
        candidate.difference_update(unc.tolist())
    currently_used.update(unc.tolist())
    used_local = set(used_indices)
    used_local.update(currently_used)
    div, _ = select_topk_latent_diversity(z, candidate, used_local, n_div)
    currently_used.update(div.tolist())
    remaining_indices.difference_update(currently_used)
    rnd, remaining_indices = random_select_indices(remaining_indices, rand_num, rng=rng)
    currently_used.update(rnd.tolist())
    return np.array(sorted(currently_used), dtype=np.int64), remaining_indices


def _fps(z, cand, first_pos, k):
    """k farthest-point picks among rows z[cand] (cand None: all rows), starting from cand[first_pos] (vaex_fps_select)."""
    if not z.is_cuda:
        raise _lib.VaexError("vaex_b200 needs CUDA tensors (no CPU fallback)")
    zc = z.detach().contiguous().float()
    if cand is None or isinstance(cand, torch.Tensor):
        cand_d = None if cand is None else cand.to(device=zc.device, dtype=torch.int64).contiguous()
    else:
        cand_d = torch.as_tensor(np.asarray(cand, dtype=np.int64), device=zc.device)
    M = zc.shape[0] if cand_d is None else cand_d.shape[0]
    sel = torch.empty(k, dtype=torch.int64, device=zc.device)
    _lib.check(_lib.load().vaex_fps_select(_lib.ptr(zc), zc.shape[0], zc.shape[1], _lib.ptr(cand_d), M, int(first_pos), int(k),
                                          _lib.ptr(sel), _lib.stream_ptr()))
    return sel


def select_init_latent_diversity(model, input_data_scaled, remaining_indices, select_num, first=None):
    """select.py:381-434: the first batch of a search -- farthest-point sampling on the raw encoder means of the whole
    pool from a random first point.  The reference's body raises UnboundLocalError (it updates remaining_indices with
    diverse_indices one line before assigning it); this is the evident intent: picks removed from remaining_indices,
    (diverse_indices, remaining_indices) returned.  first: the first pick (default: torch.randint like the reference).
    Picks are unique here; the reference's argmax over un-masked distances could return a point twice when every
    remaining point coincides with a chosen one."""
    model.eval()
    x = torch.as_tensor(np.asarray(input_data_scaled), dtype=torch.float32).cuda()
    with torch.no_grad():
        z = model.encode(x, use_mean=True)
    N = z.shape[0]
    select_num = min(select_num, N)
    if select_num == 0:
        return np.array([], dtype=np.int64), remaining_indices
    if first is None:
        first = int(torch.randint(0, N, (1,)).item())
    diverse_indices = _fps(z, None, first, select_num).cpu().numpy().astype(np.int64)
    remaining_indices.difference_update(diverse_indices.tolist())
    return diverse_indices, remaining_indices


def select_features(model, features, used_indices, fea_norm_vec, num_select=64, T_mc=20,
                    w_cost=0.5, w_unc=0.3, w_dist=0.2, topk_factor=5, mc_masks=None, mc_seed=None, flags=0):
    """scripts/pre_experiments/for_inference/tuning_improved.py:11-175: the next programs to measure, by the acquisition
    score  w_cost norm(-mu) + w_unc norm(var) + w_dist norm(d_min)  over the unlabeled pool (mu, var: T_mc MC-dropout
    passes; d_min: latent distance to the nearest labeled program; norm: min-max to [0, 1]), then farthest-point
    sampling among the num_select * topk_factor best, starting from the best.  model: the segment VAECostPredictor
    (get_latent_embedding / mc_predict); features: one [segments, input_dim] array per program.  As in the reference the
    latents are taken from features / fea_norm_vec and the MC passes from the raw features (:60 against :78-89).
    Everything after the host-side concatenation stays on the GPU; one device->host copy returns the picks.
    mc_masks / mc_seed: recorded keep-masks or the Philox seed of the MC passes; flags: VAEX_F_* of the row encoder."""
    all_n = len(features)
    used = set(used_indices)
    labeled_idx = sorted(used)
    unlabeled_idx = [i for i in range(all_n) if i not in used]
    if len(unlabeled_idx) == 0:
        return []
    if len(unlabeled_idx) <= num_select:
        return unlabeled_idx
    dev = next(model.parameters()).device
    if dev.type != "cuda":
        raise _lib.VaexError("vaex_b200 needs the model on a CUDA device (no CPU fallback)")

    def batch(indices):
        sizes = torch.as_tensor(np.array([np.shape(features[i])[0] for i in indices], dtype=np.int32), device=dev)
        rows = np.vstack([np.asarray(features[i], dtype=np.float32) for i in indices])
        return sizes, torch.from_numpy(rows).to(dev)

    model.eval()
    with torch.no_grad():
        su, fu = batch(unlabeled_idx)
        z_u = model.get_latent_embedding(su, fu, use_mean=True, fea_norm_vec=fea_norm_vec, flags=flags)
        mu_u, var_u = model.mc_predict(su, fu, T=T_mc, use_mean=True, masks=mc_masks, seed=mc_seed, flags=flags)
        if labeled_idx:
            sl, fl = batch(labeled_idx)
            z_l = model.get_latent_embedding(sl, fl, use_mean=True, fea_norm_vec=fea_norm_vec, flags=flags).contiguous()
            d_min = torch.empty(z_u.shape[0], device=dev)
            _lib.check(_lib.load().vaex_latent_min_dist(_lib.ptr(z_u), z_u.shape[0], _lib.ptr(z_l), z_l.shape[0], z_u.shape[1],
                                                       _lib.ptr(d_min), _lib.stream_ptr()))
        else:
            d_min = torch.ones(z_u.shape[0], device=dev)

        def norm01(x):
            lo, hi = x.min(), x.max()
            return (x - lo) / (hi - lo + 1e-8)

        scores = w_cost * norm01(-mu_u) + w_unc * norm01(var_u) + w_dist * norm01(d_min)
        K = min(num_select * topk_factor, len(unlabeled_idx))
        _, top = vae_model.topk(scores.float().contiguous(), K)          # score desc: position 0 is the best
        picks = _fps(z_u, top, 0, min(num_select, K))                    # ties in the score: lower index first here, np.argsort order there
    unl = np.asarray(unlabeled_idx, dtype=np.int64)
    return [int(i) for i in unl[picks.cpu().numpy()]]
