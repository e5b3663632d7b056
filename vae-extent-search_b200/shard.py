"""Multi-GPU sharding of the scoring path.

Programs are independent, so the path shards by contiguous blocks of programs with the weights
replicated and *no data-path collective*.  The only exchange is the candidate merge: every rank
contributes its k best (score, global index) pairs, one all-gather, one more top-k.  The merge
keeps the library's deterministic order (score desc, global index asc): ranks own ascending
blocks and every local list is already in that order, so "position in the gathered array" and
"global index" break ties identically.

The reference is single-GPU (utils/select.py:54-81 sorts one numpy array); this module is what
replaces that sort when the pool is spread over ranks."""
import ctypes

import numpy as np
import torch
import torch.distributed as dist

from . import _lib


class Comm:
    """The library's own NCCL communicator (include/vaex_b200.h: vaex_nccl_init / vaex_topk_allgather_merge /
    vaex_allreduce_grads / vaex_allgather).  One per process, created with the rank's CUDA device current.  The
    128-byte unique id of rank 0 travels over whatever torch.distributed group is up (any backend) or is passed in."""

    def __init__(self, rank=None, world=None, unique_id=None, group=None):
        self.lib = _lib.load()
        if world is None:
            world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
            rank = dist.get_rank(group) if world > 1 else 0
        self.rank, self.world = int(rank), int(world)
        idbuf = None
        if self.world > 1:
            if unique_id is None:
                box = [None]
                if self.rank == 0:
                    raw = (ctypes.c_uint8 * 128)()
                    _lib.check(self.lib.vaex_nccl_unique_id(raw))
                    box[0] = bytes(raw)
                dist.broadcast_object_list(box, src=0, group=group)
                unique_id = box[0]
            idbuf = (ctypes.c_uint8 * 128).from_buffer_copy(unique_id)
        self.h = _lib.vp()
        _lib.check(self.lib.vaex_nccl_init(self.rank, self.world, idbuf, ctypes.byref(self.h)))

    @staticmethod
    def unique_id():
        raw = (ctypes.c_uint8 * 128)()
        _lib.check(_lib.load().vaex_nccl_unique_id(raw))
        return bytes(raw)

    def merge_topk(self, local_values, local_indices, index_base, k, n_local=None):
        """-> (values fp32[k], global indices int64[k], count int32[1]) on the device, no host sync."""
        dev = local_values.device
        out_v = torch.empty(k, dtype=torch.float32, device=dev)
        out_i = torch.empty(k, dtype=torch.int64, device=dev)
        n_out = torch.empty(1, dtype=torch.int32, device=dev)
        lv = local_values.to(torch.float32).contiguous()
        li = local_indices.to(torch.int64).contiguous()
        if lv.shape[0] < k:                      # fewer winners than k: pad, the count says how many are real
            if n_local is None:
                n_local = torch.tensor([lv.shape[0]], dtype=torch.int32, device=dev)
            lv = torch.cat([lv, lv.new_zeros(k - lv.shape[0])])
            li = torch.cat([li, li.new_zeros(k - li.shape[0])])
        nl = None if n_local is None else n_local.to(device=dev, dtype=torch.int32).contiguous()
        _lib.check(self.lib.vaex_topk_allgather_merge(self.h, _lib.ptr(lv), _lib.ptr(li), _lib.ptr(nl), int(k), int(index_base),
                                                      _lib.ptr(out_v), _lib.ptr(out_i), _lib.ptr(n_out), _lib.stream_ptr()))
        return out_v, out_i, n_out

    def allreduce_grads(self, flat_grads):
        _lib.check(self.lib.vaex_allreduce_grads(self.h, _lib.ptr(flat_grads), flat_grads.numel(), _lib.stream_ptr()))
        return 1.0 / self.world

    def allgather(self, t):
        """every rank's contiguous tensor t -> [world, *t.shape]"""
        t = t.contiguous()
        out = torch.empty((self.world,) + tuple(t.shape), dtype=t.dtype, device=t.device)
        _lib.check(self.lib.vaex_allgather(self.h, _lib.ptr(t), _lib.ptr(out), t.numel() * t.element_size(), _lib.stream_ptr()))
        return out

    def close(self):
        if self.h:
            self.lib.vaex_nccl_destroy(self.h)
            self.h = _lib.vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def shard_range(n, rank, world):
    """Contiguous balanced block [lo, hi) of n items for `rank` (first n % world ranks get one more)."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world of %d" % (rank, world))
    base, extra = divmod(int(n), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_by_rows(sizes, world):
    """Program boundaries [world+1] that balance the number of feature *rows* per rank (the
    row encoder's work is per row, not per program).  Contiguous, monotone, covers everything."""
    sizes = np.asarray(sizes, dtype=np.int64)
    ends = np.cumsum(sizes)
    total = int(ends[-1]) if len(ends) else 0
    cuts = [0]
    for r in range(1, world):
        target = total * r / world
        c = int(np.searchsorted(ends, target, side="left"))
        # pick the nearer side of the target, keep monotone
        if c < len(ends) and c > 0 and abs(ends[c - 1] - target) <= abs(ends[c] - target):
            c -= 1
        cuts.append(max(cuts[-1], min(c + 1, len(sizes))))
    cuts.append(len(sizes))
    return np.asarray(cuts, dtype=np.int64)


def reduce_gradients(flat_grads, group=None, comm=None):
    """Data-parallel training: one sum all-reduce of the flat fp32 gradient buffer
    (NCCL over NVLink on GPUs, 0.96 MB for the SegmentSumMLP); returns the factor 1 / world that the
    optimizer kernel folds in, so every rank applies the mean gradient of the global batch.
    comm (a shard.Comm) routes it through the library's own communicator (vaex_allreduce_grads)."""
    if comm is not None:
        return comm.allreduce_grads(flat_grads)
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    if world > 1:
        dist.all_reduce(flat_grads, op=dist.ReduceOp.SUM, group=group)
    return 1.0 / world


def latent_ascent(ascent_fn, z0, comm=None, group=None):
    """Config 3 sharded: restarts are independent, so rank r runs ascent_fn on its contiguous block
    of z0's rows (ascent_fn(z_block) -> (z*, score)) and ONE all-gather of (score, z*) per restart gives every rank
    the full result.  comm (a shard.Comm) = the library's NCCL communicator; otherwise torch.distributed (any backend).
    -> (z* [R, L], score [R]) identical on every rank."""
    if comm is not None:
        rank, world = comm.rank, comm.world
    else:
        world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        rank = dist.get_rank(group) if world > 1 else 0
    R, Ld = z0.shape
    lo, hi = shard_range(R, rank, world)
    z, score = ascent_fn(z0[lo:hi].contiguous())
    if world == 1:
        return z, score
    per = -(-R // world)                                   # every rank sends the same number of rows (padded)
    send = torch.zeros(per, Ld + 1, dtype=torch.float32, device=z0.device)
    send[:hi - lo, 0] = score
    send[:hi - lo, 1:] = z
    if comm is not None:
        allr = comm.allgather(send)
    else:
        allr = torch.empty(world, per, Ld + 1, dtype=torch.float32, device=z0.device)
        dist.all_gather_into_tensor(allr.view(-1), send.view(-1), group=group)
    parts = []
    for r in range(world):
        a, b = shard_range(R, r, world)
        parts.append(allr[r, :b - a])
    full = torch.cat(parts, dim=0)
    return full[:, 1:].contiguous(), full[:, 0].contiguous()


def _default_topk(values, k, mask):
    from . import vae_model
    return vae_model.topk(values, k, mask, sync=False)


def merge_topk(local_values, local_indices, index_base, k, group=None, topk_fn=None, n_local=None, sync=True, comm=None):
    """Global top-k from per-rank winners.  With comm (a shard.Comm) the whole merge is one C-ABI call
    (vaex_topk_allgather_merge: pack, one ncclAllGather, one merge CTA); otherwise torch.distributed collectives
    (any backend -- the gloo tests on CPU) and the top-k given by topk_fn.

    local_values [m<=k] (score desc, index asc) and local_indices [m] (rank-local int64) are
    this rank's winners; index_base is the global index of the rank's first program.  When the
    winner count is only known on the device pass k-sized tensors and n_local (int32 tensor).
    Returns (values, global_indices) of the k best over all ranks, identical on every rank.
    Ranks with fewer than k winners pad with index -1; padding is masked out of the merge, so it
    never displaces a real entry (not even a real -inf score).  sync=False returns k-sized
    tensors whose unused tail has index -1 and does not wait for the GPU.
    `topk_fn(values, k, mask) -> (values, positions, count)` defaults to the CUDA radix select
    (tests on CPU inject the oracle's)."""
    m = int(local_values.shape[0])
    if m > k:
        raise ValueError("more local winners (%d) than k (%d)" % (m, k))
    if comm is not None:
        vals, idx, count = comm.merge_topk(local_values, local_indices, index_base, k, n_local)
        if not sync:
            return vals, idx
        n = int(count.item())
        return vals[:n], idx[:n]
    topk_fn = topk_fn or _default_topk
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    dev = local_values.device
    gi_local = local_indices.to(torch.int64) + int(index_base)
    if n_local is not None:
        gi_local = torch.where(torch.arange(m, device=dev) < n_local.to(dev), gi_local, gi_local.new_full((), -1))
    if m < k:
        v = torch.full((k,), float("-inf"), dtype=torch.float32, device=dev)
        i = torch.full((k,), -1, dtype=torch.int64, device=dev)
        v[:m] = local_values
        i[:m] = gi_local
    else:
        v, i = local_values.to(torch.float32).contiguous(), gi_local
    if world > 1:
        gv = torch.empty(world * k, dtype=torch.float32, device=dev)
        gi = torch.empty(world * k, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(gv, v, group=group)
        dist.all_gather_into_tensor(gi, i, group=group)
    else:
        gv, gi = v, i
    vals, pos, count = topk_fn(gv, k, (gi >= 0).to(torch.uint8))
    kk = vals.shape[0]
    live = torch.arange(kk, device=dev) < count.to(dev)
    idx = torch.where(live, gi[pos.clamp(0, gi.shape[0] - 1)], gi.new_full((), -1))
    if not sync:
        return vals, idx
    n = int(count.item())
    return vals[:n], idx[:n]
