"""N>1 host logic on CPU: block partition, row-balanced partition and the top-k merge over a
world_size-2 gloo group (the CUDA radix select is replaced by the oracle's top-k here; the GPU
version of the same merge is covered by bench.py --gpus N)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import vaex_b200
from vaex_b200 import shard
from oracle import torch_oracle as O


def test_shard_range_covers_everything():
    for n in (0, 1, 7, 8, 9, 1000003):
        for world in (1, 2, 3, 8):
            spans = [shard.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            lens = [b - a for a, b in spans]
            assert max(lens) - min(lens) <= 1
    with pytest.raises(ValueError):
        shard.shard_range(10, 2, 2)


def test_shard_by_rows_balances_rows():
    rng = np.random.default_rng(0)
    sizes = rng.integers(1, 33, size=20000)
    for world in (1, 2, 4, 8):
        cuts = shard.shard_by_rows(sizes, world)
        assert cuts[0] == 0 and cuts[-1] == len(sizes) and np.all(np.diff(cuts) >= 0)
        rows = np.add.reduceat(sizes, cuts[:-1])
        assert rows.sum() == sizes.sum()
        assert rows.max() - rows.min() <= 64          # within two maximal programs
    # degenerate: fewer programs than ranks
    cuts = shard.shard_by_rows([5, 5], 4)
    assert cuts[0] == 0 and cuts[-1] == 2 and np.all(np.diff(cuts) >= 0)


def _oracle_topk(values, k, mask):
    v = values.numpy().copy()
    keep = np.nonzero(mask.numpy())[0]
    vals, pos = O.topk_desc(v[keep], k)
    n = len(pos)                                    # k-sized outputs + count, like the CUDA call with sync=False
    ov, op = np.zeros(min(k, len(v)), np.float32), np.zeros(min(k, len(v)), np.int64)
    ov[:n], op[:n] = vals, keep[pos]
    return torch.from_numpy(ov), torch.from_numpy(op), torch.tensor([n], dtype=torch.int32)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, k, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(5)
        scores = rng.standard_normal(n).astype(np.float32)
        scores[rng.integers(0, n, size=n // 4)] = 0.5            # many exact ties across ranks
        scores[rng.integers(0, n, size=n // 50)] = -np.inf       # invalid programs
        lo, hi = shard.shard_range(n, rank, world)
        lv, li = O.topk_desc(scores[lo:hi], k)
        vals, idx = shard.merge_topk(torch.from_numpy(lv.copy()), torch.from_numpy(li), lo, k, topk_fn=_oracle_topk)
        np.savez(os.path.join(out_dir, "r%d.npz" % rank), vals=vals.numpy(), idx=idx.numpy(), scores=scores)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,k", [(5000, 320), (300, 320), (7, 4)])
def test_merge_topk_world2_gloo(tmp_path, n, k):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), n, k, str(tmp_path)), nprocs=world, join=True)
    r0, r1 = (np.load(tmp_path / ("r%d.npz" % r)) for r in range(world))
    want_v, want_i = O.topk_desc(r0["scores"], k)            # single-process answer over the whole pool
    for r in (r0, r1):
        assert np.array_equal(r["idx"], want_i)              # bit-exact incl. tie order
        assert np.array_equal(r["vals"], want_v)


def test_merge_topk_single_process_is_identity():
    v = torch.tensor([3.0, 2.0, 2.0])
    i = torch.tensor([4, 1, 9])
    vals, idx = shard.merge_topk(v, i, 100, 5, topk_fn=_oracle_topk)
    assert idx.tolist() == [104, 101, 109] and vals.tolist() == [3.0, 2.0, 2.0]
    with pytest.raises(ValueError):
        shard.merge_topk(v, i, 0, 2, topk_fn=_oracle_topk)


def test_merge_topk_device_count_and_nosync():
    v = torch.tensor([3.0, 2.0, 9.0, 9.0])          # last two are unspecified padding of the local top-k
    i = torch.tensor([4, 1, 7, 7])
    vals, idx = shard.merge_topk(v, i, 10, 4, topk_fn=_oracle_topk, n_local=torch.tensor([2], dtype=torch.int32), sync=False)
    assert idx.tolist() == [14, 11, -1, -1] and vals[:2].tolist() == [3.0, 2.0]


def _grad_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.arange(1000, dtype=torch.float32) * (rank + 1)      # rank-local gradient buffer
        scale = shard.reduce_gradients(g)
        np.savez(os.path.join(out_dir, "g%d.npz" % rank), g=g.numpy(), scale=scale)
    finally:
        dist.destroy_process_group()


def test_reduce_gradients_world2_gloo(tmp_path):
    """training's only collective: sum all-reduce of the flat gradient buffer + the 1/world factor"""
    mp.spawn(_grad_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    want = np.arange(1000, dtype=np.float32) * 3
    for r in range(2):
        z = np.load(tmp_path / ("g%d.npz" % r))
        assert np.array_equal(z["g"], want) and float(z["scale"]) == 0.5
    assert shard.reduce_gradients(torch.ones(4)) == 1.0                # no process group: identity


def _ascent_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        z0 = torch.arange(7 * 4, dtype=torch.float32).reshape(7, 4)           # 7 restarts over 2 ranks: blocks of 4 and 3
        z, score = shard.latent_ascent(lambda zb: (zb * 2 + 1, zb.sum(dim=1)), z0)
        np.savez(os.path.join(out_dir, "a%d.npz" % rank), z=z.numpy(), score=score.numpy())
    finally:
        dist.destroy_process_group()


def test_latent_ascent_world2_gloo(tmp_path):
    """config 3 sharded: each rank ascends its block of restarts, one all-gather of (score, z*) rebuilds the whole set"""
    mp.spawn(_ascent_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    z0 = np.arange(28, dtype=np.float32).reshape(7, 4)
    for r in range(2):
        a = np.load(tmp_path / ("a%d.npz" % r))
        assert np.array_equal(a["z"], z0 * 2 + 1) and np.array_equal(a["score"], z0.sum(1))
    z, s = shard.latent_ascent(lambda zb: (zb + 1, zb[:, 0]), torch.from_numpy(z0))     # no process group: identity sharding
    assert np.array_equal(z.numpy(), z0 + 1) and np.array_equal(s.numpy(), z0[:, 0])
