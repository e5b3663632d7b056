"""The C-ABI multi-GPU entry points (include/vaex_b200.h: vaex_nccl_init / vaex_topk_allgather_merge /
vaex_allreduce_grads / vaex_allgather) against the oracle's single-process top-k: world 1 on any GPU box, world 2
(no torch.distributed, no gloo: the unique id travels through a file) when two GPUs are visible."""
import os
import time

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from oracle import torch_oracle as O

import vaex_b200
from vaex_b200 import _lib, shard

pytestmark = pytest.mark.gpu


def _scores(n, seed=5):
    rng = np.random.default_rng(seed)
    s = rng.standard_normal(n).astype(np.float32)
    s[rng.integers(0, n, size=n // 4)] = 0.5            # exact ties across ranks
    s[rng.integers(0, n, size=max(1, n // 50))] = -np.inf
    return s


def _local_topk(scores_np, lo, hi, k, dev):
    sc = torch.from_numpy(scores_np[lo:hi]).to(dev)
    kk = min(k, hi - lo)
    v, i, cnt = vaex_b200.topk(sc, kk, sync=False)
    return v, i, cnt


@pytest.mark.parametrize("n,k", [(200000, 320), (300, 320), (7, 4), (5000, 1024)])
def test_merge_world1_equals_local_topk(n, k):
    comm = shard.Comm(rank=0, world=1)
    s = _scores(n)
    v, i, cnt = _local_topk(s, 0, n, k, "cuda")
    vals, idx = shard.merge_topk(v, i, 1000, k, n_local=cnt, comm=comm)
    want_v, want_i = O.topk_desc(s, k)
    assert np.array_equal(idx.cpu().numpy(), want_i + 1000)
    assert np.array_equal(vals.cpu().numpy(), want_v)
    g = torch.arange(1000, dtype=torch.float32, device="cuda")
    assert comm.allreduce_grads(g) == 1.0 and g[999].item() == 999.0
    assert torch.equal(comm.allgather(g)[0], g)
    with pytest.raises(_lib.VaexError):
        comm.merge_topk(v, i, 0, 2000)
    comm.close()


def _worker(rank, world, id_path, n, k, out_dir):
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    if rank == 0:
        with open(id_path + ".tmp", "wb") as f:
            f.write(shard.Comm.unique_id())
        os.replace(id_path + ".tmp", id_path)
    t0 = time.time()
    while not os.path.exists(id_path):
        assert time.time() - t0 < 60
        time.sleep(0.01)
    uid = open(id_path, "rb").read()
    comm = shard.Comm(rank=rank, world=world, unique_id=uid)
    s = _scores(n)
    lo, hi = shard.shard_range(n, rank, world)
    v, i, cnt = _local_topk(s, lo, hi, k, dev)
    vals, idx = shard.merge_topk(v, i, lo, k, n_local=cnt, comm=comm)
    g = torch.arange(1000, dtype=torch.float32, device=dev) * (rank + 1)
    scale = comm.allreduce_grads(g)
    both = comm.allgather(torch.full((3,), float(rank), device=dev))
    torch.cuda.synchronize()
    np.savez(os.path.join(out_dir, "r%d.npz" % rank), vals=vals.cpu().numpy(), idx=idx.cpu().numpy(), g=g.cpu().numpy(),
             scale=scale, both=both.cpu().numpy())
    comm.close()


@pytest.mark.parametrize("n,k", [(200001, 320), (300, 320), (7, 4)])
def test_merge_world2_nccl_c_abi(tmp_path, n, k):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    world = 2
    mp.spawn(_worker, args=(world, str(tmp_path / "nccl_id"), n, k, str(tmp_path)), nprocs=world, join=True)
    want_v, want_i = O.topk_desc(_scores(n), k)
    for r in range(world):
        z = np.load(tmp_path / ("r%d.npz" % r))
        assert np.array_equal(z["idx"], want_i) and np.array_equal(z["vals"], want_v)
        assert np.array_equal(z["g"], np.arange(1000, dtype=np.float32) * 3) and float(z["scale"]) == 0.5
        assert np.array_equal(z["both"], np.array([[0, 0, 0], [1, 1, 1]], np.float32))
