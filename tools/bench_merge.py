"""Cost of the candidate merge (vaex_topk_allgather_merge) and of the bare all-gather on N GPUs.
torchrun --nproc-per-node N tools/bench_merge.py"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vaex_b200
from vaex_b200 import _lib, shard


def timed(fn, iters=50, warm=5):
    for _ in range(warm):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    comm = shard.Comm()
    lib = _lib.load()
    P, K = 1_000_000, 320
    scores = torch.randn(P, device="cuda")
    tv, ti = torch.empty(K, device="cuda"), torch.empty(K, dtype=torch.int64, device="cuda")
    n = torch.zeros(1, dtype=torch.int32, device="cuda")
    mv, mi, mn = torch.empty(K, device="cuda"), torch.empty(K, dtype=torch.int64, device="cuda"), torch.zeros(1, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    topk = lambda: _lib.check(lib.vaex_topk(_lib.ptr(scores), P, K, None, _lib.ptr(tv), _lib.ptr(ti), _lib.ptr(n), st))
    merge = lambda: _lib.check(lib.vaex_topk_allgather_merge(comm.h, _lib.ptr(tv), _lib.ptr(ti), _lib.ptr(n), K, rank * P, _lib.ptr(mv), _lib.ptr(mi), _lib.ptr(mn), st))
    small = torch.zeros(1280, device="cuda")
    out = torch.empty(world * 1280, device="cuda")
    res = {"world": world,
           "topk_us": timed(topk), "merge_us": timed(merge), "topk+merge_us": timed(lambda: (topk(), merge())),
           "vaex_allgather_5KB_us": timed(lambda: _lib.check(lib.vaex_allgather(comm.h, _lib.ptr(small), _lib.ptr(out), 5120, st))),
           "torch_allgather_5KB_us": timed(lambda: dist.all_gather_into_tensor(out, small)),
           "python_merge_us": timed(lambda: shard.merge_topk(tv, ti, rank * P, K, n_local=n, sync=False))}
    if rank == 0:
        print(json.dumps(res))
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
