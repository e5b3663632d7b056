"""vaex_legalise_extents alone on config 5's shape (2 M rows x 48 slots, task-1264-like loop bounds).  The kernel is
idempotent on its own output, so the timed launches repeat the same work on the same buffer.
Usage on the GPU box: PYTHONPATH=. python tools/bench_legalise.py [rows]"""
import json
import sys

import numpy as np
import torch

from vaex_b200 import _lib

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
d = 48
rng = np.random.default_rng(2023)
bounds = [1, 3, 4, 28, 56, 57, 64, 128]
ext, last = [], []
while len(ext) < 44:
    ext += [bounds[rng.integers(len(bounds))]] * 4
    last += [0, 0, 0, 1]
ext, last = ext + [0] * 4, last + [0] * 4
e_d = torch.tensor(ext, dtype=torch.int32, device="cuda")
l_d = torch.tensor(last, dtype=torch.uint8, device="cuda")
raw = torch.from_numpy(np.exp(rng.normal(1.5, 1.0, size=(n, d))).round().clip(1, 1 << 20).astype(np.int32)).cuda()
lib, st = _lib.load(), _lib.stream_ptr()


def run(t):
    _lib.check(lib.vaex_legalise_extents(_lib.ptr(t), n, d, _lib.ptr(e_d), _lib.ptr(l_d), None, 0, 64, st))


res = {"rows": n, "slots": d}
for name in ("first_pass", "repeat"):
    t = raw.clone()
    if name == "repeat":
        run(t)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    iters = 1 if name == "first_pass" else 20
    e0.record()
    for _ in range(iters):
        run(t)
    e1.record()
    torch.cuda.synchronize()
    res[name + "_ms"] = round(e0.elapsed_time(e1) / iters, 4)
res["GB_per_s"] = round(2 * n * d * 4 / res["repeat_ms"] / 1e6, 1)
res["hbm_floor_ms"] = round(2 * n * d * 4 / 6543.7e9 * 1e3, 4)
print(json.dumps(res))
