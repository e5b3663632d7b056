"""End-to-end throughput of vaex_segsum_mlp_predict_blob (TVM's packed per-store feature bytes in host memory -> scores
in host memory) with the narrowing feeders and with fp32 staging (VAEX_F_NO_HOST_PACK).
Usage on the GPU box: PYTHONPATH=. python tools/bench_blob.py [records]"""
import json
import sys
import time

import numpy as np
import torch

import vaex_b200
from vaex_b200 import _lib

n = int(sys.argv[1]) if len(sys.argv) > 1 else 400_000
F = 164
rng = np.random.default_rng(0)
nst = rng.integers(1, 33, size=n).astype(np.int64)
rec = 1 + nst * F                                   # floats per record: n_stmts, then the rows
start = np.concatenate([[0], np.cumsum(rec)])
body = rng.random(int(start[-1]), dtype=np.float32) * 24
body[start[:-1]] = nst.astype(np.float32)
head = np.concatenate([[n], rec, [n, n, 1]]).astype(np.int32)
blob = head.tobytes() + body.tobytes() + np.zeros(n, np.float32).tobytes() + np.zeros(n, np.int32).tobytes() + np.zeros(1, np.float32).tobytes()
del body
rows = int(nst.sum())
torch.manual_seed(0)
net = vaex_b200.SegmentSumMLPModule(F, 256, 1).cuda().eval()
m = net.packed(None)
lib = _lib.load()
buf = np.frombuffer(blob, dtype=np.uint8)
out = np.empty(n, np.float32)


def run(flags):
    _lib.check(lib.vaex_segsum_mlp_predict_blob(m.h, buf.ctypes.data, buf.nbytes, out.ctypes.data, n, flags))


res = {"records": n, "rows": rows, "blob_bytes": buf.nbytes}
ref = None
for name, fl in (("fp32_staging", _lib.F_NO_HOST_PACK), ("narrowing_feeders", 0)):
    run(fl)
    c0 = (m.counter(0), m.counter(1))
    ts = []
    for _ in range(4):
        t0 = time.perf_counter()
        run(fl)
        ts.append(time.perf_counter() - t0)
    res[name] = {"ms": round(min(ts) * 1e3, 2), "records_per_s": n / min(ts), "blob_GB_per_s": buf.nbytes / min(ts) / 1e9,
                 "chunks_fp16": (m.counter(0) - c0[0]) // 4, "chunks_fp32": (m.counter(1) - c0[1]) // 4}
    if ref is None:
        ref = out.copy()
    else:
        res["bit_identical"] = bool(np.array_equal(ref.view(np.uint32), out.view(np.uint32)))
print(json.dumps(res))
