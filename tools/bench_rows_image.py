"""A/B of the row encoder fed from fp32 rows (loader warps) and from the packed row image (TMA), alternating
(the power cap makes back-to-back legs incomparable: profiles/r01_ncu_segsum_tc_v8_notes.md).
PYTHONPATH=. python tools/bench_rows_image.py [programs]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import vaex_b200
from vaex_b200 import _lib


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    dev = torch.device("cuda", 0)
    feats, sizes, norm, bad = bench.gen_workload(P, 1000, dev)
    rows = feats.shape[0]
    torch.manual_seed(0)
    net = vaex_b200.SegmentSumMLPModule(164, 256, 1).to(dev).eval()
    model = net.packed(norm.cpu())
    lib = _lib.load()
    st = torch.cuda.current_stream().cuda_stream
    offsets = torch.empty(P + 1, dtype=torch.int32, device=dev)
    _lib.check(lib.vaex_segment_offsets(_lib.ptr(torch.from_numpy(sizes).to(dev)), P, _lib.ptr(offsets), st))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    img = vaex_b200.RowsImage(feats)
    e1.record()
    torch.cuda.synchronize()
    pack_ms = e0.elapsed_time(e1)
    seg = torch.empty(P, 256, device=dev)
    legs = {"fp32": (feats, 0), "image": (img.data, _lib.F_FEATS_IMAGE)}
    out = {"programs": P, "rows": rows, "pack_ms_first_call": pack_ms, "rounds": []}
    ref = None
    for rnd in range(4):
        rec = {}
        for name, (buf, fl) in legs.items():
            for _ in range(2):
                _lib.check(lib.vaex_segsum_encode(model.h, _lib.ptr(buf), _lib.ptr(offsets), P, rows, _lib.ptr(seg), fl, st))
            torch.cuda.synchronize()
            e0.record()
            for _ in range(5):
                _lib.check(lib.vaex_segsum_encode(model.h, _lib.ptr(buf), _lib.ptr(offsets), P, rows, _lib.ptr(seg), fl, st))
            e1.record()
            torch.cuda.synchronize()
            rec[name] = e0.elapsed_time(e1) / 5
            if rnd == 0:
                if ref is None:
                    ref = seg.clone()
                else:
                    rec["bit_identical"] = bool(torch.equal(ref.view(torch.int32), seg.view(torch.int32)))
        out["rounds"].append(rec)
    e0.record()
    for _ in range(3):
        img = vaex_b200.RowsImage(feats)
    e1.record()
    torch.cuda.synchronize()
    out["pack_ms"] = e0.elapsed_time(e1) / 3
    flop = rows * bench.FLOP_PER_ROW
    out["tflops"] = {k: flop / (min(r[k] for r in out["rounds"]) * 1e-3) / 1e12 for k in legs}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
