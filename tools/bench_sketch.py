"""Times vaex_sketch_features (n4) and the device-resident pipeline it closes:
   z -> decode + snap + divisor legalise (D = 23 tile extents of the conv2d sketch) -> 3 feature rows per candidate ->
   SegmentSum-MLP -> top-k.   python tools/bench_sketch.py [--n 1048576]"""
import argparse
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import vaex_b200                                        # noqa: E402
from vaex_b200 import _lib, sketch_features as SF       # noqa: E402


def timed(fn, it=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(it):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / it


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1 << 20)
    args = ap.parse_args()
    n = args.n
    lib = _lib.load()
    dev = torch.device("cuda")
    t = SF.conv2d_nhwc_cpu_template(derive_outer=True)
    sk = SF.SketchFeatures(t)
    slot_ext, slot_last = SF.conv2d_slot_tables()
    torch.manual_seed(2023)
    vae = vaex_b200.VAE_feature_head(23, None, 64, 256).eval().to(dev)
    rng = np.random.default_rng(1)
    mean = torch.from_numpy(rng.uniform(0.5, 3, 23).astype(np.float32)).to(dev)
    scale = torch.from_numpy(rng.uniform(0.3, 1.5, 23).astype(np.float32)).to(dev)
    z = torch.randn(n, 64, device=dev)
    ext = torch.empty(n, 23, dtype=torch.int32, device=dev)
    se, sl_ = torch.from_numpy(slot_ext).to(dev), torch.from_numpy(slot_last).to(dev)
    st = _lib.stream_ptr()
    vm = vae.packed()

    def decode():
        _lib.check(lib.vaex_decode_snap(vm.h, _lib.ptr(z), n, None, _lib.ptr(ext), _lib.ptr(mean), _lib.ptr(scale), _lib.SNAP_EXPM1 | _lib.SNAP_DIVISOR,
                                        _lib.ptr(se), _lib.ptr(sl_), 64, st))
    decode()
    e = ext.cpu().numpy()
    rows = torch.empty(n * sk.n_rows, 164, device=dev)
    ms_dec = timed(decode)
    ms_rows = timed(lambda: sk.rows(ext, out=rows))
    rows16 = torch.empty(n * sk.n_rows, 164, device=dev, dtype=torch.float16)
    ms_rows16 = timed(lambda: sk.rows(ext, out=rows16))
    # scoring: the SegmentSum-MLP on the generated rows
    out = {"n": n, "rows_per_candidate": sk.n_rows, "program_words": int(sk.program.numel()), "ms_decode_snap_legalise": ms_dec,
           "ms_sketch_rows": ms_rows, "ms_sketch_rows_f16": ms_rows16, "candidates_per_s_rows": n / (ms_rows * 1e-3), "rows_GBps": n * sk.n_rows * 656 / (ms_rows * 1e-3) / 1e9,
           "extent_sample": e[:3].tolist(), "distinct_candidates_in_first_4096": int(len({tuple(r) for r in e[:4096].tolist()}))}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
