"""config 3 on one GPU: both ascent graphs, 256 restarts x 200 steps.  PYTHONPATH=. python tools/bench_ascent.py"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import vaex_b200


def timed(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    dev = torch.device("cuda", 0)
    vae, reg, *_ = bench.extent_models(dev)
    torch.manual_seed(0)
    net = vaex_b200.SegmentSumMLPModule(164, 256, 1).to(dev).eval()
    out = {}
    for R in (32, 128, 256):
        z0 = torch.randn(R, 64, device=dev)
        out["predictor_R%d_us_per_step" % R] = timed(lambda: reg.latent_ascent(z0, steps=200, lr=1e-2)) * 1e3 / 200
        out["decoder_head_R%d_us_per_step" % R] = timed(lambda: vae.latent_ascent_through(net, z0, steps=200, lr=1e-2)) * 1e3 / 200
    print(json.dumps(out))


if __name__ == "__main__":
    main()
