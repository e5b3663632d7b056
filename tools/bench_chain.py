"""Extent-VAE scoring chain (BASELINE configs 3/5): latents/s of decode(+snap) and predict_cost on one GPU,
tcgen05 chain vs fp32 SIMT.  Run on the GPU box: PYTHONPATH=. python tools/bench_chain.py [n]"""
import json
import sys

import numpy as np
import torch

import vaex_b200


def timed(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
    torch.manual_seed(0)
    vae = vaex_b200.VAE_feature_head(48, None, 64, 256).cuda().eval()
    reg = vaex_b200.VAECostPredictor(48, hidden_dim=256, latent_dim=64).cuda().eval()
    z = torch.randn(n, 64, device="cuda")
    xin = torch.randn(n, 48, device="cuda")
    mean = np.random.default_rng(0).uniform(0.5, 3, 48).astype(np.float32)
    std = np.random.default_rng(1).uniform(0.3, 1.5, 48).astype(np.float32)
    md, sd = torch.from_numpy(mean).cuda(), torch.from_numpy(std).cuda()
    out = {"n": n}
    for tc in (True, False):
        vae.set_tensor_cores(tc), reg.set_tensor_cores(tc)
        tag = "tc" if tc else "simt"
        iters = 20 if tc else 3
        out["encode_mean_ms_" + tag] = timed(lambda: vae.encode(xin, use_mean=True), iters)
        out["encode_ms_" + tag] = timed(lambda: vae.encode(xin), iters)
        out["decode_ms_" + tag] = timed(lambda: vae.decode(z), iters)
        out["decode_snap_ms_" + tag] = timed(lambda: vae.decode_and_snap(z, md, sd, want_x=False), iters)
        out["predict_cost_ms_" + tag] = timed(lambda: reg.predict_cost(z), iters)
    # algorithmic flops per latent: decoder 2*(64*256 + 2*256*256 + 256*48), predictor 2*(64*256 + 256*256 + 256)
    fd, fp = 2 * (64 * 256 + 2 * 256 * 256 + 256 * 48), 2 * (64 * 256 + 256 * 256 + 256)
    out["decode_snap_latents_per_s_tc"] = n / out["decode_snap_ms_tc"] * 1e3
    out["predict_latents_per_s_tc"] = n / out["predict_cost_ms_tc"] * 1e3
    out["decode_snap_tflops_tc"] = fd * n / out["decode_snap_ms_tc"] / 1e9
    out["predict_tflops_tc"] = fp * n / out["predict_cost_ms_tc"] / 1e9
    print(json.dumps(out))


if __name__ == "__main__":
    main()
