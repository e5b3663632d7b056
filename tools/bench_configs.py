"""BASELINE.json configs other than the headline one (configs[1] is bench.py), on one B200:
  config 1  4096 latents: decode -> snap + legalise -> predictor -> top-64        (+ the oracle on the host cores)
  config 3  latent ascent, 256 restarts x 200 steps through the cost predictor
  config 4  SegmentSumMLP training step (LambdaRank, Adam, clip), batch 512 from a device-resident set
  config 5  2 M latents (one GPU's shard of 16 M): decode + snap + predict + top-320
Usage on the GPU box: PYTHONPATH=. python tools/bench_configs.py > gpurun_out/configs.json
CUDA events on the current stream, warm-up first, inputs resident in HBM.  One JSON object per line."""
import json
import os
import sys
import time

import numpy as np
import torch

import vaex_b200
from vaex_b200 import train

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timed(fn, iters, warm=3):
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


def models():
    torch.manual_seed(2023)
    vae = vaex_b200.VAE_feature_head(48, None, 64, 256).cuda().eval()
    reg = vaex_b200.VAECostPredictor(48, hidden_dim=256, latent_dim=64).cuda().eval()
    rng = np.random.default_rng(2023)
    mean = rng.uniform(0.5, 3, 48).astype(np.float32)
    scale = rng.uniform(0.3, 1.5, 48).astype(np.float32)
    bounds = [1, 3, 4, 28, 56, 57, 64, 128]
    ext, last = [], []
    while len(ext) < 44:
        ext += [bounds[rng.integers(len(bounds))]] * 4
        last += [0, 0, 0, 1]
    return vae, reg, mean, scale, ext + [0] * 4, last + [0] * 4


def score_pipeline(vae, reg, z, mean_d, scale_d, ext_d, last_d, k):
    extents = vae.decode_snap(z, mean_d, scale_d, slot_extent=ext_d, slot_last=last_d)
    cost = reg.predict_cost(z)
    vals, idx, cnt = vaex_b200.topk(cost, k, sync=False)
    return extents, vals, idx


def config1_and_5(out):
    vae, reg, mean, scale, ext, last = models()
    mean_d, scale_d = torch.from_numpy(mean).cuda(), torch.from_numpy(scale).cuda()
    ext_d = torch.tensor(ext, dtype=torch.int32, device="cuda")
    last_d = torch.tensor(last, dtype=torch.uint8, device="cuda")
    for name, n, k, iters in (("config1", 4096, 64, 200), ("config5", 2_000_000, 320, 20)):
        z = torch.randn(n, 64, generator=torch.Generator().manual_seed(2023)).cuda()
        ms = timed(lambda: score_pipeline(vae, reg, z, mean_d, scale_d, ext_d, last_d, k), iters)
        rec = {"config": name, "latents": n, "topk": k, "ms": ms, "latents_per_s": n / ms * 1e3,
               "pipeline": "decode (tcgen05 chain) + snap + divisor legalise -> predict_cost (tcgen05 chain) -> top-k"}
        if name == "config1":      # the same five launches replayed from a CUDA graph: what the kernels cost without Python between them
            try:
                side = torch.cuda.Stream()
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    for _ in range(3):
                        score_pipeline(vae, reg, z, mean_d, scale_d, ext_d, last_d, k)
                torch.cuda.current_stream().wait_stream(side)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    g_ext, g_vals, g_idx = score_pipeline(vae, reg, z, mean_d, scale_d, ext_d, last_d, k)
                want = score_pipeline(vae, reg, z, mean_d, scale_d, ext_d, last_d, k)
                graph.replay()
                torch.cuda.synchronize()
                same = bool(torch.equal(g_ext, want[0]) and torch.equal(g_idx, want[2]))
                rec["cuda_graph_replay_ms"] = round(timed(graph.replay, iters), 5)
                rec["cuda_graph_same_result"] = same
            except Exception as e:      # capture is a convenience of the caller, not part of the path
                rec["cuda_graph_replay_ms"] = None
                rec["cuda_graph_error"] = str(e)[:200]
        if name == "config5":      # where the time goes (each stage timed alone, back to back launches)
            cost5 = reg.predict_cost(z)
            rec["stages_ms"] = {
                "decode+snap": round(timed(lambda: vae.decode_snap(z, mean_d, scale_d), iters), 4),
                "decode+snap+legalise": round(timed(lambda: vae.decode_snap(z, mean_d, scale_d, slot_extent=ext_d, slot_last=last_d), iters), 4),
                "predict_cost": round(timed(lambda: reg.predict_cost(z), iters), 4),
                "top-k": round(timed(lambda: vaex_b200.topk(cost5, k, sync=False), iters), 4),
            }
        if name == "config1":
            from oracle import torch_oracle as O
            Pv = {k_: v.detach().cpu() for k_, v in vae.state_dict().items()}
            Pr = {k_: v.detach().cpu() for k_, v in reg.state_dict().items()}
            zc = z.cpu().numpy()
            torch.set_num_threads(os.cpu_count())

            def cpu():
                x = O.vae_decode(Pv, zc).numpy()
                e = O.legalise_divisor(O.snap_round_clamp(x, mean, scale).numpy(), ext, last)
                c = O.predict_cost(Pr, zc).numpy()
                return e, O.topk_desc(c, 64)
            cpu()
            t0 = time.perf_counter()
            for _ in range(3):
                cpu()
            cms = (time.perf_counter() - t0) / 3 * 1e3
            rec.update({"cpu_ms": cms, "cpu_latents_per_s": n / cms * 1e3, "cpu_cores": os.cpu_count(),
                        "cpu_kind": "oracle port (torch CPU fp32 + numpy legalise)"})
        out(rec)


def config3(out):
    _, reg, _, _, _, _ = models()
    z0 = torch.randn(256, 64, generator=torch.Generator().manual_seed(3)).cuda()
    ms = timed(lambda: reg.latent_ascent(z0, steps=200, lr=1e-2), 5, warm=2)
    out({"config": "config3", "restarts": 256, "steps": 200, "ms_per_search": ms, "us_per_step": ms * 1e3 / 200,
         "restart_steps_per_s": 256 * 200 / ms * 1e3, "graph": "predictor(z) (the reference's graph, utils/select.py:128-135)"})


def config4(out):
    rng = np.random.default_rng(4)
    n = 200_000
    sizes = rng.integers(1, 33, size=n).astype(np.int32)
    rows = int(sizes.sum())
    feats = torch.rand(rows, 164, device="cuda") * 24
    labels = torch.rand(n, device="cuda")
    offs = torch.zeros(n + 1, dtype=torch.int32, device="cuda")
    offs[1:] = torch.cumsum(torch.from_numpy(sizes).cuda(), 0)
    net = vaex_b200.SegmentSumMLPModule(164, 256, 1).cuda()
    tr = train.Trainer(net, "lambdaRankLoss")
    lib = vaex_b200._lib.load()
    sizes_host = sizes.astype(np.int64)
    perm = torch.randperm(n)
    state = {"i": 0}

    def step():
        i = state["i"] % (n // 512)
        state["i"] += 1
        idx = perm[i * 512:(i + 1) * 512]
        idx_d = idx.cuda()
        r = int(sizes_host[idx.numpy()].sum())
        off = torch.empty(513, dtype=torch.int32, device="cuda")
        bf = torch.empty(r, 164, device="cuda")
        bl = torch.empty(512, device="cuda")
        vaex_b200._lib.check(lib.vaex_gather_segments(feats.data_ptr(), 164, offs.data_ptr(), idx_d.data_ptr(), 512, off.data_ptr(),
                                                      bf.data_ptr(), labels.data_ptr(), bl.data_ptr(), vaex_b200._lib.stream_ptr()))
        tr.step(off, bf, bl, 7e-4, 1e-6, 0.5)

    ms = timed(step, 100, warm=5)
    out({"config": "config4", "model": "SegmentSumMLP 164-256, LambdaRank + Adam + clip 0.5", "batch": 512, "ms_per_step": ms,
         "records_per_s": 512 / ms * 1e3, "note": "one GPU; data parallel adds one 0.96 MB NCCL all-reduce per step"})


def main():
    def out(rec):
        print(json.dumps(rec), flush=True)
    config1_and_5(out)
    config3(out)
    config4(out)


if __name__ == "__main__":
    main()
