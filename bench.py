#!/usr/bin/env python
"""bench.py -- candidate schedules scored/sec on BASELINE.json configs[1]:
TenSet SegmentSumMLP inference, 1M synthetic programs x <=32 stmts x 164 features per B200.

One step = one pass of the hot path over the rank's batch: fused row-encoder + segment-sum
(tcgen05) -> residual head -> scores, then the deterministic top-k (k=320) and, at N>1, the
all-gather merge of every rank's k winners (the only collective on the path).
  value : whole-job programs/s, inputs resident in HBM (10.8 GB/rank > L2, no flush needed)
  e2e   : same metric through vaex_segsum_mlp_predict_host with pinned HOST buffers
          (H2D of the features and D2H of the scores inside the timed region)
  --impl reference : the reference's CPU implementation of the path (oracle port of
          MLPModelInternal.predict, torch CPU fp32, all host threads) on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "candidate schedules scored/sec (decode+legalise+cost model) at 1/2/4/8 B200"
UNIT = "programs/s"
F, H, SMAX, TOPK = 164, 256, 32, 320
FLOP_PER_ROW = 2 * (F * H + H * H)            # 215 040  
FLOP_PER_PROGRAM_HEAD = 2 * (2 * H * H + H)   # 262 656
WORKLOAD = "TenSet SegmentSumMLP inference, 1M synthetic programs x <=32 stmts x 164 features, 1 B200"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="vaex", choices=["vaex", "reference"])
    ap.add_argument("--programs", type=int, default=1_000_000, help="programs per GPU")
    ap.add_argument("--e2e-programs", type=int, default=0, help="0 = same as --programs if host RAM allows")
    ap.add_argument("--cpu-sample", type=int, default=30000)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


# ---------------------------------------------------------------------------- synthetic input
COL_SCALE = None


def col_scale():
    """Value ranges per column, after src/auto_scheduler/feature.cc:1063-1187:
    slog columns ~U(0,24), one-hot / flag columns in {0,1}, AI curve raw ~U(0,8)."""
    global COL_SCALE
    if COL_SCALE is None:
        s = np.full(F, 24.0, np.float32)
        onehot = list(range(19, 27)) + list(range(30, 38)) + list(range(41, 49)) + [49]
        for b in range(5):
            base = 57 + 18 * b
            onehot += [base, base + 1, base + 2, base + 7, base + 8, base + 9]
        s[onehot] = -1.0                      # marker: binary column
        s[147:157] = 8.0
        COL_SCALE = s
    return COL_SCALE


def gen_workload(n_programs, seed, device):
    """Seeded synthetic config-2 input generated on the device: sizes ~U{1..32}; 15 % of
    columns zero per program block, 0.1 % of programs all-zero (invalid -> -inf)."""
    import torch
    rng = np.random.default_rng(seed)
    sizes = rng.integers(1, SMAX + 1, size=n_programs).astype(np.int32)
    rows = int(sizes.sum())
    g = torch.Generator(device=device).manual_seed(seed)
    scale = torch.from_numpy(col_scale()).to(device)
    feats = torch.empty(rows, F, device=device)
    chunk = 1 << 20
    for r0 in range(0, rows, chunk):
        r1 = min(rows, r0 + chunk)
        u = torch.rand(r1 - r0, F, device=device, generator=g)
        v = torch.where(scale < 0, (u > 0.6).float(), u * scale.abs())
        keep = torch.rand(r1 - r0, F, device=device, generator=g) > 0.15
        feats[r0:r1] = v * keep
    offs = np.concatenate([[0], np.cumsum(sizes, dtype=np.int64)])
    bad = np.nonzero(rng.random(n_programs) < 0.001)[0]
    for p in bad:
        feats[offs[p]:offs[p + 1]] = 0.0
    norm = torch.where(scale < 0, torch.ones_like(scale), scale.abs())   # per-column max
    return feats, sizes, norm, bad


class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None

    def run(self):
        """In-process NVML sampling every 5 ms (a query costs microseconds and takes no lock the launches need);
        spawning nvidia-smi instead took ~100 ms per sample and could land inside a 40 ms timed region."""
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            import pynvml as nv
            import torch
            nv.nvmlInit()
            try:
                uuid = str(torch.cuda.get_device_properties(self.index).uuid)
                handle = nv.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:
                handle = nv.nvmlDeviceGetHandleByIndex(self.index)
            bits = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(handle, nv.NVML_CLOCK_SM))
            while not self.stop_flag:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(handle, nv.NVML_CLOCK_SM)))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(handle)
                for n, b in bits.items():
                    if r & b:
                        self.reasons.add(n)
                time.sleep(0.005)
            return
        except Exception:
            pass
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                self.samples.append(float(out[0]))
                self.max_mhz = float(out[1])
                for n, v in zip(names, out[2:]):
                    if v.strip().lower().startswith("active"):
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ---------------------------------------------------------------------------- reference arm
def cpu_reference_run(n_sample, steps, warmup, seed=0):
    """Oracle port of MLPModelInternal.predict (SegmentDataLoader flatten + normalise + batched
    SegmentSumMLPModule.forward + the -inf rule), torch CPU fp32 with all host threads."""
    import torch
    from oracle import torch_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    rng = np.random.default_rng(seed)
    sizes = rng.integers(1, SMAX + 1, size=n_sample)
    scale = np.abs(col_scale())
    progs = [(rng.random((int(s), F), dtype=np.float32) * scale).astype(np.float32) for s in sizes]
    torch.manual_seed(0)
    import vaex_b200
    net = vaex_b200.SegmentSumMLPModule(F, H, 1)
    P = {k: v.detach() for k, v in net.state_dict().items()}
    norm = torch.from_numpy(scale)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        O.predict_programs(P, progs, norm, None, batch=4096)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    t = float(np.mean(times))
    return n_sample / t, t, cores


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = args.cpu_sample
    steps, warmup = max(1, min(args.steps, 3)), max(1, min(args.warmup, 1))
    v, t, cores = cpu_reference_run(n, steps, warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample_programs": n},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": "%d programs of the same workload per step; oracle port of MLPModelInternal.predict "
                                   "(python flatten + normalise + batch-4096 forward + -inf rule), torch CPU fp32" % n},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


# ---------------------------------------------------------------------------- vaex arm
def main_vaex(args):
    import torch
    import torch.distributed as dist
    import vaex_b200
    from vaex_b200 import _lib, shard

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()

    P = args.programs
    feats, sizes, norm, bad = gen_workload(P, 1000 + rank, dev)
    rows = feats.shape[0]
    torch.manual_seed(0)
    net = vaex_b200.SegmentSumMLPModule(F, H, 1).to(dev).eval()
    model = net.packed(norm.cpu())
    sizes_d = torch.from_numpy(sizes).to(dev)
    offsets = torch.empty(P + 1, dtype=torch.int32, device=dev)
    scores = torch.empty(P, device=dev)
    segsum = None
    top_v, top_i = torch.empty(TOPK, device=dev), torch.empty(TOPK, dtype=torch.int64, device=dev)
    n_out = torch.zeros(1, dtype=torch.int32, device=dev)
    merged = [None]
    stream = torch.cuda.current_stream()
    st = stream.cuda_stream
    flags = _lib.F_MASK_INVALID

    def step():
        _lib.check(lib.vaex_segment_offsets(_lib.ptr(sizes_d), P, _lib.ptr(offsets), st))
        _lib.check(lib.vaex_segsum_mlp_forward(model.h, _lib.ptr(feats), _lib.ptr(offsets), P, rows, _lib.ptr(scores), flags, st))
        _lib.check(lib.vaex_topk(_lib.ptr(scores), P, TOPK, None, _lib.ptr(top_v), _lib.ptr(top_i), _lib.ptr(n_out), st))
        if world > 1:       # candidate merge: the path's only exchange (vaex_b200/shard.py)
            merged[0] = shard.merge_topk(top_v, top_i, rank * P, TOPK, n_local=n_out, sync=False)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    torch.cuda.cudart().cudaProfilerStart()     # no-op unless `ncu --profile-from-start off` (profiles/ launch list)
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    # sanity: invalid programs are -inf, everything else finite
    sc = scores.cpu().numpy()
    assert np.isneginf(sc[bad]).all() and np.isfinite(np.delete(sc, bad)).all(), "bad scores"

    sampler = ClockSampler(local)
    sampler.start()
    l0 = lib.vaex_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    launches = lib.vaex_launch_count() - l0
    ms = e0.elapsed_time(e1) / args.steps
    torch.cuda.cudart().cudaProfilerStop()

    # dominant kernel alone (fused row encoder + segment sum), CUDA events on its stream
    segsum = torch.empty(P, H, device=dev)
    for _ in range(2):
        _lib.check(lib.vaex_segsum_encode(model.h, _lib.ptr(feats), _lib.ptr(offsets), P, rows, _lib.ptr(segsum), 0, st))
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    k0.record(stream)
    for _ in range(args.steps):
        _lib.check(lib.vaex_segsum_encode(model.h, _lib.ptr(feats), _lib.ptr(offsets), P, rows, _lib.ptr(segsum), 0, st))
    k1.record(stream)
    torch.cuda.synchronize()
    kms = k0.elapsed_time(k1) / args.steps
    del segsum
    sampler.stop_flag = True
    sampler.join(timeout=2)

    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = world * P / (ms_max * 1e-3)

    # ---- e2e through the host C-ABI call (rank-local pinned buffers) ------------------------
    e2e = None
    if not args.no_e2e:
        import psutil
        Pe = args.e2e_programs or P
        need = rows * F * 4 * Pe / P
        if psutil.virtual_memory().available < 3 * need * world:
            Pe = max(1, int(Pe * psutil.virtual_memory().available / (3 * need * world)))
        offs = np.concatenate([[0], np.cumsum(sizes[:Pe], dtype=np.int64)])
        re_ = int(offs[-1])
        host = torch.empty(re_, F, dtype=torch.float32, pin_memory=True)
        host.copy_(feats[:re_])
        out = torch.empty(Pe, dtype=torch.float32, pin_memory=True)
        hs = np.ascontiguousarray(sizes[:Pe])
        n_e2e = max(2, min(args.steps, 5))

        # several ranks share the host cores: give each its share of narrowing threads (the library default, cores - 1,
        # is for one process per host); with one core or less per rank the call simply sends more chunks as raw fp32
        e2e_flags = flags
        if world > 1 and "VAEX_PACK_THREADS" not in os.environ:
            os.environ["VAEX_PACK_THREADS"] = str(max(1, (os.cpu_count() or 1) // world - 1))
        if os.environ.get("VAEX_BENCH_NO_PACK") or (os.cpu_count() or 1) // world < 8:
            # fewer than 8 host cores per rank: the host's memory system, not one GPU's PCIe link, bounds the copy, and the
            # narrowing threads only add traffic (measured on 16 cores, aggregate programs/s raw vs narrowed: 1 rank 5.1e6 vs
            # 8.3e6, 2 ranks 1.01e7 vs 1.12e7, 4 ranks 1.94e7 vs 1.51e7, 8 ranks 1.69e7 vs 1.57e7) -> send the rows as they are
            e2e_flags = flags | _lib.F_NO_HOST_PACK

        def e2e_step():
            _lib.check(lib.vaex_segsum_mlp_predict_host(model.h, host.data_ptr(), hs.ctypes.data, Pe, re_, out.data_ptr(), e2e_flags))

        e2e_step()
        barrier()
        cnt0 = (model.counter(0), model.counter(1))
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            e2e_step()
        barrier()
        dt = (time.perf_counter() - t0) / n_e2e
        tt = torch.tensor([dt], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        fin = np.isfinite(sc[:Pe])
        assert np.array_equal(np.isfinite(out.numpy()), fin) and \
            np.abs(out.numpy()[fin] - sc[:Pe][fin]).max() < 1e-3 * max(1.0, np.abs(sc[:Pe][fin]).max()), \
            "host path disagrees with device path"
        # chunks that crossed PCIe narrowed to fp16 by the host cores vs as the caller's fp32 rows (same scores bit for bit)
        nar, raw = model.counter(0) - cnt0[0], model.counter(1) - cnt0[1]
        frac16 = nar / max(1, nar + raw)
        e2e = {"value": world * Pe / float(tt.item()), "unit": UNIT,
               "h2d_bytes_per_step": int(re_ * F * 4 * (1 - frac16) + re_ * F * 2 * frac16 + Pe * 4),
               "d2h_bytes_per_step": int(Pe * 4), "programs_per_gpu": Pe, "steps": n_e2e,
               "host_input_bytes_per_step": int(re_ * F * 4 + Pe * 4),
               "chunks_fp16_narrowed_on_host": int(nar), "chunks_fp32_raw": int(raw)}
        del host

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = peaks.get("bf16_tflops_sustained")
        peak_src = "MEASURED_PEAKS.json bf16_tflops_sustained (fp16 operands run at the bf16 rate)"
        if not peak:
            peak, peak_src = 1400.0, "fallback (B200_PROFILING.md: ~1.4 PFLOP/s sustained)"
        ach = rows * FLOP_PER_ROW / (kms * 1e-3) / 1e12
        traffic = None          # dram read+write bytes per launch of the same kernel/workload from the committed ncu capture
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json")))["segsum_tc_kernel"]
            if tr["rows"] == rows and tr["programs"] == P:
                traffic = tr["dram_bytes_read"] + tr["dram_bytes_write"]
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_max, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16xf16->f32",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "programs_per_gpu": P, "rows_per_gpu": rows, "features": F, "hidden": H,
                       "topk": TOPK, "l2": "inputs (%.1f GB/GPU) exceed L2; no flush needed" % (rows * F * 4 / 1e9),
                       "parallelism": "programs sharded over %d GPU(s); top-k all-gather only" % world},
            "e2e": e2e, "gpu_launches": int(launches), "clocks": sampler.summary(),
            "roofline": {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                         "traffic": traffic, "kernel": "segsum_tc_kernel (fused row encoder + segment sum)",
                         "kernel_ms": kms, "flop_per_launch": rows * FLOP_PER_ROW, "peak_source": peak_src,
                         "hbm_gbs_algorithmic": (rows * F * 4 + P * H * 4) / (kms * 1e-3) / 1e9},
        }
        if not args.no_cpu and world == 1:
            v, tcpu, cores = cpu_reference_run(args.cpu_sample, 1, 1)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "%d programs; oracle port of MLPModelInternal.predict, torch CPU fp32" % args.cpu_sample}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        main_reference(a)
    else:
        main_vaex(a)
