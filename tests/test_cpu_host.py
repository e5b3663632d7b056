"""CPU-only checks: the C-ABI library loads and exports every symbol the header declares,
host-side logic matches the oracle, and the product fails loudly without a GPU."""
import os
import pickle
import re

import numpy as np
import pytest
import torch

from conftest import ROOT, load_golden
from oracle import torch_oracle as O

import vaex_b200
from vaex_b200 import _lib


def header_symbols():
    src = open(os.path.join(ROOT, "include", "vaex_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vaex_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    syms = header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), "missing export " + s
        assert s in _lib.SIGNATURES, "no ctypes signature for " + s
    assert lib.vaex_abi_version() == 1


def test_no_silent_cpu_fallback():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = _lib.load()
    assert lib.vaex_device_count() < 0 or lib.vaex_device_count() == 0
    net = vaex_b200.SegmentSumMLPModule(164, 256, 1).eval()
    with pytest.raises(_lib.VaexError):
        net(torch.ones(2, dtype=torch.int32), torch.zeros(2, 164))
    with pytest.raises(_lib.VaexError):
        net.packed()  # packing needs the device
    with pytest.raises(_lib.VaexError):
        vaex_b200.MLPModelInternal(device="cpu")


def test_model_create_argument_errors():
    import ctypes
    lib = _lib.load()
    h = ctypes.c_void_p()
    dims = (ctypes.c_int32 * 2)(164, 256)
    assert lib.vaex_model_create(_lib.KIND_SEGSUM_MLP, dims, 2, ctypes.byref(h)) == -1
    assert b"dims" in lib.vaex_last_error()
    m = _lib.Model(_lib.KIND_SEGSUM_MLP, [164, 256, 0, 0])
    w = np.zeros((3, 3), np.float32)
    assert lib.vaex_model_load_weights(m.h, b"x", w.ctypes.data, 0, 3) == -1
    m.close()


def test_flatten_matches_oracle_and_reference_loader():
    g = load_golden("segsum_mlp_f174.npz")
    offs = np.concatenate([[0], np.cumsum(g["sizes"])])
    progs = np.empty(len(g["sizes"]), dtype=object)
    for i in range(len(progs)):
        progs[i] = g["raw"][offs[i]:offs[i + 1]]
    emb = np.zeros(10, np.float32)
    flat, sizes = vaex_b200.flatten_features(progs, emb)
    oflat, osizes = O.flatten_programs(list(progs), emb)
    assert np.array_equal(flat, oflat) and np.array_equal(sizes, osizes)
    ds = vaex_b200.Dataset.create_one_task(vaex_b200.LearningTask("wk", "cuda"), progs, np.zeros(len(progs)))
    loader = vaex_b200.SegmentDataLoader(ds, 40, "cpu", True, False, {}, None)
    norm = loader.normalize()
    assert np.array_equal(norm.numpy(), g["norm"])
    assert np.array_equal(loader.features[:64].numpy(), g["flat_norm_head"])
    assert np.array_equal(loader.segment_sizes.numpy(), g["sizes"])
    batches = list(loader)
    assert sum(len(b[0]) for b in batches) == len(progs)
    assert sum(b[1].shape[0] for b in batches) == flat.shape[0]


def test_dataset_container():
    t = vaex_b200.LearningTask("k", "cuda")
    f = np.empty(2, dtype=object)
    f[0], f[1] = np.zeros((2, 164), np.float32), np.ones((3, 164), np.float32)
    ds = vaex_b200.Dataset.create_one_task(t, f, np.array([0.5, 1.0]), min_latency=1.0)
    ds.load_task_data(t, f, np.array([0.25, 1.0]), min_latency=0.5)
    assert len(ds) == 4 and ds.tasks() == [t] and ds.min_latency[t] == 0.5
    np.testing.assert_allclose(ds.throughputs[t], [0.25, 0.5, 0.25, 1.0])
    ds2 = pickle.loads(pickle.dumps(ds))
    assert len(ds2) == 4


def test_state_dict_keys_match_reference_goldens():
    g = load_golden("segsum_mlp_f164.npz")
    ref_keys = sorted(k[2:] for k in g if k.startswith("w."))
    assert sorted(vaex_b200.SegmentSumMLPModule(164, 256, 1).state_dict().keys()) == ref_keys
    g = load_golden("extent_vae.npz")
    assert sorted(vaex_b200.VAE_feature_head(48, None, 64, 256).state_dict().keys()) == \
        sorted(k[4:] for k in g if k.startswith("vae."))
    assert sorted(vaex_b200.VAECostPredictor(48, hidden_dim=256, latent_dim=64).state_dict().keys()) == \
        sorted(k[4:] for k in g if k.startswith("reg."))


def test_reference_pickle_format_roundtrip(tmp_path):
    m = vaex_b200.MLPModelInternal.__new__(vaex_b200.MLPModelInternal)
    m.base_model = vaex_b200.SegmentSumMLPModule(164, 256, 1)
    m.local_model, m.few_shot_learning, m.fea_norm_vec = {}, "base_only", torch.ones(164)
    p = str(tmp_path / "m.pkl")
    m.save(p)
    base, local, fsl, norm = pickle.load(open(p, "rb"))
    assert fsl == "base_only" and isinstance(base, vaex_b200.SegmentSumMLPModule)
    m2 = vaex_b200.MLPModelInternal.__new__(vaex_b200.MLPModelInternal)
    m2.load(p)
    for k, v in m.base_model.state_dict().items():
        assert torch.equal(v, m2.base_model.state_dict()[k])


def test_select_budget_arithmetic_cpu():
    # the integer budget split of select_programs (select.py:309-315) is host logic
    for total, g, r, w in [(64, 2, 0, (0.5, 0.3, 0.2)), (48, 2, 0, (0.4, 0.3, 0.3)), (64, 4, 2, (1.0, 0.0, 0.0))]:
        budget = total - g - r
        c, u, d = int(budget * w[0]), int(budget * w[1]), int(budget * w[2])
        if c + u + d < budget:
            c += budget - (c + u + d)
        assert c + u + d == budget


def test_pack_features_f16_is_the_kernels_rounding():
    """vaex_pack_features_f16 (host, AVX2 + F16C) == clamp to +-65504 then round-to-nearest-even, the function
    cvt.rn.satfinite.f16.f32 computes in the row encoder: bitwise against numpy's float16 cast."""
    from vaex_b200.mlp_model import pack_features_f16
    rng = np.random.default_rng(11)
    x = np.concatenate([
        rng.standard_normal(100003).astype(np.float32) * 30,
        np.float32([0.0, -0.0, 65504, 65519.9, 65520, 1e6, -1e6, np.inf, -np.inf, 6.1e-5, 5.96e-8, 3.1e-8, 1e-7, -2.5e-7]),
        (rng.standard_normal(5000) * 1e-6).astype(np.float32),                     # fp16 subnormal range
        np.float32(2048) + np.arange(64, dtype=np.float32),                        # ties
    ])
    for shift in (0, 1, 3, 7):                                                    # unaligned heads and tails
        got, lossy = pack_features_f16(x[shift:])
        want = np.clip(x[shift:], -65504, 65504).astype(np.float16)
        assert got.dtype == np.float16 and np.array_equal(got.view(np.uint16), want.view(np.uint16))
        assert lossy == bool(((x[shift:] != 0) & (want == 0)).any())
    assert not pack_features_f16(x[:100017])[1] and pack_features_f16(x[100017:])[1]
    got, lossy = pack_features_f16(np.float32([[1.0, 2.9e-8, 3.0]]))
    assert lossy and got[0, 1] == 0                                               # 2^-25 = 2.98e-8 and below round to 0
    got, lossy = pack_features_f16(np.float32([np.nan, 1.0]))
    assert np.isnan(got[0]) and got[1] == 1.0
    assert pack_features_f16(np.zeros((0, 164), np.float32))[0].shape == (0, 164)
