"""Extent parser + scaler + the way back to measure records.

Goldens (tests/golden/extent_text.npz, oracle/gen_golden.py::gen_extent_text): the reference's own
state_to_records (utils/extent.py:94-136) on printed states, sklearn's StandardScaler (what the reference calls,
vae_extent_search.py:394-397), and the program printed in the reference README (:86-108) as a known answer.
The parser and the record writer are host code (no GPU needed); the scaler and the legaliser are CUDA kernels."""
import json

import numpy as np
import pytest

from conftest import load_golden
from oracle import extent_oracle as XO

import vaex_b200
from vaex_b200 import extent


def _strs(g, tag):
    raw, off = g[tag + "_text"].tobytes(), g[tag + "_offsets"]
    return [raw[off[i]:off[i + 1]].decode() for i in range(len(off) - 1)]


def _ragged(flat, lens):
    return np.split(flat, np.cumsum(lens)[:-1])


# ------------------------------------------------------------------ oracle pins (CPU)
@pytest.mark.parametrize("tag", ["rag", "pool"])
def test_oracle_parser_matches_reference(tag):
    g = load_golden("extent_text.npz")
    rows = XO.state_to_extents(_strs(g, tag))
    assert np.array_equal(np.array([len(r) for r in rows], np.int32), g[tag + "_len"])
    assert np.array_equal(np.concatenate(rows), g[tag + "_all"])


def test_oracle_scaler_matches_sklearn():
    g = load_golden("extent_text.npz")
    mean, var, scale = XO.scaler_fit(g["scaler_raw"])
    np.testing.assert_allclose(mean, g["scaler_mean"], rtol=1e-13)
    np.testing.assert_allclose(var, g["scaler_var"], rtol=1e-11, atol=1e-18)
    np.testing.assert_allclose(scale, g["scaler_scale"], rtol=1e-11)
    assert scale[-1] == 1.0 and scale[2] == 1.0                      # constant columns
    assert np.array_equal(XO.scaler_transform(g["scaler_raw"], g["scaler_mean"], g["scaler_scale"]), g["scaler_out"])
    back = XO.scaler_inverse(g["scaler_out"], g["scaler_mean"], g["scaler_scale"])
    assert np.array_equal(back, np.maximum(g["scaler_raw"], 1).astype(np.int64))


# ------------------------------------------------------------------ host parser through the C-ABI (CPU)
@pytest.mark.parametrize("tag", ["rag", "pool"])
def test_parser_matches_reference(tag):
    g = load_golden("extent_text.npz")
    strs = _strs(g, tag)
    mat, row_len, counts = extent.parse_extents(strs)
    assert mat.dtype == np.float32 and mat.shape == (len(strs), int(g[tag + "_len"].max()))
    assert np.array_equal(row_len, g[tag + "_len"]) and np.array_equal(counts[:, 0], g[tag + "_n_ext"])
    for i, want in enumerate(_ragged(g[tag + "_all"], g[tag + "_len"])):
        assert np.array_equal(mat[i, :row_len[i]], want), i
        assert not mat[i, row_len[i]:].any()


def test_parser_readme_program_known_answer():
    g = load_golden("extent_text.npz")
    readme = _strs(g, "rag")[400]
    assert "Conv2dOutput auto_unroll: 16" in readme
    mat, row_len, _ = extent.parse_extents([readme])
    assert row_len[0] == 15 and np.array_equal(mat[0], g["readme_extents"])      # D = 15 (README.md:86-108)


def test_parser_edge_cases():
    mat, row_len, counts = extent.parse_extents([])
    assert mat.shape == (0, 1) and len(row_len) == 0
    cases = ["", "Placeholder", "Placeholder: a\n", "(0,5)", "xx(0,5)", "Placeholder: a\nfor i (0,5",
             "Placeholder: a\nfor i (0,)\nfor j (0,(0,6))", "Placeholder x Placeholder: b\n(0,9) auto_unroll:\n 12 auto_unroll:x",
             "auto_unroll: 4 auto_unroll:5auto_unroll:", "pp(00,3) (0,0) (0,007)", "Placeholder: a\nfor i (0,\x1c8)\n"]
    mat, row_len, counts = extent.parse_extents(cases)
    for i, want in enumerate(XO.state_to_extents(cases)):
        assert np.array_equal(mat[i, :row_len[i]], want), (cases[i], mat[i], want)
    with pytest.raises(vaex_b200._lib.VaexError):
        extent.parse_extents(["Placeholder: a\n(0,1) (0,2) (0,3)"], d=2)


def test_state_to_records_matches_reference():
    g = load_golden("extent_text.npz")
    for tag in ("rag", "pool"):
        strs = _strs(g, tag)
        rec = extent.state_to_records(strs)
        assert set(rec) >= {"schedules", "extents", "unroll", "all", "cleaned_schedules"}
        assert [len(a) for a in rec["all"]] == list(g[tag + "_len"])
        assert np.array_equal(np.concatenate(rec["all"]), g[tag + "_all"])
        assert all(isinstance(v, float) for v in rec["extents"][0] + rec["unroll"][0])
        assert rec["cleaned_schedules"] == _strs(g, tag + "_clean")
    assert rec["matrix"].shape == (3000, 15)
    with pytest.raises(vaex_b200._lib.VaexError):
        extent.gen_program_pool(None, 10, 10, 10)


# ------------------------------------------------------------------ extents -> measure records (host)
def _template():
    steps = [["CHW", 2, "local"], ["SP", 2, 0, 56, [1, 2, 7], 1], ["SP", 2, 4, 64, [1, 4, 8], 1], ["SP", 2, 8, 3, [1], 1],
             ["RE", 2, [0, 4, 1, 5, 2, 6, 8, 9, 3, 7]], ["FSP", 3, 0, 1, 2], ["AN", 3, 0, 3],
             ["PR", 2, 0, "auto_unroll_max_step$512"], ["AN", 2, 9, 2]]
    return {"i": [["[\"conv2d\", 1, 56, 56, 64]", "llvm -keys=cpu -link-params=0 -mcpu=core-avx2"], [[], steps]],
            "r": [[0.00123], 0, 1.5, 1620000000], "v": "v0.6"}


# printed order interleaves the two big splits: i.1 j.1 i.2 j.2 (k outer: derived) k.1 i.3 j.3 unroll
_SLOTS = [None, ("SP", 1, 0), ("SP", 2, 0), ("SP", 1, 1), ("SP", 2, 1), None, ("SP", 3, 0), ("SP", 1, 2), ("SP", 2, 2), ("PR", 7)]


def test_records_from_extents_round_trip():
    sk = extent.SketchSlots(json.dumps(_template()), _SLOTS)
    ext = np.array([[9, 2, 4, 7, 2, 3, 1, 4, 8, 64], [1, 1, 1, 1, 1, 1, 3, 56, 64, 0]], dtype=np.int32)
    lines = sk.records_from_extents(ext)
    for row, ln in zip(ext, lines):
        rec = json.loads(ln)
        assert set(rec) == {"i", "r", "v"} and rec["i"][0] == _template()["i"][0] and rec["v"] == "v0.6"
        st = rec["i"][1][1]
        assert st[1][:4] == ["SP", 2, 0, 56] and st[1][4] == [int(row[1]), int(row[3]), int(row[7])] and st[1][5] == 1
        assert st[2][4] == [int(row[2]), int(row[4]), int(row[8])] and st[3][4] == [int(row[6])]
        assert st[7][3] == "auto_unroll_max_step$%d" % row[9]
        assert st[0] == ["CHW", 2, "local"] and st[4] == _template()["i"][1][1][4]        # untouched steps
        got = sk.extents_from_record(ln)
        mask = np.array([s is not None for s in _SLOTS])
        assert np.array_equal(got[mask], row[mask]) and not got[~mask].any()
    bad = ext.copy()
    bad[0, 1] = 5                                                    # 5*7*4 does not divide 56
    with pytest.raises(ValueError):
        sk.records_from_extents(bad)
    assert sk.records_from_extents(bad, strict=False)[0] is None
    se, sl, so = sk.legalise_tables()
    assert list(so) == [1, 3, 7, 2, 4, 8, 6] and list(se[[1, 2, 6, 0]]) == [56, 64, 3, 0] and list(np.nonzero(sl)[0]) == [6, 7, 8]
    with pytest.raises(ValueError):
        extent.SketchSlots(_template(), [("SP", 0, 0)])
    with pytest.raises(ValueError):
        extent.SketchSlots(_template(), [("PR", 8)])


# ------------------------------------------------------------------ GPU: scaler, legaliser
@pytest.mark.gpu
def test_scaler_fit_transform_matches_sklearn():
    import torch
    g = load_golden("extent_text.npz")
    raw = g["scaler_raw"]
    sc = extent.ExtentScaler()
    out = sc.fit_transform(raw)
    np.testing.assert_allclose(sc.mean_, g["scaler_mean"], rtol=1e-6)     # CUDA log1pf vs numpy's: <= 1 ulp of fp32 per value
    np.testing.assert_allclose(sc.scale_, g["scaler_scale"], rtol=1e-6)
    np.testing.assert_allclose(sc.var_, g["scaler_var"], rtol=2e-6, atol=1e-12)
    assert sc.scale_[-1] == 1.0 and sc.scale_[2] == 1.0 and sc.n_samples_seen_ == raw.shape[0]
    np.testing.assert_allclose(out.cpu().numpy(), g["scaler_out"], rtol=0, atol=2e-6)
    # with the reference's statistics the transform is sklearn's arithmetic step for step
    sc2 = extent.ExtentScaler()
    sc2.mean_, sc2.scale_ = g["scaler_mean"], g["scaler_scale"]
    got = sc2.transform(torch.from_numpy(raw).cuda()).cpu().numpy()
    assert np.abs(got - g["scaler_out"]).max() <= 2e-6 and (got != g["scaler_out"]).mean() < 0.02
    # same data twice -> bit-identical statistics (fixed reduction order)
    sc3 = extent.ExtentScaler().fit(raw)
    assert np.array_equal(sc3.mean_, sc.mean_) and np.array_equal(sc3.var_, sc.var_)
    # inverse: the snap kernel gives the raw integers back (clamped at 1)
    back = sc.inverse_transform(out).cpu().numpy()
    assert np.array_equal(back, np.maximum(raw, 1).astype(np.int32))


@pytest.mark.gpu
@pytest.mark.parametrize("n,d", [(1, 1), (7, 256), (100003, 48), (5, 45)])
def test_scaler_sizes(n, d):
    rng = np.random.default_rng(n + d)
    raw = rng.integers(0, 2000, size=(n, d)).astype(np.float32)
    sc = extent.ExtentScaler().fit(raw)
    mean, var, scale = XO.scaler_fit(raw)
    np.testing.assert_allclose(sc.mean_, mean, rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(sc.scale_, scale, rtol=1e-5)
    got = sc.transform(raw).cpu().numpy()
    np.testing.assert_allclose(got, XO.scaler_transform(raw, sc.mean_, sc.scale_), atol=3e-6 * max(1.0, float(1 / scale.min())))


@pytest.mark.gpu
def test_legalise_with_slot_order_and_records():
    import torch
    sk = extent.SketchSlots(_template(), _SLOTS)
    rng = np.random.default_rng(5)
    ext = torch.from_numpy(rng.integers(1, 70, size=(4096, 10)).astype(np.int32)).cuda()
    before = ext.clone()
    sk.legalise(ext, max_innermost=0)
    got = ext.cpu().numpy()
    assert np.array_equal(got[:, [0, 5, 9]], before.cpu().numpy()[:, [0, 5, 9]])        # unlisted slots untouched
    for cols, E in (([1, 3, 7], 56), ([2, 4, 8], 64), ([6], 3)):
        assert (E % np.prod(got[:, cols], axis=1) == 0).all()
    # host restatement of the walk (nearest divisor of what is left, ties to the smaller)
    def nearest(rem, v):
        v = min(max(v, 1), rem)
        for dlt in range(rem + 1):
            if v - dlt >= 1 and rem % (v - dlt) == 0:
                return v - dlt
            if v + dlt <= rem and rem % (v + dlt) == 0:
                return v + dlt
    b = before.cpu().numpy()
    for i in range(0, 4096, 37):
        for cols, E in (([1, 3, 7], 56), ([2, 4, 8], 64), ([6], 3)):
            rem = E
            for c in cols:
                q = nearest(rem, int(b[i, c]))
                assert got[i, c] == q
                rem //= q
    lines = sk.records_from_extents(got[:64])                      # strict: every row is a legal record
    assert len(lines) == 64 and all(json.loads(ln)["i"][1][1][1][3] == 56 for ln in lines)
    # already legal rows are a fixed point
    again = ext.clone()
    sk.legalise(again)
    assert torch.equal(again, ext)
