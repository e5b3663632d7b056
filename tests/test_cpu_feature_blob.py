"""Host-side parser of Ansor's packed per-store feature blob against the reference's own
unpack_feature output (tests/golden/feature_blob.npz, produced by importing
python/tvm/auto_scheduler/feature.py) and the struct-by-struct restatement in oracle/feature_blob.py.
Host code only: runs without a GPU."""
import struct

import numpy as np
import pytest

from conftest import load_golden
from oracle import feature_blob as FB

import vaex_b200
from vaex_b200 import _lib, feature


def test_oracle_unpack_matches_reference_golden():
    g = load_golden("feature_blob.npz")
    feats, thr, ids, mins = FB.unpack_feature(g["blob"].tobytes())
    assert np.array_equal(np.array([len(f) for f in feats], np.int32), g["sizes"])
    assert np.array_equal(np.concatenate(list(feats), 0), g["flat"])
    assert np.array_equal(thr, g["thr"]) and np.array_equal(ids, g["task_ids"]) and np.array_equal(mins, g["min_costs"])


@pytest.mark.parametrize("container", [bytes, bytearray, lambda b: np.frombuffer(b, np.uint8)])
def test_parser_matches_reference_golden(container):
    g = load_golden("feature_blob.npz")
    blob = container(g["blob"].tobytes())
    assert feature.scan(blob) == (41, 608, 41, 41, 5)
    flat, sizes, lowered, thr, ids, mins = feature.unpack_feature_flat(blob)
    assert flat.dtype == np.float32 and np.array_equal(flat.astype(np.float64), g["flat"])
    assert np.array_equal(sizes, g["sizes"])
    assert np.array_equal(np.nonzero(~lowered)[0], [0, 7, 40])          # the size-0 records: one zero row each
    assert np.array_equal(thr.astype(np.float64), g["thr"]) and np.array_equal(ids, g["task_ids"])
    assert np.array_equal(mins.astype(np.float64), g["min_costs"])
    # drop-in tuple (feature.py:151-156)
    feats, thr2, ids2, mins2 = feature.unpack_feature(blob)
    assert feats.dtype == object and len(feats) == 41 and feats[0].shape == (1, 164) and not feats[0].any()
    assert np.array_equal(np.concatenate(list(feats), 0).astype(np.float64), g["flat"])
    assert np.array_equal(thr2, g["thr"]) and np.array_equal(ids2, g["task_ids"]) and np.array_equal(mins2, g["min_costs"])


def test_parser_large_blob_uses_threads_and_matches_oracle():
    rng = np.random.default_rng(3)
    feats = [None if rng.random() < 0.01 else rng.random((int(rng.integers(1, 33)), 164)).astype(np.float32) for _ in range(3500)]
    blob = FB.serialize_features(feats, rng.random(3500), rng.integers(0, 9, 3500), rng.random(9))
    assert len(blob) > 32 << 20                                           # > 8 MiB per copy thread
    want, thr, ids, mins = FB.unpack_feature(blob)
    flat, sizes, lowered, thr2, ids2, mins2 = feature.unpack_feature_flat(blob, pinned=False)
    assert np.array_equal(sizes, [len(f) for f in want])
    assert np.array_equal(flat, np.concatenate(list(want), 0).astype(np.float32))
    assert np.array_equal(lowered, [f is not None for f in feats])
    assert np.array_equal(thr2, thr.astype(np.float32)) and np.array_equal(ids2, ids) and np.array_equal(mins2, mins.astype(np.float32))


def test_parser_edge_cases_and_errors():
    empty = FB.serialize_features([], [], [], [])
    assert feature.scan(empty) == (0, 0, 0, 0, 0)
    flat, sizes, lowered, thr, ids, mins = feature.unpack_feature_flat(empty)
    assert flat.shape == (0, 164) and len(sizes) == 0
    feats, _, _, _ = feature.unpack_feature(empty)
    assert feats.shape == (0,)
    only_failed = FB.serialize_features([None, None], [0.0, 0.0], [0, 0], [1.0])
    flat, sizes, lowered, _, _, _ = feature.unpack_feature_flat(only_failed)
    assert flat.shape == (2, 164) and not flat.any() and not lowered.any() and np.array_equal(sizes, [1, 1])
    good = FB.serialize_features([np.ones((2, 164), np.float32)], [1.0], [0], [1.0])
    with pytest.raises(_lib.VaexError, match="vs"):                      # feature.py:157 "%d vs %d"
        feature.scan(good[:-4])
    with pytest.raises(_lib.VaexError, match="vs"):
        feature.scan(good + b"\0\0\0\0")
    with pytest.raises(_lib.VaexError, match="The length of feature vector is wrong. Expected 164 but got 100"):
        feature.scan(FB.serialize_features([np.ones((2, 100), np.float32)], [1.0], [0], [1.0]))
    with pytest.raises(_lib.VaexError, match="past the end"):
        feature.scan(good[:400])
    with pytest.raises(_lib.VaexError):
        feature.scan(b"")
    with pytest.raises(struct.error):                                     # the restatement fails where the reference does
        FB.unpack_feature(good[:-4])
    with pytest.raises(AssertionError, match="vs"):
        FB.unpack_feature(good + b"\0\0\0\0")
