"""a10 pin: the oracle's divisor legaliser (oracle/torch_oracle.py::legalise_divisor, which the CUDA legaliser is
bit-exact against in tests/test_gpu_parity.py) against the reference's OWN SplitFactorizationMemo
(src/auto_scheduler/search_policy/utils.cc:414-463), compiled from the reference's source lines into
oracle/_ref/libsplitmemo.so by oracle/Makefile.  CPU only."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT
from oracle import torch_oracle as O

SO = os.path.join(ROOT, "oracle", "_ref", "libsplitmemo.so")
MAXI = 64      # max_innermost_split_factor, python/tvm/auto_scheduler/search_policy.py:191


@pytest.fixture(scope="module")
def ref():
    if not os.path.isfile(SO) and os.path.isdir("/root/reference"):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
    if not os.path.isfile(SO):
        pytest.skip("oracle/_ref/libsplitmemo.so not built (needs /root/reference)")
    lib = ctypes.CDLL(SO)
    lib.ref_get_factors.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    lib.ref_factorization_schemes.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int64]
    lib.ref_factorization_schemes.restype = ctypes.c_int64
    lib.ref_scheme_members.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]
    return lib


def schemes(ref, extent, n, maxi=MAXI):
    cnt = ref.ref_factorization_schemes(extent, n, maxi, None, 0)
    out = np.empty((cnt, n), np.int32)
    ref.ref_factorization_schemes(extent, n, maxi, out.ctypes.data, cnt)
    return out


def members(ref, extent, n, rows, maxi=MAXI):
    rows = np.ascontiguousarray(rows, np.int32)
    m = np.zeros(len(rows), np.uint8)
    assert ref.ref_scheme_members(extent, n, maxi, rows.ctypes.data, len(rows), m.ctypes.data) == 0
    return m.astype(bool)


def legalise(extent, n, raw):
    ext = np.full(n, extent, np.int32)
    last = np.zeros(n, np.uint8)
    last[-1] = 1
    return O.legalise_divisor(raw, ext, last, MAXI)


def test_get_factors_equals_reference(ref):
    buf = np.empty(4096, np.int32)
    for n in list(range(1, 4097)) + [65536, 65537, 99991, 2 ** 20, 3 * 5 * 7 * 11 * 13 * 17]:
        c = ref.ref_get_factors(n, buf.ctypes.data, len(buf))
        assert buf[:c].tolist() == O.get_factors(n), n


def test_known_scheme_counts(ref):
    # hand-checkable anchors for the compiled reference: 12 = 2^2 * 3 with one length -> its six divisors;
    # two lengths (a, b): a | 12, b | 12 / a  ->  sum over a of d(12 / a) = 6 + 4 + 3 + 2 + 2 + 1 = 18
    assert schemes(ref, 12, 1)[:, 0].tolist() == [1, 2, 3, 4, 6, 12]
    assert len(schemes(ref, 12, 2)) == 18
    assert schemes(ref, 128, 1, 64)[:, 0].tolist() == [1, 2, 4, 8, 16, 32, 64]      # innermost <= 64 drops 128
    s = schemes(ref, 56, 3)
    assert ((56 % np.prod(s, axis=1)) == 0).all() and (s[:, -1] <= 64).all()


def test_legalised_rows_are_reference_schemes_every_extent(ref):
    """Every extent <= 4096, 1-4 split lengths: each legalised row is a member of the reference's enumerated set, the
    reference's schemes are fixed points, and each length is the nearest choice (ties -> smaller) among the values the
    reference allows in that slot given the lengths already fixed outside it."""
    rng = np.random.default_rng(11)
    for extent in range(1, 4097):
        for n in (1, 2, 3, 4):
            raw = np.maximum(1, np.round(np.exp(rng.uniform(0, np.log(extent + 2.0), size=(6, n))))).astype(np.int32)
            raw[0, :] = rng.integers(1, extent + 3, size=n)
            out = legalise(extent, n, raw)
            assert members(ref, extent, n, out).all(), (extent, n, raw, out)
    # fixed points + greedy-nearest on the full enumeration for a spread of extents (primes, powers, highly composite,
    # task 1264's loop bounds {1,3,4,28,56,57,64,128})
    for extent in [1, 2, 3, 4, 7, 28, 56, 57, 64, 97, 128, 224, 360, 512, 720, 1000, 1024, 2310, 2520, 3600, 4093, 4096]:
        for n in (1, 2, 3, 4):
            S = schemes(ref, extent, n)
            if len(S) > 20000:
                S = S[rng.choice(len(S), 20000, replace=False)]
            assert np.array_equal(legalise(extent, n, S), S), (extent, n)
            full = schemes(ref, extent, n)
            raw = rng.integers(1, extent + 3, size=(40, n)).astype(np.int32)
            out = legalise(extent, n, raw)
            for r, o in zip(raw, out):
                cand = full
                for j in range(n):
                    allowed = np.unique(cand[:, j])
                    d = np.abs(allowed.astype(np.int64) - int(r[j]))
                    best = allowed[np.flatnonzero(d == d.min())[0]]       # ties -> smaller (allowed is sorted)
                    assert o[j] == best, (extent, n, r, o, j)
                    cand = cand[cand[:, j] == o[j]]
