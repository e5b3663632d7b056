"""The allocation-free entry point vaex_segment_offsets_noalloc (include/vaex_b200.h) against the existing
vaex_segment_offsets, through the C-ABI.  Collected last on purpose: a failure here must not
hide any other test.  Passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest
import torch

import vaex_b200
from vaex_b200 import _lib

pytestmark = pytest.mark.gpu

NOALLOC_MAX = 16384 * 2048      # VAEX_SEGMENT_OFFSETS_NOALLOC_MAX


def _both(lib, s):
    n = s.numel()
    ref = torch.empty(n + 1, dtype=torch.int32, device="cuda")
    got = torch.full((n + 1,), -7, dtype=torch.int32, device="cuda")
    _lib.check(lib.vaex_segment_offsets(_lib.ptr(s), n, _lib.ptr(ref), _lib.stream_ptr()))
    _lib.check(lib.vaex_segment_offsets_noalloc(_lib.ptr(s), n, _lib.ptr(got), _lib.stream_ptr()))
    return ref, got


def test_noalloc_scan_matches_existing_entry_point_twice_in_a_row():
    lib = _lib.load()
    g = torch.Generator(device="cuda").manual_seed(5)
    # two calls in a row share the static scratch: a large call first, then a smaller one that must not see its leftovers
    for n in [1000003, 4099, 1000003, 1, 0, 2048, 2049]:
        s = torch.randint(0, 33, (n,), dtype=torch.int32, device="cuda", generator=g)
        ref, got = _both(lib, s)
        assert torch.equal(ref, got), "n = %d" % n                  # integer work: bit-exact
        want = np.concatenate([[0], np.cumsum(s.cpu().numpy().astype(np.int64))])
        assert np.array_equal(got.cpu().numpy(), want)
    # the same input twice, back to back
    s = torch.randint(1, 33, (300000,), dtype=torch.int32, device="cuda", generator=g)
    a = _both(lib, s)[1].clone()
    b = _both(lib, s)[1]
    assert torch.equal(a, b)


def test_noalloc_scan_allocates_nothing_and_rejects_oversize():
    lib = _lib.load()
    s = torch.ones(50000, dtype=torch.int32, device="cuda")
    out = torch.empty(50001, dtype=torch.int32, device="cuda")
    _lib.check(lib.vaex_segment_offsets_noalloc(_lib.ptr(s), 50000, _lib.ptr(out), _lib.stream_ptr()))   # address resolved
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(20):
        _lib.check(lib.vaex_segment_offsets_noalloc(_lib.ptr(s), 50000, _lib.ptr(out), _lib.stream_ptr()))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0                    # no device memory taken by the calls
    assert int(out[-1]) == 50000
    # beyond the stated maximum: a clear error, nothing launched (the pointers are never dereferenced)
    rc = lib.vaex_segment_offsets_noalloc(_lib.ptr(s), NOALLOC_MAX + 1, _lib.ptr(out), _lib.stream_ptr())
    assert rc == -1
    msg = lib.vaex_last_error().decode()
    assert "exceeds" in msg and str(NOALLOC_MAX) in msg
    with pytest.raises(_lib.VaexError):
        _lib.check(rc)
