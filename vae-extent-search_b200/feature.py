"""Drop-in for the unpacking half of python/tvm/auto_scheduler/feature.py.

`unpack_feature(byte_arr)` keeps the reference's signature and return tuple (feature.py:49-158) but
parses the packed blob (SerializeFeatures, src/auto_scheduler/feature.cc:1562-1622) in one C++ pass
instead of a struct.unpack_from per row; `unpack_feature_flat` returns the layout the kernels want
(one [rows, 164] fp32 matrix + int32 rows-per-record) without building per-record Python objects."""
import ctypes

import numpy as np

from . import _lib

DEFAULT_FEATURE_VEC_LEN = 164      # feature.py:41


def _as_bytes(byte_arr):
    """bytes / bytearray / memoryview / TVMByteArray-like -> (buffer object kept alive, address, length)."""
    if isinstance(byte_arr, np.ndarray):
        a = np.ascontiguousarray(byte_arr).view(np.uint8).reshape(-1)
    else:
        a = np.frombuffer(byte_arr, dtype=np.uint8)
    return a, a.ctypes.data, a.size


def scan(byte_arr, vec_len=DEFAULT_FEATURE_VEC_LEN):
    """(records, rows, throughputs, task ids, min costs) after validating the blob."""
    keep, addr, n = _as_bytes(byte_arr)
    counts = (ctypes.c_int64 * 5)()
    _lib.check(_lib.load().vaex_feature_blob_scan(addr, n, vec_len, counts))
    return tuple(int(c) for c in counts)


def unpack_feature_flat(byte_arr, vec_len=DEFAULT_FEATURE_VEC_LEN, pinned=False):
    """-> (features fp32 [rows, vec_len], sizes int32 [n], lowered bool [n], throughputs fp32, task_ids int32,
    min_costs fp32).  A record that failed lowering is one all-zero row with lowered[i] == False.
    pinned=True puts the feature matrix in page-locked memory (a torch tensor's buffer) for async H2D."""
    keep, addr, nbytes = _as_bytes(byte_arr)
    n, rows, n_thr, n_task, n_min = scan(keep, vec_len)
    if pinned:
        import torch
        holder = torch.empty((rows, vec_len), dtype=torch.float32, pin_memory=True)
        feats = holder.numpy()
    else:
        feats = np.empty((rows, vec_len), dtype=np.float32)
    sizes = np.empty(n, dtype=np.int32)
    lowered = np.empty(n, dtype=np.uint8)
    thr, task_ids, min_costs = np.empty(n_thr, np.float32), np.empty(n_task, np.int32), np.empty(n_min, np.float32)
    _lib.check(_lib.load().vaex_feature_blob_unpack(addr, nbytes, vec_len, feats.ctypes.data, sizes.ctypes.data,
                                                    lowered.ctypes.data, thr.ctypes.data, task_ids.ctypes.data,
                                                    min_costs.ctypes.data))
    return feats, sizes, lowered.astype(bool), thr, task_ids, min_costs


def unpack_feature(byte_arr):
    """feature.py:49-158: -> (features: object array of [n_stmts, 164] arrays, normalized_throughputs,
    task_ids, min_costs).  The per-record arrays are fp32 views into one flat matrix (the reference builds
    float64 copies of the same fp32 values)."""
    feats, sizes, _, thr, task_ids, min_costs = unpack_feature_flat(byte_arr)
    out = np.empty(len(sizes), dtype=object)
    ends = np.cumsum(sizes, dtype=np.int64)
    for i, e in enumerate(ends):
        out[i] = feats[e - sizes[i]:e]
    return out, thr.astype(np.float64), task_ids.astype(np.int64), min_costs.astype(np.float64)


DEFAULT_MAX_N_BUFS = 5             # feature.py:38


def get_per_store_features_from_measure_pairs(inputs, results, skip_first_n_feature_extraction=0, max_n_bufs=None):
    """feature.py:192-225: the extraction itself is TVM's C++ (GetPerStoreFeaturesFromMeasurePairs lowers every
    program, src/auto_scheduler/feature.cc:1425-1470) and stays there; the bytes it returns are parsed by the C++
    blob parser of this library instead of the per-row struct.unpack_from loop.  Needs an importable `tvm`."""
    try:
        from tvm.auto_scheduler import _ffi_api
    except ImportError as e:
        raise _lib.VaexError("feature extraction from MeasureInput/MeasureResult pairs needs TVM "
                             "(auto_scheduler._ffi_api.GetPerStoreFeaturesFromMeasurePairs): %s" % e)
    byte_arr = _ffi_api.GetPerStoreFeaturesFromMeasurePairs(inputs, results, skip_first_n_feature_extraction,
                                                            max_n_bufs or DEFAULT_MAX_N_BUFS)
    return unpack_feature(byte_arr)
