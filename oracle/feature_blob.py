"""TEST INFRASTRUCTURE ONLY -- CPU restatement of Ansor's packed per-store feature format.

serialize_features  follows SerializeFeatures, src/auto_scheduler/feature.cc:1562-1622 (the writer lives in
                    TVM's C++ and cannot be built here; tests need blobs, so the layout is restated);
unpack_feature      follows unpack_feature, python/tvm/auto_scheduler/feature.py:49-158, struct call by struct
                    call.  Pinned against the reference's own unpack_feature by tests/golden/feature_blob.npz
                    (oracle/gen_golden.py: the reference function is imported and run on the same bytes)."""
import struct

import numpy as np

VEC_LEN = 164


def serialize_features(features, normalized_throughputs, task_ids, min_costs):
    """features: list of fp32 arrays [n_stmts, vec_len], or None for a record that failed lowering (size 0).
    Per record the C++ side stores `float n_stmts` followed by the rows (feature.cc:1058)."""
    recs = []
    for f in features:
        if f is None:
            recs.append(np.zeros(0, np.float32))
        else:
            f = np.asarray(f, np.float32)
            recs.append(np.concatenate([[np.float32(f.shape[0])], f.reshape(-1)]).astype(np.float32))
    n = len(recs)
    sizes = [n] + [len(r) for r in recs] + [len(normalized_throughputs), len(task_ids), len(min_costs)]
    out = bytearray(struct.pack("%di" % len(sizes), *sizes))
    for r in recs:
        out += r.tobytes()
    out += np.asarray(normalized_throughputs, np.float32).tobytes()
    out += np.asarray(task_ids, np.int32).tobytes()
    out += np.asarray(min_costs, np.float32).tobytes()
    return bytes(out)


def unpack_feature(byte_arr, vec_len=VEC_LEN):
    offset = 0
    n = struct.unpack_from("1i", byte_arr, offset=offset)[0]
    offset += 4
    sizes = struct.unpack_from("%di" % (n + 3), byte_arr, offset=offset)
    offset += 4 * (n + 3)
    features = []
    for size in sizes[:-3]:
        if size == 0:
            features.append(np.zeros((1, vec_len)))
            continue
        n_stmts = int(struct.unpack_from("f", byte_arr, offset=offset)[0] + 0.5)
        offset += 4
        tmp = (size - 1) // n_stmts
        assert tmp == vec_len, "The length of feature vector is wrong. Expected %d but got %d." % (vec_len, tmp)
        assert tmp * n_stmts == size - 1
        rows = np.frombuffer(byte_arr, dtype=np.float32, count=n_stmts * vec_len, offset=offset)
        offset += 4 * n_stmts * vec_len
        features.append(rows.reshape(n_stmts, vec_len).astype(np.float64))
    m = sizes[-3]
    thr = np.array(struct.unpack_from("%df" % m, byte_arr, offset=offset))
    offset += 4 * m
    m = sizes[-2]
    task_ids = np.array(struct.unpack_from("%di" % m, byte_arr, offset=offset))
    offset += 4 * m
    m = sizes[-1]
    min_costs = np.array(struct.unpack_from("%df" % m, byte_arr, offset=offset))
    offset += 4 * m
    assert offset == len(byte_arr), "%d vs %d" % (offset, len(byte_arr))
    out = np.empty(len(features), dtype=object)
    for i, f in enumerate(features):
        out[i] = f
    return out, thr, task_ids, min_costs
