"""MLPModel.update / update_from_file / Dataset.update_from_measure_pairs / shuffled SegmentDataLoader iteration
(reference: python/tvm/auto_scheduler/cost_model/mlp_model.py:107-144,824-851, dataset.py:39-57), driven on the CPU
through stub MeasureInput / MeasureResult objects and a stub of TVM's feature FFI that returns a packed blob."""
import collections
import sys
import types

import numpy as np
import pytest
import torch

import vaex_b200
from vaex_b200 import _lib
from vaex_b200.dataset import Dataset, LearningTask, input_to_learning_task
from oracle import feature_blob as FB

Task = collections.namedtuple("Task", ["workload_key", "target"])
MeasureInput = collections.namedtuple("MeasureInput", ["task", "state"])
MeasureResult = collections.namedtuple("MeasureResult", ["costs"])


def make_pairs(rng, task, n):
    inputs = [MeasureInput(task, rng.random((int(rng.integers(1, 9)), 164), dtype=np.float32)) for _ in range(n)]
    results = [MeasureResult([float(rng.uniform(1e-4, 1e-2))]) for _ in range(n)]
    return inputs, results


def install_stub_tvm(monkeypatch, calls):
    """tvm.auto_scheduler._ffi_api.GetPerStoreFeaturesFromMeasurePairs -> the packed bytes TVM's C++ would return
    (SerializeFeatures layout): one record per pair, throughput = min cost / cost, task id 0."""
    def get_pairs(inputs, results, skip, max_n_bufs):
        calls.append((len(inputs), skip, max_n_bufs))
        costs = np.array([np.mean(r.costs) for r in results])
        return FB.serialize_features([i.state for i in inputs], costs.min() / costs, np.zeros(len(inputs), np.int32), [costs.min()])
    tvm, autos, ffi = types.ModuleType("tvm"), types.ModuleType("tvm.auto_scheduler"), types.ModuleType("tvm.auto_scheduler._ffi_api")
    ffi.GetPerStoreFeaturesFromMeasurePairs = get_pairs
    tvm.auto_scheduler, autos._ffi_api = autos, ffi
    for name, mod in (("tvm", tvm), ("tvm.auto_scheduler", autos), ("tvm.auto_scheduler._ffi_api", ffi)):
        monkeypatch.setitem(sys.modules, name, mod)


def test_update_from_measure_pairs_groups_by_task(monkeypatch):
    calls = []
    install_stub_tvm(monkeypatch, calls)
    rng = np.random.default_rng(0)
    ta, tb = Task("wk_a", "cuda"), Task("wk_b", "llvm")
    ia, ra = make_pairs(rng, ta, 5)
    ib, rb = make_pairs(rng, tb, 3)
    ds = Dataset()
    ds.update_from_measure_pairs(ia[:2] + ib + ia[2:], ra[:2] + rb + ra[2:])
    assert [c[0] for c in calls] == [5, 3] and calls[0][1:] == (0, 5)
    la, lb = LearningTask("wk_a", "cuda"), LearningTask("wk_b", "llvm")
    assert input_to_learning_task(ia[0]) == la
    assert ds.tasks() == [la, lb] and len(ds) == 8
    for got, inp in zip(ds.features[la], ia):
        assert np.array_equal(got, inp.state)
    ca = np.array([r.costs[0] for r in ra])
    np.testing.assert_allclose(ds.throughputs[la], ca.min() / ca, rtol=1e-6)
    assert ds.min_latency[la] == pytest.approx(ca.min(), rel=1e-6)
    # a second round for the same task appends and renormalises to the combined minimum (dataset.py:66-95)
    ia2, ra2 = make_pairs(rng, ta, 4)
    ra2[1] = MeasureResult([ca.min() / 2])
    ds.update_from_measure_pairs(ia2, ra2)
    assert len(ds.features[la]) == 9
    call = np.concatenate([ca, [r.costs[0] for r in ra2]])
    np.testing.assert_allclose(ds.throughputs[la], call.min() / call, rtol=1e-5)


def test_mlp_model_update_refits_on_the_whole_dataset(monkeypatch):
    calls = []
    install_stub_tvm(monkeypatch, calls)
    rng = np.random.default_rng(1)
    model = vaex_b200.MLPModel()
    assert isinstance(model, vaex_b200.mlp_model.PythonBasedModel)
    fits = []
    monkeypatch.setattr(model.model, "fit_base", lambda ds, *a: fits.append(len(ds)))
    task = Task("wk", "cuda")
    model.update([], [])
    assert not fits and not calls                              # mlp_model.py:825: nothing to do
    model.update(*make_pairs(rng, task, 6))
    model.update(*make_pairs(rng, task, 4))
    assert fits == [6, 10]                                     # re-fit on everything seen so far
    frozen = vaex_b200.MLPModel(disable_update=True)
    monkeypatch.setattr(frozen.model, "fit_base", lambda ds, *a: fits.append(-1))
    frozen.update(*make_pairs(rng, task, 3))
    assert fits == [6, 10] and len(frozen.dataset) == 0
    # without a GPU the re-fit itself fails loudly (no CPU fallback)
    if not torch.cuda.is_available():
        real = vaex_b200.MLPModel()
        with pytest.raises(_lib.VaexError):
            real.update(*make_pairs(rng, task, 4))
    with pytest.raises(_lib.VaexError):                         # the stub tvm has no measure_record.RecordReader
        model.update_from_file("records.json")


def test_update_from_file_reads_records_then_updates(monkeypatch):
    calls = []
    install_stub_tvm(monkeypatch, calls)
    rng = np.random.default_rng(2)
    pairs = make_pairs(rng, Task("wk", "cuda"), 7)
    mr = types.ModuleType("tvm.auto_scheduler.measure_record")

    class RecordReader:
        def __init__(self, file_name):
            self.file_name = file_name

        def read_lines(self, n_lines=None):
            return (pairs[0][:n_lines], pairs[1][:n_lines]) if n_lines else pairs
    mr.RecordReader = RecordReader
    monkeypatch.setitem(sys.modules, "tvm.auto_scheduler.measure_record", mr)
    model = vaex_b200.MLPModel()
    fits = []
    monkeypatch.setattr(model.model, "fit_base", lambda ds, *a: fits.append(len(ds)))
    model.update_from_file("log.json", n_lines=5)
    assert fits == [5] and calls[0][0] == 5


def test_shuffled_loader_matches_reference_order():
    """SegmentDataLoader with shuffle=True draws torch.randperm per epoch and gathers the programs' rows in that order
    (mlp_model.py:107-141)."""
    rng = np.random.default_rng(3)
    feats = np.empty(23, dtype=object)
    for i in range(23):
        feats[i] = rng.random((int(rng.integers(1, 6)), 164), dtype=np.float32)
    thr = rng.random(23).astype(np.float32)
    ds = Dataset.create_one_task(LearningTask("wk", "cuda"), feats, thr)
    loader = vaex_b200.SegmentDataLoader(ds, 5, "cpu", False, False, {}, None, shuffle=True)
    torch.manual_seed(7)
    batches = list(loader)
    torch.manual_seed(7)
    order = torch.randperm(23).numpy()
    assert len(batches) == 5 and sum(len(b[0]) for b in batches) == 23
    pos = 0
    for sizes, rows, labels in batches:
        idx = order[pos:pos + len(sizes)]
        pos += len(sizes)
        assert sizes.tolist() == [len(feats[i]) for i in idx]
        assert np.array_equal(rows.numpy(), np.concatenate([feats[i] for i in idx]))
        assert np.array_equal(labels.numpy(), thr[idx])
    torch.manual_seed(8)
    assert [b[0].tolist() for b in loader] != [b[0].tolist() for b in batches]      # a new permutation every epoch
    plain = vaex_b200.SegmentDataLoader(ds, 5, "cpu", False, False, {}, None)
    assert np.array_equal(torch.cat([b[1] for b in plain]).numpy(), np.concatenate(list(feats)))
