"""Input containers of the cost-model path.  Mirrors the interface of the reference's
python/tvm/auto_scheduler/dataset.py:13-211 (LearningTask, Dataset) for the members the hot
path touches; record-file parsing (make_dataset_from_log_file, :214-287) is TVM-side and out
of scope."""
from collections import OrderedDict, namedtuple

import numpy as np

LearningTask = namedtuple("LearningTask", ["workload_key", "target"])


def input_to_learning_task(inp):
    """dataset.py:16-17."""
    return LearningTask(inp.task.workload_key, str(inp.task.target))


class Dataset:
    def __init__(self):
        self.raw_files = None
        self.features = OrderedDict()      # LearningTask -> object array of [S_i, F] float32
        self.throughputs = OrderedDict()   # LearningTask -> normalised throughputs
        self.min_latency = {}
        self.measure_records = {}

    @staticmethod
    def create_one_task(task, features, throughputs, min_latency=None):
        ret = Dataset()
        ret.load_task_data(task, features, throughputs, min_latency)
        return ret

    def load_task_data(self, task, features, throughputs, min_latency=None):
        if task not in self.features:
            self.features[task] = features
            self.throughputs[task] = throughputs
            self.min_latency[task] = min_latency
            return
        merged = list(self.features[task]) + list(features)
        arr = np.empty(len(merged), dtype=object)
        for i, f in enumerate(merged):
            arr[i] = f
        self.features[task] = arr
        assert min_latency is not None
        comb = min(self.min_latency[task], min_latency)
        self.throughputs[task] = np.concatenate([
            np.asarray(self.throughputs[task]) * (comb / self.min_latency[task]),
            np.asarray(throughputs) * (comb / min_latency)])
        self.min_latency[task] = comb

    def update_from_measure_pairs(self, inputs, results, featurize=None):
        """dataset.py:39-57: group the measured pairs by task, extract their per-store features and append them.
        featurize(inputs, results) -> (features, normalized_throughputs, task_ids, min_latency) defaults to
        feature.get_per_store_features_from_measure_pairs (TVM's extractor + this library's blob parser)."""
        if featurize is None:
            from .feature import get_per_store_features_from_measure_pairs as featurize
        new_data = {}
        for inp, res in zip(inputs, results):
            store = new_data.setdefault(input_to_learning_task(inp), ([], []))
            store[0].append(inp)
            store[1].append(res)
        for task, (t_inputs, t_results) in new_data.items():
            features, normalized_throughputs, task_ids, min_latency = featurize(t_inputs, t_results)
            assert not np.any(task_ids)   # all task ids should be zero
            assert len(min_latency) == 1  # should have only one task
            self.load_task_data(task, features, normalized_throughputs, min_latency[0])

    def update_from_dataset(self, dataset):
        for task in dataset.features:
            if task not in self.features:
                self.load_task_data(task, dataset.features[task], dataset.throughputs[task],
                                    dataset.min_latency[task])

    def tasks(self):
        return list(self.features.keys()) if self.features else list(self.measure_records.keys())

    def targets(self):
        return list({t.target for t in self.tasks()})

    def extract_subset(self, tasks):
        ret = Dataset()
        for task in tasks:
            if task in self.features:
                ret.load_task_data(task, self.features[task], self.throughputs[task], self.min_latency[task])
        return ret

    def __getstate__(self):
        return self.raw_files, self.features, self.throughputs, self.min_latency, 0.1

    def __setstate__(self, value):
        self.raw_files, self.features, self.throughputs, self.min_latency, _ = value
        self.measure_records = {}

    def __len__(self):
        return sum(len(x) for x in self.features.values())
