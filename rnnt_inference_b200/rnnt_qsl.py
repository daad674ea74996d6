"""Query sample library (reference: models/rnnt_qsl.py:5-34).

Holds the pre-extracted features in host memory — page-locked when a GPU is present, so that the per-batch
host->device copies of the SUT are asynchronous DMA.  `dataset_dir` is the reference's on-disk format, a torch-saved
dict {"x": [T_i x 240 fp32 tensors], "x_lens": [int tensors]} (datasets/convert_librispeech.py:135-163), or that
dict itself.
"""
import os

import torch

try:
    import mlperf_loadgen as lg
except ImportError:   # not installed offline
    from . import loadgen_shim as lg


class RNNTQSL:
    def __init__(self, dataset_dir, perf_count=None, pin_memory=None):
        data = torch.load(dataset_dir) if isinstance(dataset_dir, (str, os.PathLike)) else dataset_dir
        self.dataset = list(data.values())      # [x list, x_lens list]
        self.count = len(self.dataset[0])
        perf_count = self.count if perf_count is None else perf_count
        self.pin_memory = torch.cuda.is_available() if pin_memory is None else pin_memory
        self._pinned = {}                        # id(tensor) -> pinned copy (replicated samples share storage)
        self.sample_id_to_sample = {}
        self.qsl = lg.ConstructQSL(self.count, perf_count, self.load_query_samples, self.unload_query_samples)
        self.load_query_samples(range(self.count))

    def load_query_samples(self, sample_list):
        for sample_id in sample_list:
            if sample_id not in self.sample_id_to_sample:
                self.sample_id_to_sample[sample_id] = self._load_sample(sample_id)

    def unload_query_samples(self, sample_list):
        for sample_id in sample_list:
            self.sample_id_to_sample.pop(sample_id, None)

    def _load_sample(self, index):
        x, x_len = self.dataset[0][index], self.dataset[1][index]
        if self.pin_memory:
            key = id(x)
            if key not in self._pinned:
                self._pinned[key] = x.contiguous().pin_memory()
            x = self._pinned[key]
        return (x, x_len)

    def __getitem__(self, index):
        return self.sample_id_to_sample[index]

    def __del__(self):
        try:
            lg.DestroyQSL(self.qsl)
        except Exception:
            pass
