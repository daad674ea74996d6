"""Query sample library (reference: models/rnnt_qsl.py:5-34).

Holds the pre-extracted features in host memory — page-locked when a GPU is present, so that the per-batch
host->device copies of the SUT are asynchronous DMA.  The page-locked copy is ONE arena in which the loaded samples
lie longest first (ties by sample id): an Offline query sorted by length then reads contiguous arena ranges, so the
SUT moves a whole batch with a single DMA instead of one per sample.  Samples are still handed out one by one
(`qsl[i]` -> (features view, length)), exactly as the reference does.  `dataset_dir` is the reference's on-disk format, a torch-saved
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
    def __init__(self, dataset_dir, perf_count=None, pin_memory=None, preload=True):
        data = torch.load(dataset_dir) if isinstance(dataset_dir, (str, os.PathLike)) else dataset_dir
        self.dataset = list(data.values())      # [x list, x_lens list]
        self.count = len(self.dataset[0])
        perf_count = self.count if perf_count is None else perf_count
        self.pin_memory = torch.cuda.is_available() if pin_memory is None else pin_memory
        self.arena = None                        # pinned [sum T_i, 240] fp32, samples longest first
        self.arena_off = {}                      # sample id -> first arena row
        # lengths come from the dataset's x_lens like the reference (pytorch_sut.py:88; the sort key too); a sample whose
        # tensor holds more rows than x_lens says is used up to x_lens only (the encoder is causal and everything
        # downstream reads t < x_lens, so the tail can never reach a transcript); fewer rows than x_lens is an error
        self.x_lens_host = [int(n) for n in self.dataset[1]]
        for i, (x, n) in enumerate(zip(self.dataset[0], self.x_lens_host)):
            if int(x.shape[0]) < n:
                raise RuntimeError(f"RNNTQSL: sample {i} has {int(x.shape[0])} feature rows but x_lens says {n}")
        self.rows = list(self.x_lens_host)                           # feature rows used per sample, plain ints
        self._keys = [(-n, i) for i, n in enumerate(self.x_lens_host)]
        self.sample_id_to_sample = {}
        self.qsl = lg.ConstructQSL(self.count, perf_count, self.load_query_samples, self.unload_query_samples)
        if preload:          # the reference loads the whole set at construction (rnnt_qsl.py:11-14); a rank of a multi-GPU front
            self.load_query_samples(range(self.count))     # loads only the samples it will be dealt (preload=False)

    def load_query_samples(self, sample_list):
        new = [i for i in sample_list if i not in self.sample_id_to_sample]
        for sample_id in new:
            self.sample_id_to_sample[sample_id] = (self.dataset[0][sample_id], self.dataset[1][sample_id])
        if new and self.pin_memory:
            self._build_arena()

    def unload_query_samples(self, sample_list):
        for sample_id in sample_list:
            self.sample_id_to_sample.pop(sample_id, None)
            self.arena_off.pop(sample_id, None)

    def _build_arena(self):
        """(Re)pack every loaded sample into one page-locked arena, longest first.  Runs when samples are loaded, i.e.
        outside LoadGen's timed region (LoadSamplesToRam)."""
        ids = sorted(self.sample_id_to_sample, key=self._keys.__getitem__)
        rows = [self.rows[i] for i in ids]
        first = self.dataset[0][ids[0]]
        shape = (sum(rows),) if first.dim() == 1 else (sum(rows), int(first.shape[1]))     # 1-D: waveforms (--enable_process)
        arena = torch.empty(shape, dtype=torch.float32, pin_memory=True)
        off = 0
        for i, n in zip(ids, rows):
            arena[off:off + n].copy_(self.dataset[0][i][:n])
            self.arena_off[i] = off
            self.sample_id_to_sample[i] = (arena[off:off + n], self.dataset[1][i])
            off += n
        self.arena = arena

    def sort_key(self, index):
        """Offline sort key: feature length descending (pytorch_sut.py:58-65); equal lengths by sample id = arena order,
        so that a sorted query walks the arena front to back.  Plain tuples of ints, made once."""
        return self._keys[index]

    def __getitem__(self, index):
        return self.sample_id_to_sample[index]

    def __del__(self):
        try:
            lg.DestroyQSL(self.qsl)
        except Exception:
            pass
