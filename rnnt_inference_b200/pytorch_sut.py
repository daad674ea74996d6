"""System under test (reference: models/pytorch_sut.py:20-125).

PytorchSUT(model_path, dataset_dir, batch_size, run_mode, args) with the reference's issue/complete contract:
  issue_queries(samples)      Offline: sort by feature length descending, cut into batch_size chunks, run each,
                              respond per chunk (pytorch_sut.py:58-65)
  inference(batch_idx)        pad T to the batch maximum, pad N up to a multiple of 32 with zero-length lanes
                              (:85-104), run the greedy decoder, return (res, res_len)
  query_samples_complete(...) one QuerySampleResponse(id, ptr, n_tokens * 4) + one QuerySamplesComplete per sample
                              (:107-118); tokens are int32
What changes underneath on the B200: the batch travels as the UNPADDED concatenation of its utterances (one async
copy per pinned sample on a copy stream), padding/quantisation happen in one device kernel, and the Offline loop is
software-pipelined — batch i+1 is staged while batch i computes and batch i-1 is answered, chunks running shortest
first so that the one exposed copy is the smallest, and the short end of the query cut by frames so that the first
chunk's compute covers the next chunk's copy (sharding.plan_chunks; head_fraction=0 restores the reference's plain
batch_size chunks) — so the GPU never waits for the host.
"""
import math
import os
from types import SimpleNamespace

import numpy as np
import torch

try:
    import mlperf_loadgen as lg
except ImportError:   # not installed offline
    from . import loadgen_shim as lg

from . import _lib, sharding
from .decoder import GreedyDecoder
from .modeling_rnnt import RNNT
from .rnnt_qsl import RNNTQSL
from .utils import logger, seq_to_sen

_DEFAULT_ARGS = dict(enable_bf16=False, split_len=-1, scenario="Offline", load_jit=False, save_jit=False,
                     enable_process=False, toml_path="", processor_path="", exact_order=False)


class _Batch:
    """One staged / in-flight batch of the pipelined Offline loop."""
    __slots__ = ("samples", "n_actual", "n_padded", "max_len", "packed", "meta", "ready", "done", "res", "res_len", "tok_host",
                 "len_host", "cap", "h2d_bytes", "keep", "lens_host", "status_host", "stream")


class PytorchSUT:
    def __init__(self, model_path, dataset_dir, batch_size=1, run_mode=None, args=None, **kwargs):
        merged = dict(_DEFAULT_ARGS)
        if args is not None:
            merged.update({k: v for k, v in vars(args).items()})
        merged.update(kwargs)
        args = SimpleNamespace(**merged)
        self.processor = None
        if args.enable_process:
            # pytorch_sut.py:24-35: the featurizer in front of the model, configured by [input_eval] of the toml file (the
            # MLPerf RNN-T values when no file is given); the samples of the QSL are then 1-D waveforms
            from .features import AudioProcessing, MLPERF_INPUT_EVAL
            cfg = dict(MLPERF_INPUT_EVAL)
            if args.toml_path and os.path.exists(args.toml_path):
                import tomllib
                with open(args.toml_path, "rb") as fh:
                    cfg = tomllib.load(fh)["input_eval"]
            self.processor = AudioProcessing(run_mode, pad_batch_size=True, **cfg).eval()
        self.enable_process = self.processor is not None
        self.batch_size = batch_size
        self.run_mode = run_mode
        self.device = kwargs.get("device")
        rnnt = RNNT(model_path, run_mode, args.enable_bf16, args.load_jit, device=self.device).eval()
        self.model = GreedyDecoder(rnnt, run_mode, args.enable_bf16, args.split_len, self.batch_size,
                                   exact_order=args.exact_order)
        self.scenario = args.scenario if run_mode != "calib" else None
        self.batch_sort = self.scenario == "Offline"
        self.pipelined = run_mode == "quant" and args.split_len == -1 and not self.enable_process
        self.qsl = RNNTQSL(dataset_dir, preload=kwargs.get("preload", True))
        self.sut = lg.ConstructSUT(self.issue_queries, self.flush_queries)
        self.response_hook = None   # callable(samples, tokens, counts) replacing lg.QuerySamplesComplete
        self.h2d_bytes = 0          # running totals, read by bench.py
        self.d2h_bytes = 0
        self.batches_run = 0
        # pipelined Offline path: share of the short end's frames that runs first (sharding.plan_chunks); 0 = plain chunks
        self.head_fraction = float(kwargs.get("head_fraction", os.environ.get("RNNT_SUT_HEAD_FRACTION", 0.15)))
        self.head_min_frames = int(kwargs.get("head_min_frames", 65536))
        if self.pipelined:
            dev = self.model.device
            self._copy_stream = torch.cuda.Stream(dev)
            # Consecutive chunks alternate between two compute streams (each with its own scratch, ops.Workspace): the
            # GEMMs and byte kernels of one chunk run on the SMs the other chunk's recurrent kernel leaves idle (128 of 148
            # CTAs).  The library orders the COOPERATIVE kernels (recurrent, decode) of one device whatever stream they come
            # from (api.cu, CoopOrder), so two of them are never resident together.
            n_streams = max(1, int(kwargs.get("compute_streams", os.environ.get("RNNT_SUT_COMPUTE_STREAMS", 2))))
            self._compute_streams = [torch.cuda.Stream(dev) for _ in range(n_streams)]
            self._launched = 0
            self._packed = [None, None]         # double-buffered device staging of the ragged batch
            self._slot = 0
            self._prev_done = [None, None]

    # --------------------------------------------------------------------------------------------- LoadGen entry
    def issue_queries(self, samples):
        if self.batch_sort:
            keys = self.qsl._keys
            samples.sort(key=lambda s: keys[s.index])
        if not self.pipelined:
            chunks = [samples[i:i + self.batch_size] for i in range(0, len(samples), self.batch_size)]
            for chunk in chunks:
                results, results_idx = self.inference([s.index for s in chunk])
                self.query_samples_complete(chunk, results, results_idx)
            return
        # Shortest chunk first: its host->device copy (the only one nothing can hide) is the smallest, and every later,
        # larger copy overlaps the previous chunk's compute; sharding.plan_chunks sizes the first chunk so that its
        # compute covers the next copy.  Chunks are independent and each is answered whole, so the order in which they
        # run changes no transcript (LoadGen accepts completions in any order).
        if self.batch_sort:
            x_lens = self.qsl.x_lens_host
            plan = sharding.plan_chunks([x_lens[s.index] for s in samples], self.batch_size, self.head_fraction,
                                        self.head_min_frames)
            chunks = [samples[lo:hi] for lo, hi in plan]
        else:
            chunks = [samples[i:i + self.batch_size] for i in range(0, len(samples), self.batch_size)][::-1]
        staged = self._stage(chunks[0]) if chunks else None
        in_flight = None
        for i in range(len(chunks)):
            current = staged
            self._launch(current)
            staged = self._stage(chunks[i + 1]) if i + 1 < len(chunks) else None
            if in_flight is not None:
                self._finish(in_flight)
            in_flight = current
        if in_flight is not None:
            self._finish(in_flight)

    def flush_queries(self):
        pass

    # --------------------------------------------------------------------------------------------- pipelined path
    def _stage(self, chunk):
        """Host->device copy of one batch on the copy stream: every pinned sample goes straight into its slice of a
        packed [sum T_i, 240] device buffer; offsets and lengths travel as one small tensor."""
        dev = self.model.device
        b = _Batch()
        b.samples = chunk
        b.n_actual = len(chunk)
        b.n_padded = math.ceil(b.n_actual / 32) * 32          # pytorch_sut.py:90-104: pad N to a multiple of 32
        x_lens = self.qsl.x_lens_host        # the dataset's x_lens like the reference (pytorch_sut.py:88), = rows (checked by the QSL)
        lens = [x_lens[s.index] for s in chunk]
        b.max_len = max(lens)
        b.lens_host = lens
        meta = torch.zeros((2, b.n_padded), dtype=torch.int64)
        meta[0, :b.n_actual] = torch.tensor(np.concatenate([[0], np.cumsum(lens[:-1])]), dtype=torch.int64)
        meta[1, :b.n_actual] = torch.tensor(lens, dtype=torch.int64)
        total = int(sum(lens))
        slot = self._slot
        self._slot ^= 1
        with torch.cuda.stream(self._copy_stream):
            if self._prev_done[slot] is not None:
                self._copy_stream.wait_event(self._prev_done[slot])   # the buffer's previous batch must have been consumed
            buf = self._packed[slot]
            if buf is None or buf.shape[0] < total:
                grow = max(total, self.batch_size * 512)
                buf = torch.empty((grow, self.qsl[chunk[0].index][0].shape[1]), dtype=torch.float32, device=dev)
                self._packed[slot] = buf
            # one DMA per run of samples that are neighbours in the QSL's pinned arena (a sorted Offline query: one run
            # per batch); samples outside the arena are copied one by one
            arena, aoff = self.qsl.arena, self.qsl.arena_off
            off = i = 0
            while i < len(chunk):
                start = aoff.get(chunk[i].index) if arena is not None else None
                if start is None:
                    buf[off:off + lens[i]].copy_(self.qsl[chunk[i].index][0], non_blocking=True)
                    off += lens[i]
                    i += 1
                    continue
                rows = lens[i]
                j = i + 1
                while j < len(chunk) and aoff.get(chunk[j].index) == start + rows:
                    rows += lens[j]
                    j += 1
                buf[off:off + rows].copy_(arena[start:start + rows], non_blocking=True)
                off += rows
                i = j
            meta_host = meta.pin_memory()
            b.meta = meta_host.to(dev, non_blocking=True)
            b.packed = buf
            b.ready = torch.cuda.Event()
            b.ready.record(self._copy_stream)
        b.keep = (meta_host, slot)
        b.h2d_bytes = total * buf.shape[1] * 4 + meta.numel() * 8
        return b

    def _launch(self, b):
        """Encoder + fused decode of a staged batch on the compute stream, then the (small) result copy back."""
        b.stream = self._compute_streams[self._launched % len(self._compute_streams)]
        self._launched += 1
        with torch.cuda.stream(b.stream), torch.no_grad():
            b.stream.wait_event(b.ready)
            offsets = b.meta[0]
            lens = b.meta[1].to(torch.int32)
            b.res, b.res_len = self.model.forward_ragged(b.packed, offsets, lens, b.max_len, b.lens_host)
            # tokens: almost always fewer than one per stacked frame; the rare longer row is fetched in _finish
            b.cap = min(b.res.shape[1], max(32, (b.max_len + 1) // 2))
            b.tok_host = torch.empty((b.n_padded, b.cap), dtype=torch.int32, pin_memory=True)
            b.len_host = torch.empty((b.n_padded,), dtype=torch.int32, pin_memory=True)
            b.tok_host.copy_(b.res[:, :b.cap], non_blocking=True)
            b.len_host.copy_(b.res_len, non_blocking=True)
            # device-side watchdog word of this batch's kernels, copied behind them on the same stream (checked in _finish)
            b.status_host = torch.zeros(1, dtype=torch.int32).pin_memory()
            _lib.check(_lib.lib.rnnt_peek_device_status(_lib.ptr(b.status_host), _lib.stream_ptr()))
            b.done = torch.cuda.Event()
            b.done.record(b.stream)
        self._prev_done[b.keep[1]] = b.done

    def _finish(self, b):
        b.done.synchronize()
        self._raise_on_device_error(int(b.status_host[0]))
        lens = b.len_host.numpy()
        tokens = b.tok_host.numpy()
        self.h2d_bytes += b.h2d_bytes
        self.d2h_bytes += tokens.nbytes + lens.nbytes
        self.batches_run += 1
        longest = int(lens[:b.n_actual].max(initial=0))
        if longest > b.cap:
            # a few rows (a frame that ran into the 30-symbol cap) hold more tokens than the prefix copied back: fetch just
            # those rows, at the width they need, and widen the host array
            rows = np.nonzero(lens[:b.n_actual] > b.cap)[0]
            with torch.cuda.stream(b.stream):
                sel = torch.from_numpy(rows).to(b.res.device)
                extra = b.res.index_select(0, sel)[:, :longest].cpu().numpy()
            wide = np.full((b.n_actual, longest), -1, dtype=np.int32)
            wide[:, :b.cap] = tokens[:b.n_actual]
            wide[rows] = extra
            tokens = wide
            self.d2h_bytes += extra.nbytes
        self._respond(b.samples, tokens, lens)

    def _raise_on_device_error(self, code):
        """A kernel's bounded wait fired (protocol error): res / res_len of the batch are garbage and must not reach
        LoadGen as transcripts.  Clears the (sticky, per-device) word so that a caller who handles the error can go on."""
        if code != 0:
            _lib.check(_lib.lib.rnnt_clear_device_status(_lib.stream_ptr()))
            raise RuntimeError(f"librnnt_b200: device-side watchdog/protocol error {code} in this batch; no response sent")

    def _respond(self, samples, tokens, lens):
        """query_samples_complete (pytorch_sut.py:107-118) on host arrays: one QuerySampleResponse per sample (pointer into
        its row, n_tokens * 4 bytes); a chunk's responses are handed to LoadGen in one QuerySamplesComplete call."""
        base, stride = tokens.ctypes.data, tokens.strides[0]
        counts = lens[:len(samples)].tolist()
        if self.response_hook is not None:      # a front that answers LoadGen itself (sharding.DistributedOfflineSUT)
            self.response_hook(samples, tokens, counts)
            return
        lg.QuerySamplesComplete([lg.QuerySampleResponse(s.id, base + i * stride, n * 4)
                                 for i, (s, n) in enumerate(zip(samples, counts))])
        if logger.isEnabledFor(10):
            for i, s in enumerate(samples):
                logger.debug(f"{s.index}::{seq_to_sen(tokens[i], counts[i])}")

    # --------------------------------------------------------------------------------------------- synchronous path
    def inference(self, batch_idx):
        """pytorch_sut.py:67-105: returns (res {N_padded, L}, res_len {N_padded}) for the given sample indices."""
        batch_idx = list(batch_idx)
        self.actual_batch_size = len(batch_idx)
        if self.enable_process:
            return self._inference_from_waveforms(batch_idx)
        with torch.no_grad():
            feas = torch.nn.utils.rnn.pad_sequence([self.qsl[i][0] for i in batch_idx])          # {T, N, 240}
            fea_lens = torch.tensor([int(self.qsl[i][1]) for i in batch_idx], dtype=torch.int64)
            if self.actual_batch_size % 32 != 0:
                pad_n = math.ceil(self.actual_batch_size / 32) * 32 - self.actual_batch_size
                feas = torch.nn.functional.pad(feas, (0, 0, 0, pad_n, 0, 0), "constant", 0.0)
                fea_lens = torch.nn.functional.pad(fea_lens, (0, pad_n), "constant", 0)
            self.h2d_bytes += feas.numel() * 4 + fea_lens.numel() * 8
            results = self.model(feas, fea_lens)
            _lib.check_device_status()          # synchronises; raises instead of returning garbage transcripts
        self.batches_run += 1
        return results

    def _inference_from_waveforms(self, batch_idx):
        """pytorch_sut.py:70-82: waveforms -> features on the GPU (features.AudioProcessing) -> {T, N, C} -> decoder."""
        dev = self.model.device
        if next(self.processor.buffers()).device != dev:
            self.processor.to(dev)
        with torch.no_grad():
            wavs = torch.nn.utils.rnn.pad_sequence([self.qsl[i][0] for i in batch_idx], batch_first=True)
            wav_lens = torch.tensor([int(self.qsl[i][1]) for i in batch_idx])
            self.h2d_bytes += wavs.numel() * 4 + wav_lens.numel() * 8
            feas, fea_lens = self.processor(wavs.to(dev, non_blocking=True), wav_lens.to(torch.int32),
                                            pad_batch_size=(self.scenario == "Offline"))
            feas = feas.permute(2, 0, 1).contiguous()                                             # {N, C, T} -> {T, N, C}
            results = self.model(feas, fea_lens)
            _lib.check_device_status()
        self.batches_run += 1
        return results

    def query_samples_complete(self, samples, results, results_idx):
        lens = results_idx[:len(samples)].to(torch.int32).cpu().numpy()
        width = max(1, int(lens.max(initial=0)))
        tokens = np.ascontiguousarray(results[:len(samples), :width].to(torch.int32).cpu().numpy())
        self.d2h_bytes += tokens.nbytes + lens.nbytes
        self._respond(samples, tokens, lens)

    def __del__(self):
        try:
            lg.DestroySUT(self.sut)
        except Exception:
            pass
