"""Server-scenario SUT: continuous (slot) batching on one GPU.

Reference: csrc/torch_sut.cpp:354-571 (ServerSUT producer / consumer threads, QoS parking of long samples, early
response), csrc/metadata.cpp:97-194 (PipelineState::update / next: finished slots are refilled, unfinished ones carry
their encoder and prediction state), csrc/rnnt_model.hpp:63-124 (encode runs `split_len`-frame chunks until at least
`response_size` slots have finished, then ONE decode over the frames encoded in this iteration).

The same scheme on the B200 kernels — nothing new on the device, only per-slot state management on the host:
  * `batch_size` slots; per slot the features [T_cap, 240] (device), the int8 / fp16 encoder state of the five layers,
    the prediction state (pre_g, hg, cg) and a token row; a slot that is (re)filled has all of it reset;
  * one iteration = refill finished slots from the queue -> chunks of `split_len` frames (every slot at its own
    offset, gathered with one indexing op) through Transcription with carried state until `response_size` more slots
    have finished -> hoisted f.W_trans + ONE fused greedy decode over the stacked frames of the iteration (its time
    index restarts at 0, state carried — decoder.py:80-91 semantics) -> responses for the finished slots;
  * samples longer than `length_threshold` frames are parked in a QoS queue and only served when the main queue is
    empty (torch_sut.cpp:384-414), so that they do not hold short samples' latency hostage.
LoadGen calls issue_queries from its own thread (one sample at a time in Server mode); a worker thread runs the
iterations; flush_queries waits until everything issued has been answered.
"""
import collections
import threading

import numpy as np
import torch

try:
    import mlperf_loadgen as lg
except ImportError:   # not installed offline
    from . import loadgen_shim as lg

from . import _C as P
from . import _lib
from .config import RNNTParam
from .modeling_rnnt import RNNT
from .rnnt_qsl import RNNTQSL


class ServerSUT:
    def __init__(self, model_path, dataset_dir, batch_size=128, split_len=8, response_size=8, length_threshold=0,
                 exact_order=False, device=None):
        if split_len <= 0 or split_len % RNNTParam.stack_time_factor:
            raise RuntimeError("ServerSUT: split_len must be a positive multiple of the stack-time factor (2)")
        self.rnnt = RNNT(model_path, "quant", True, device=device).eval()
        self.device = next(self.rnnt.parameters()).device
        self.qsl = RNNTQSL(dataset_dir)
        self.slots = int(batch_size)
        self.split_len = int(split_len)
        self.response_size = max(1, int(response_size))
        self.length_threshold = int(length_threshold)
        self.exact_order = exact_order
        self.sut = lg.ConstructSUT(self.issue_queries, self.flush_queries)
        longest = max(self.qsl.x_lens_host)
        self.t_cap = (longest + self.split_len - 1) // self.split_len * self.split_len       # metadata.cpp:99-102
        self.k_in = int(self.qsl[0][0].shape[1])
        self._alloc()
        self._queue, self._qos = collections.deque(), collections.deque()
        self._cv = threading.Condition()
        self._stop = False
        self._busy = False
        self.iterations = 0
        self._stream = torch.cuda.Stream(self.device)
        self._worker = threading.Thread(target=self._run, daemon=True)
        self._worker.start()

    # ------------------------------------------------------------------------------------------------- state
    def _alloc(self):
        dev, b = self.device, self.slots
        hid, pred = RNNTParam.trans_hidden_size, RNNTParam.pred_hidden_size
        z = lambda shape, dt: torch.zeros(shape, dtype=dt, device=dev)
        self.F = z((self.t_cap, b, self.k_in), torch.float32)
        self.pre_hx = [z((b, hid), torch.int8) for _ in range(RNNTParam.pre_num_layers)]
        self.pre_cx = [z((b, hid), torch.float16) for _ in range(RNNTParam.pre_num_layers)]
        self.post_hx = [z((b, hid), torch.int8) for _ in range(RNNTParam.post_num_layers)]
        self.post_cx = [z((b, hid), torch.float16) for _ in range(RNNTParam.post_num_layers)]
        self.pre_g = torch.full((b,), RNNTParam.SOS, dtype=torch.int32, device=dev)
        self.pre_hg = [z((b, pred), torch.bfloat16) for _ in range(RNNTParam.pred_num_layers)]
        self.pre_cg = [z((b, pred), torch.float32) for _ in range(RNNTParam.pred_num_layers)]
        self.res = torch.full((b, RNNTParam.max_symbols_per_step * self.t_cap // 2), -1, dtype=torch.int32, device=dev)
        self.res_idx = torch.full((b,), -1, dtype=torch.int32, device=dev)
        self.sample = [None] * b            # QuerySample held by each slot
        self.f_len = [0] * b                # frames of the slot's utterance
        self.remain = [0] * b               # frames not encoded yet

    def _reset_slots(self, idx):
        """metadata.cpp:111-139: zero the encoder / prediction state and the token cursor of refilled slots."""
        sel = torch.tensor(idx, dtype=torch.int64, device=self.device)
        for t in (*self.pre_hx, *self.pre_cx, *self.post_hx, *self.post_cx, *self.pre_hg, *self.pre_cg):
            t.index_fill_(0, sel, 0)
        self.pre_g.index_fill_(0, sel, RNNTParam.SOS)
        self.res_idx.index_fill_(0, sel, -1)

    # ------------------------------------------------------------------------------------------------- LoadGen entry
    def issue_queries(self, samples):
        with self._cv:
            for s in samples:
                long = self.length_threshold > 0 and self.qsl.x_lens_host[s.index] > self.length_threshold
                (self._qos if long else self._queue).append(s)
            self._cv.notify()

    def flush_queries(self):
        with self._cv:
            while self._queue or self._qos or self._busy or any(r > 0 for r in self.remain):
                self._cv.wait(0.01)

    def close(self):
        with self._cv:
            self._stop = True
            self._cv.notify()
        self._worker.join(timeout=10)

    # ------------------------------------------------------------------------------------------------- worker
    def _take(self, n):
        """Up to n samples: the main queue first, parked long samples only when it is empty (torch_sut.cpp:384-414)."""
        out = []
        while len(out) < n and self._queue:
            out.append(self._queue.popleft())
        while len(out) < n and not self._queue and self._qos:
            out.append(self._qos.popleft())
        return out

    def _run(self):
        torch.cuda.set_device(self.device)
        while True:
            with self._cv:
                while not self._stop and not (self._queue or self._qos or any(r > 0 for r in self.remain)):
                    self._busy = False
                    self._cv.notify_all()
                    self._cv.wait(0.05)
                if self._stop:
                    return
                self._busy = True
                free = [i for i in range(self.slots) if self.remain[i] == 0]
                new = self._take(len(free))
            with torch.cuda.stream(self._stream), torch.no_grad():
                self._iteration(free[:len(new)], new)
            self.iterations += 1

    def _iteration(self, slots, new):
        dev, b, sl = self.device, self.slots, self.split_len
        # ---- refill (PipelineState::update)
        if new:
            for i, s in zip(slots, new):
                x, n = self.qsl[s.index]
                n = int(n)
                self.F[:n, i].copy_(x[:n], non_blocking=True)
                self.sample[i], self.f_len[i], self.remain[i] = s, n, n
            self._reset_slots(slots)
        active = [i for i in range(b) if self.remain[i] > 0]
        if not active:
            return
        # ---- chunked encode until `response_size` more slots have finished (rnnt_model.hpp:63-78, PipelineState::next)
        target = min(len(active), self.response_size)
        infer = [0] * b
        chunks, finished = [], 0
        ar = torch.arange(sl, device=dev).unsqueeze(1)
        lanes = torch.arange(b, device=dev).unsqueeze(0)
        while finished < target:
            begin = [self.f_len[i] - self.remain[i] if self.remain[i] > 0 else 0 for i in range(b)]
            clen = [min(sl, self.remain[i]) for i in range(b)]
            idx = (torch.tensor(begin, device=dev).unsqueeze(0) + ar).clamp_(max=self.t_cap - 1)
            x = self.F[idx, lanes]                                           # {split_len, B, 240}: every slot at its own offset
            lens = torch.tensor(clen, dtype=torch.int32, device=dev)
            f, *_ = self.rnnt.transcription(x, lens, self.pre_hx, self.pre_cx, self.post_hx, self.post_cx)
            chunks.append(f)
            for i in range(b):
                self.remain[i] -= clen[i]
                infer[i] += clen[i]
            finished = sum(1 for i in active if self.remain[i] == 0)
        f = torch.cat(chunks, 0) if len(chunks) > 1 else chunks[0]
        fac = RNNTParam.stack_time_factor
        f_lens = torch.tensor([(v + fac - 1) // fac for v in infer], dtype=torch.int32, device=dev)
        # ---- one fused decode over the frames of this iteration, prediction state carried (rnnt_model.hpp:80-124)
        y1 = self.rnnt.joint.hoist_trans(f, None, self.exact_order)
        symbols_added = torch.zeros(b, dtype=torch.int32, device=dev)
        time_idx = torch.zeros(b, dtype=torch.int32, device=dev)
        P.greedy_decode(y1, f_lens, self.rnnt.decode_weights(), self.res, self.res_idx, symbols_added, time_idx, self.pre_g,
                        self.pre_hg, self.pre_cg, None, exact_order=self.exact_order)
        # ---- early response for the slots that finished (ServerSUT::QuerySamplesComplete)
        done = [i for i in active if self.remain[i] == 0]
        sel = torch.tensor(done, dtype=torch.int64, device=dev)
        counts = (self.res_idx.index_select(0, sel) + 1).cpu().numpy()
        width = max(1, int(counts.max(initial=0)))
        tokens = np.ascontiguousarray(self.res.index_select(0, sel)[:, :width].cpu().numpy())
        _lib.check_device_status()
        base, stride = tokens.ctypes.data, tokens.strides[0]
        lg.QuerySamplesComplete([lg.QuerySampleResponse(self.sample[i].id, base + k * stride, int(counts[k]) * 4)
                                 for k, i in enumerate(done)])
        for i in done:
            self.sample[i] = None

    def __del__(self):
        try:
            self.close()
            lg.DestroySUT(self.sut)
        except Exception:
            pass
