"""Greedy decoder (reference: models/decoder.py).

GreedyDecoder(model, run_mode, enable_bf16, split_len, batch_size).forward(x {T,N,C}, x_lens {N}) -> (res {N,L},
res_len {N}) with L = 30 * max(x_lens) (decoder.py:30-35) and max_symbols_per_step = 30.

run_mode "quant": encoder on the int8 kernels, then ONE launch of the device-resident decode kernel
(`fused=True`, default) or the reference's per-iteration loop of four operators (`fused=False`; same
arithmetic, kept for cross-checks).  The fp32 modes run the torch-mask algorithm of decoder.py:102-169.
"""
import torch
from torch import Tensor

from . import _C as P
from .config import RNNTParam


class GreedyDecoder(torch.nn.Module):
    def __init__(self, model, run_mode, enable_bf16=False, split_len=-1, batch_size=1, fused=True, exact_order=False,
                 keep_intermediates=False):
        """fused / exact_order / keep_intermediates are extensions (quant mode): fused=False runs the reference's
        per-iteration operator loop; exact_order=True computes every bf16 contraction (the hoisted f.W_trans and the
        decode loop's) on CUDA cores in the CPU oracle's accumulation order instead of on the tensor cores (bit-exact
        cross-checks); keep_intermediates leaves the encoder
        output and the hoisted product of the last call in self.last_f / self.last_y1 (tests, reports)."""
        super().__init__()
        self.rnnt = model
        self.split_len = split_len
        self.enable_bf16 = enable_bf16
        self.run_mode = run_mode
        self.fused = fused
        self.exact_order = exact_order
        self.keep_intermediates = keep_intermediates
        self.last_f = self.last_y1 = None
        self.padded_batch_size = batch_size   # the AMX thread-count padding of decoder.py:18-19 has no GPU meaning
        self.last_iters = None                # device int32[1]: lock-step iterations of the last fused decode

    @property
    def device(self):
        return next(self.rnnt.parameters()).device

    def _alloc_state(self, n, max_len, dev):
        quant = self.run_mode == "quant"
        res_dtype = torch.int32 if quant else torch.int64
        hx_dtype = torch.int8 if quant else torch.float32
        cx_dtype = torch.float16 if quant else torch.float32
        hid, pred = RNNTParam.trans_hidden_size, RNNTParam.pred_hidden_size
        z = lambda shape, dt: torch.zeros(shape, dtype=dt, device=dev)
        self.res = torch.full((n, RNNTParam.max_symbols_per_step * max_len), -1, dtype=res_dtype, device=dev)
        self.res_idx = torch.full((n,), -1, dtype=res_dtype, device=dev)
        self.pre_hx = [z((n, hid), hx_dtype) for _ in range(RNNTParam.pre_num_layers)]
        self.pre_cx = [z((n, hid), cx_dtype) for _ in range(RNNTParam.pre_num_layers)]
        self.post_hx = [z((n, hid), hx_dtype) for _ in range(RNNTParam.post_num_layers)]
        self.post_cx = [z((n, hid), cx_dtype) for _ in range(RNNTParam.post_num_layers)]
        self.pre_g = torch.full((1, n), RNNTParam.SOS, dtype=res_dtype, device=dev)
        self.pre_hg = [z((n, pred), torch.bfloat16 if self.enable_bf16 else torch.float32)
                       for _ in range(RNNTParam.pred_num_layers)]
        self.pre_cg = [z((n, pred), torch.float32) for _ in range(RNNTParam.pred_num_layers)]

    def forward(self, x: Tensor, x_lens: Tensor):
        """x {T, N, C} fp32 features (host or device), x_lens {N}.  Returns (res, res_len) on the model's device."""
        dev = self.device
        max_len = int(x_lens.max().item())          # read on the host side before the copy
        x = x.to(dev, non_blocking=True)
        x_lens = x_lens.to(dev, non_blocking=True)
        self.batch_size = x_lens.size(0)
        self._alloc_state(self.batch_size, max_len, dev)
        if self.split_len != -1:
            # chunked encoder with carried (h, c) and decode state (decoder.py:80-91)
            for start in range(0, max_len, self.split_len):
                xi_lens = torch.clamp(x_lens - start, min=0, max=self.split_len).to(x_lens.dtype)
                self.greedy_decode(x[start:start + self.split_len], xi_lens)
        else:
            self.greedy_decode(x, x_lens)
        return self.res, self.res_idx + 1

    def forward_ragged(self, packed: Tensor, offsets: Tensor, lens: Tensor, max_len: int, lens_host=None):
        """SUT fast path (quant mode): the batch as the unpadded concatenation of its utterances, all on the device.
        packed {sum T_i, 240} fp32, offsets {N} int64, lens {N} int32 (0 for batch-padding lanes).  Produces the same
        int8 {T, N, 256} tensor that forward() derives from the padded fp32 batch, without materialising that batch.
        lens_host (host copy of lens, rows sorted by length as PytorchSUT.issue_queries does) lets the encoder skip the
        (timestep, 128-row tile) items past a tile's longest utterance; transcripts are unchanged."""
        if self.run_mode != "quant" or self.split_len != -1:
            raise RuntimeError("forward_ragged serves run_mode='quant' without split_len")
        dev = packed.device
        self.batch_size = lens.numel()
        self._alloc_state(self.batch_size, max_len, dev)
        pre = self.rnnt.transcription.pre_rnn
        k = packed.shape[-1]
        x = P.quantize_ragged(packed, offsets, lens, float(pre.in_scale[0].item()), max_len, (k + 63) // 64 * 64)
        self.greedy_decode_quant(x, lens, lens_host)
        return self.res, self.res_idx + 1

    def greedy_decode(self, f: Tensor, f_lens: Tensor):
        if self.run_mode in ("f32", "calib", None):
            return self.greedy_decode_f32(f, f_lens)
        if self.run_mode == "quant":
            return self.greedy_decode_quant(f, f_lens)
        return self.greedy_decode_f32(f, f_lens)     # fake_quant: torch modules, fp32 state

    # ------------------------------------------------------------------------------------------- int8 / bf16
    def greedy_decode_quant(self, f: Tensor, f_lens: Tensor, lens_host=None):
        n = self.batch_size
        dev = f.device
        self.symbols_added = torch.zeros(n, dtype=torch.int32, device=dev)
        self.time_idx = torch.zeros(n, dtype=torch.int32, device=dev)
        f_lens = f_lens.to(torch.int32)
        f, self.pre_hx, self.pre_cx, self.post_hx, self.post_cx = self.rnnt.transcription(
            f, f_lens, self.pre_hx, self.pre_cx, self.post_hx, self.post_cx, lens_host)
        fac = RNNTParam.stack_time_factor
        f_lens = torch.div(f_lens + (fac - 1), fac, rounding_mode="floor").to(torch.int32)
        if self.fused:
            steps2 = None if lens_host is None else P.tile_steps_for(lens_host, n, fac)
            y1 = self.rnnt.joint.hoist_trans(f, steps2, self.exact_order)
            if self.keep_intermediates:
                self.last_f, self.last_y1 = f, y1
            if self.last_iters is None or self.last_iters.device != dev:
                self.last_iters = torch.zeros(1, dtype=torch.int32, device=dev)
            P.greedy_decode(y1, f_lens, self.rnnt.decode_weights(), self.res, self.res_idx, self.symbols_added, self.time_idx,
                            self.pre_g.view(-1), self.pre_hg, self.pre_cg, self.last_iters, exact_order=self.exact_order)
            return self.res
        # reference loop, one operator per step (decoder.py:186-211); fi aliases f[0] like the reference
        fi = f[0]
        while True:
            g, hg, cg = self.rnnt.prediction(self.pre_g, self.pre_hg, self.pre_cg)
            y = self.rnnt.joint(fi, g[0])
            symbols = torch.argmax(y, dim=1)
            if self.rnnt.update(symbols, self.symbols_added, self.res, self.res_idx, f, f_lens, self.time_idx, fi,
                                self.pre_g.view(-1), self.pre_hg, self.pre_cg, hg, cg):
                break
        return self.res

    # ------------------------------------------------------------------------------------------- fp32
    def greedy_decode_f32(self, f: Tensor, f_lens: Tensor):
        """decoder.py:102-169: boolean-mask formulation; finished lanes are masked out (unlike the quant path)."""
        n = self.batch_size
        dev = f.device
        blank, cap = RNNTParam.BLANK, RNNTParam.max_symbols_per_step
        self.symbols_added = torch.zeros(n, dtype=torch.int64, device=dev)
        self.time_idx = torch.zeros(n, dtype=torch.int64, device=dev)
        self.finish = f_lens.eq(0)
        lanes = torch.arange(n, dtype=torch.int64, device=dev)
        f, self.pre_hx, self.pre_cx, self.post_hx, self.post_cx = self.rnnt.transcription(
            f, f_lens, self.pre_hx, self.pre_cx, self.post_hx, self.post_cx)
        f_lens = torch.ceil(f_lens / RNNTParam.stack_time_factor).to(torch.int32)
        self.eos_idx = torch.clamp(f_lens - 1, min=0)
        if self.enable_bf16:
            f = f.to(torch.bfloat16)
        fi = f[0]
        while True:
            g, hg, cg = self.rnnt.prediction(self.pre_g, self.pre_hg, self.pre_cg)
            y = self.rnnt.joint(fi, g[0], self.padded_batch_size)
            symbols = torch.argmax(y, dim=1)
            emit = symbols.ne(blank) & self.symbols_added.ne(cap) & ~self.finish
            if bool(emit.any()):
                self.res_idx += emit
                self.res[emit, self.res_idx[emit]] = symbols[emit]
                self.symbols_added += emit
                self.pre_g[0][emit] = symbols[emit]
                for layer in range(RNNTParam.pred_num_layers):
                    self.pre_hg[layer][emit, :] = hg[layer][emit, :]
                    self.pre_cg[layer][emit, :] = cg[layer][emit, :]
            advance = ~emit & ~self.finish
            if bool(advance.any()):
                self.time_idx += advance
                self.finish |= self.time_idx.ge(f_lens)
                self.time_idx = self.time_idx.min(self.eos_idx)
                if bool(self.finish.all()):
                    break
                fi = f[self.time_idx, lanes, :]
                self.symbols_added *= ~advance
        return self.res
