"""Reference-named operator surface (stands in for models/_C.py:15-51, i.e. torch.ops.intel_mlperf.*).

Same names, argument order and in-place/ownership behaviour as the schemas in
mlperf_plugins/csrc/init.cpp:20-144, on CUDA tensors, backed by librnnt_b200.so (C ABI, include/rnnt_b200.h).
Differences a caller can observe:
  * tensors must live on a CUDA device (there is no CPU path; a CPU tensor raises RuntimeError);
  * weights are plain K-major [out_features, in_features] (see quant_modules.transpose_tile_weight*), not the
    AMX VNNI tiles;
  * `greedy_decode` is a new fused entry point: the whole prediction/joint/argmax/update loop of
    decoder.py:188-211 in one device-resident kernel.  The per-iteration ops of that loop (`lstm_amx_bf16`,
    `amx_linear_bf16_accum_relu`, `amx_linear_i16o32`, `greedy_decode_update`) are also provided so that the
    reference's Python loop runs unmodified and can be cross-checked against the fused kernel.
"""
from typing import List, Tuple

import torch
from torch import Tensor

from . import ops


def _scalar(v, i=None):
    if isinstance(v, Tensor):
        return float(v[i].item()) if i is not None else float(v.item())
    return float(v if i is None else v[i])


def lstm_layer_amx_int8(x: Tensor, hx: Tensor, cx: Tensor, w_ih: Tensor, w_hh: Tensor, b_ih, b_hh: Tensor, rb_scale,
                        i_scale, o_scale, skip_quant_y: bool) -> Tuple[Tensor, Tensor, Tensor]:
    """init.cpp:129-131 / lstm_amx.cpp:59-158.  hx, cx are mutated in place and returned."""
    y = ops.lstm_layer_i8(x, hx, cx, w_ih, w_hh, b_hh, _scalar(rb_scale), _scalar(i_scale), _scalar(o_scale),
                          bool(skip_quant_y))
    return y, hx, cx


def lstm_amx_int8(x: Tensor, hx: List[Tensor], cx: List[Tensor], all_weights: List[List[Tensor]], rb_scale, i_scale,
                  o_scale, skip_quant_y: bool, tile_steps=None, true_rows=None) -> Tuple[Tensor, List[Tensor], List[Tensor]]:
    """init.cpp:132-134 / lstm_amx.cpp:24-57: quantise fp32 input with i_scale[0], zero-pad K to a multiple of 64,
    then run the layers; only the last layer honours skip_quant_y.  tile_steps (extension, default off): see
    ops.lstm_layer_i8 — ragged batches sorted by length skip the dead (timestep, 128-row tile) items; their rows of
    the result are left unwritten (every consumer on the path — the next layer's projection, stack_time, the hoisted
    joint product, the decode loop — skips them by the same lengths)."""
    num_layers = len(all_weights)
    if x.dtype == torch.float32:
        k = x.shape[-1]
        x = ops.quantize_pad(x.contiguous(), _scalar(i_scale, 0), (k + 63) // 64 * 64)
    elif x.dtype == torch.int8 and x.shape[-1] % 64 != 0:
        x = torch.nn.functional.pad(x, (0, 64 - x.shape[-1] % 64))
    for layer in range(num_layers):
        w_ih, w_hh, _b_ih, b_hh = all_weights[layer]
        last = layer == num_layers - 1
        x = ops.lstm_layer_i8(x, hx[layer], cx[layer], w_ih, w_hh, b_hh, _scalar(rb_scale, layer), _scalar(i_scale, layer),
                              _scalar(o_scale, layer), bool(skip_quant_y) and last, tile_steps=tile_steps, zero_dead=False,
                              true_rows=true_rows)
    return x, hx, cx


def stack_time(x: Tensor, x_lens: Tensor, factor) -> Tensor:
    """init.cpp:96-98 / stack_time.cpp:27-59."""
    return ops.stack_time(x, x_lens, int(factor))


def amx_linear_i8o32(x: Tensor, weight: Tensor, bias: Tensor, scale) -> Tensor:
    """init.cpp:30-32 / amx_linear.cpp:18-92 (i8o32b32): y_f32 = (float(x.W^T) + bias) * scale.  The integer
    contraction runs on the tcgen05 GEMM; the dequant is two separately rounded fp32 ops like the reference."""
    acc = ops.gemm_i8_i32(x, weight)
    return ops.dequant_i32_f32(acc, bias, _scalar(scale))


def linear(x: Tensor, weight: Tensor, bias: Tensor, scale, zero=None) -> Tensor:
    """P.linear as used by iLinear (quant_linear.py:62-64): int8 x int8 -> fp32 with scale (oneDNN int8 inner
    product in the reference, asserted equal to amx_linear_i8o32 by test_amx_linear.py:118-126)."""
    return amx_linear_i8o32(x, weight, bias, scale)


def lstm_postop(it: Tensor, ft: Tensor, gt: Tensor, ot: Tensor, ct: Tensor, i_scale, o_scale, skip_quant_y: bool):
    """init.cpp:126-128 / lstm_postop.cpp:14-72: the fp16 gate math of one LSTM step as a stand-alone operator; ct is
    mutated in place.  Returns [y_bf16, y_q, h_q, ct]."""
    return ops.lstm_postop(it, ft, gt, ot, ct, _scalar(i_scale), _scalar(o_scale), bool(skip_quant_y))


def amx_linear_bf16_accum_relu(x0: Tensor, w0: Tensor, x1: Tensor, w1: Tensor, bias: Tensor) -> Tensor:
    """init.cpp:33-35 / amx_linear.cpp:119-170: bf16(relu((x1.W1^T + float(bf16(x0.W0^T))) + bias))."""
    y1 = ops.linear_bf16_bf16(x0, w0)
    return ops.linear_bf16_accum_relu(y1, x1, w1, bias)


def amx_linear_i16o32(x: Tensor, weight: Tensor, bias: Tensor) -> Tensor:
    """init.cpp:36-38 / amx_linear.cpp:113-117: fp32 = x.W^T + bias (weight rows padded to 32)."""
    return ops.linear_bf16_f32(x, weight, bias)


def lstm_amx_bf16(x: Tensor, hx: List[Tensor], cx: List[Tensor], all_weights: List[List[Tensor]]):
    """init.cpp:138-140 / lstm_amx.cpp:160-185.  x {U=1, N, P} bf16.  Returns NEW hy/cy lists (inputs untouched)."""
    if x.shape[0] != 1:
        raise RuntimeError("lstm_amx_bf16: only seq_len == 1 is supported (the decode loop's shape)")
    hy, cy = [], []
    cur = x[0]
    for layer, (w_ih, w_hh, _b_ih, b_sum) in enumerate(all_weights):
        h, c = ops.lstm_cell_bf16(cur, hx[layer], cx[layer], w_ih, w_hh, b_sum)
        hy.append(h)
        cy.append(c)
        cur = h
    return cur.unsqueeze(0), hy, cy


def greedy_decode_update(symbols: Tensor, symbols_added: Tensor, res: Tensor, res_idx: Tensor, f: Tensor, f_lens: Tensor,
                         time_idx: Tensor, fi: Tensor, pre_g: Tensor, pre_hg: List[Tensor], pre_cg: List[Tensor],
                         hg: List[Tensor], cg: List[Tensor]) -> bool:
    """init.cpp:141-143 / greedy_decode_update.cpp:13-96.  Mutates its state arguments in place; returns True when
    every lane has time_idx >= f_lens (one 4-byte device->host read, like the reference's bool return)."""
    return ops.greedy_decode_update(symbols, symbols_added, res, res_idx, f, f_lens, time_idx, fi, pre_g, pre_hg, pre_cg,
                                    hg, cg)


def prepack_linear_weight(weight: Tensor) -> Tensor:
    """Layout-only in the reference (linear.cpp:377-449, oneDNN blocked format); identity here."""
    return weight


def prepack_lstm_weights(w_ih: Tensor, w_hh: Tensor):
    """Layout-only in the reference (lstm.cpp:89-185); identity here."""
    return w_ih, w_hh


# ---- fused entry points (new; used by GreedyDecoder / Transcription of this package) ------------------------
linear_bf16_bf16 = ops.linear_bf16_bf16
gemm_bf16_bf16 = ops.gemm_bf16_bf16
make_decode_weights = ops.make_decode_weights
greedy_decode = ops.greedy_decode
quantize_pad = ops.quantize_pad


def lstm(x: Tensor, hx: List[Tensor], cx: List[Tensor], all_weights: List[List[Tensor]]):
    """init.cpp:105-107 / lstm.cpp:187-307 (oneDNN fp32 lstm_forward in the reference; its source is not vendored).
    fp32 configuration only (BASELINE config C1, run_mode f32/calib): textbook LSTM in plain torch, gate order
    i,f,g,o, on whatever device the tensors live.  Returns NEW hy/cy lists."""
    hy, cy = [], []
    for layer, (w_ih, w_hh, b_ih, b_hh) in enumerate(all_weights):
        h, c = hx[layer], cx[layer]
        steps = []
        for t in range(x.shape[0]):
            gates = torch.nn.functional.linear(x[t], w_ih, b_ih) + torch.nn.functional.linear(h, w_hh, b_hh)
            i_g, f_g, g_g, o_g = gates.chunk(4, 1)
            c = torch.sigmoid(f_g) * c + torch.sigmoid(i_g) * torch.tanh(g_g)
            h = torch.sigmoid(o_g) * torch.tanh(c)
            steps.append(h)
        x = torch.stack(steps, 0)
        hy.append(h)
        cy.append(c)
    return x, hy, cy
quantize_ragged = ops.quantize_ragged
tile_steps_for = ops.tile_steps_for


# ------------------------------------------------------------------------------------------------ audio front-end operators
def preemphasis(input: Tensor, length: Tensor, coeff=0.97, pad_size=0) -> Tensor:
    """intel_mlperf::preemphasis (preemphasis.cpp:10-46)."""
    return ops.preemphasis(input, length, coeff, pad_size)


def power_spectrum(input: Tensor, length: Tensor) -> Tensor:
    """intel_mlperf::power_spectrum (power_spectrum.cpp:10-33)."""
    return ops.power_spectrum(input, length)


def frame_splicing(input: Tensor, length: Tensor, factor) -> Tensor:
    """intel_mlperf::frame_splicing (frame_splicing.cpp:12-45); like the reference, factor 3 only."""
    if int(factor) != 3:
        raise RuntimeError("Unsupported frame splicing factor.")
    return ops.frame_splicing(input, length, 3)


def i_layernorm_pad(input: Tensor, weight, bias, length: Tensor, eps=1e-12, unbiased=False, output_shape=None):
    """intel_mlperf::i_layernorm_pad (normalization.cpp:10-64): returns (output [padded_batch, out_feat, max(length)], length
    padded with zeros to padded_batch).  weight / bias [1, C, T'] or None."""
    n, c, _t = input.shape
    n_pad, c_out = n, c
    out_len = length
    if output_shape is not None:
        n_pad, c_out = int(output_shape[0]), int(output_shape[1])
        if n_pad == 1:
            n_pad = n
        else:
            out_len = torch.nn.functional.pad(length, (0, n_pad - n), "constant", 0)
    w = b = None
    if weight is not None and not bool((weight == 1).all()):
        w = weight[0, :c, :input.shape[2]].to(input.device).contiguous()
    if bias is not None and not bool((bias == 0).all()):
        b = bias[0, :c, :input.shape[2]].to(input.device).contiguous()
    return ops.layernorm_pad(input, length, eps, unbiased, n_pad, c_out, None, w, b), out_len
