"""Functional wrappers over the C ABI: torch CUDA tensors in, torch CUDA tensors out.

Plain row-major weights ([out_features, in_features]).  The reference-named operator surface (`_C.py`) and the
modules are built on these.  Every function runs on the current CUDA stream and does not synchronise.
"""
import ctypes

import torch

from . import _lib as L

H = 1024          # trans_hidden_size (models/config.py:4)
P = 320           # pred_hidden_size
J = 512           # joint_hidden_size
V_PAD = 32        # num_labels 29 padded to 32 (modeling_rnnt.py:241-250)


class Workspace:
    """Grow-only cache of device scratch buffers; the C ABI never allocates.  Keyed by (name, device, stream): two
    streams (the pipelined SUT's compute stream and a caller on the default stream, or two SUT instances) never share
    the recurrent kernel's flags / h ping-pong or the decode scratch."""

    def __init__(self):
        self._bufs = {}

    def get(self, name, nbytes, device):
        nbytes = int(nbytes)
        key = (name, str(device), torch.cuda.current_stream(device).cuda_stream)
        buf = self._bufs.get(key)
        if buf is None or buf.numel() < nbytes + 256:
            buf = torch.empty(max(nbytes, 256) + 256, dtype=torch.uint8, device=device)
            self._bufs[key] = buf
        off = (-buf.data_ptr()) % 256
        return buf[off:off + nbytes]


_ws = Workspace()


class Telemetry:
    """Launch counter and optional per-kernel CUDA-event timing (bench.py's roofline leg).  `launches` counts the
    kernels of librnnt_b200.so launched through this module.  With `events` set to a dict, the recurrent kernel and
    the projection GEMM are bracketed by CUDA events on the launching stream: events[name] = [(start, end, units)]."""

    def __init__(self):
        self.launches = 0
        self.events = None

    def bracket(self, name, units, fn):
        if self.events is None:
            return fn()
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record(torch.cuda.current_stream())
        out = fn()
        end.record(torch.cuda.current_stream())
        self.events.setdefault(name, []).append((start, end, units))
        return out


telemetry = Telemetry()


def quantize_pad(x, scale, k_out=None):
    """fp32 [..., k_in] -> int8 [..., k_out] (zero padded). quant_tpp.cpp:6-17 + lstm_amx.cpp:48-50."""
    L.require_cuda(x)
    if x.dtype != torch.float32:
        raise RuntimeError("quantize_pad expects float32 input")
    k_in = x.shape[-1]
    k_out = k_in if k_out is None else k_out
    out = torch.empty(x.shape[:-1] + (k_out,), dtype=torch.int8, device=x.device)
    rows = x.numel() // k_in
    telemetry.launches += 1
    L.check(L.lib.rnnt_quantize_pad_f32_i8(L.ptr(x), L.ptr(out), float(scale), rows, k_in, k_out, L.stream_ptr()))
    return out


def quantize_ragged(packed, offsets, lens, scale, t_max, k_out):
    """Ragged batch -> padded time-major int8 {t_max, N, k_out} (pad_sequence + quantise + K pad in one pass)."""
    L.require_cuda(packed, offsets, lens)
    if packed.dtype != torch.float32 or offsets.dtype != torch.int64 or lens.dtype != torch.int32:
        raise RuntimeError("quantize_ragged expects float32 features, int64 offsets, int32 lens")
    n = lens.numel()
    k_in = packed.shape[-1]
    out = torch.empty((t_max, n, k_out), dtype=torch.int8, device=packed.device)
    telemetry.launches += 1
    L.check(L.lib.rnnt_quantize_ragged_f32_i8(L.ptr(packed), L.ptr(offsets), L.ptr(lens), L.ptr(out), float(scale), int(t_max), n,
                                              k_in, int(k_out), L.stream_ptr()))
    return out


def gemm_i8_i32(a, w, out=None, simt=False):
    """int8 [..., K] x int8 [N, K]^T -> int32 [..., N], exact."""
    L.require_cuda(a, w)
    k = a.shape[-1]
    m = a.numel() // k
    n = w.shape[0]
    if w.shape[1] != k:
        raise RuntimeError(f"gemm_i8_i32: K mismatch {k} vs {w.shape[1]}")
    if out is None:
        out = torch.empty(a.shape[:-1] + (n,), dtype=torch.int32, device=a.device)
    fn = L.lib.rnnt_gemm_i8_i32_simt if simt else L.lib.rnnt_gemm_i8_i32
    L.check(fn(L.ptr(a), L.ptr(w), L.ptr(out), m, n, k, L.stream_ptr()))
    return out


def tile_steps_for(lens_host, n, divisor=1):
    """Host int32 array for rnnt_lstm_recurrent_i8_ragged: per 128-row batch tile, max(ceil(len / divisor)) of its rows."""
    import numpy as np
    lens = np.zeros((n + 127) // 128 * 128, dtype=np.int64)
    lens[:len(lens_host)] = np.asarray(lens_host, dtype=np.int64)
    steps = (lens.reshape(-1, 128).max(axis=1) + divisor - 1) // divisor
    return np.ascontiguousarray(steps.astype(np.int32))


def lstm_layer_i8(x, hx, cx, w_ih, w_hh, bias, rb_scale, in_scale, out_scale, skip_quant_y, tile_steps=None, zero_dead=True,
                  true_rows=None):
    """lstm_layer_amx_int8 (lstm_amx.cpp:59-158). x int8 [T,N,K]; hx int8 [N,H], cx fp16 [N,H] updated IN PLACE.
    Returns y: int8 [T,N,H] or bf16 [T,N,H] when skip_quant_y.  tile_steps (host int32 per 128-row tile, see
    tile_steps_for): ragged batch sorted by length — (t, tile) items past a tile's last step are skipped and y is zero
    there (zero_dead=False leaves those rows unwritten: for callers whose consumers skip the same items, it saves the
    fill of the whole output)."""
    L.require_cuda(x, hx, cx, w_ih, w_hh, bias)
    t, n, k = x.shape
    if hx.dtype != torch.int8 or cx.dtype != torch.float16 or x.dtype != torch.int8:
        raise RuntimeError("lstm_layer_i8: x/hx must be int8 and cx float16")
    alloc = torch.zeros if (tile_steps is not None and zero_dead) else torch.empty
    y = alloc((t, n, H), dtype=torch.bfloat16 if skip_quant_y else torch.int8, device=x.device)
    if t == 0:
        return y
    g_ws = _ws.get("lstm_g", t * n * 4 * H * 4, x.device)
    ws = _ws.get("lstm_rec", L.lib.rnnt_lstm_recurrent_workspace_bytes(n), x.device)
    st = L.stream_ptr()
    ts_ptr = None
    if tile_steps is not None:
        if tile_steps.dtype.name != "int32" or len(tile_steps) != (n + 127) // 128:
            raise RuntimeError("lstm_layer_i8: tile_steps must be int32 with ceil(N/128) entries")
        ts_ptr = ctypes.c_void_p(tile_steps.ctypes.data)
    # phase A: one projection GEMM over all timesteps; phase B: one persistent recurrent kernel, issued separately so
    # that bench.py can time each kernel with CUDA events.  The projection scratch between them is gate-blocked
    # ([32 slices][4 gates][4 unit octets][T*N rows][8] int32): the rows a warp of the recurrent kernel reads are
    # contiguous there, 8 lines per load instead of 32 of the row-major [T*N][4096] result.
    gemm_ts = ts_ptr                               # dead (t, tile) rows of the projection are skipped as well
    telemetry.bracket("gemm_i8", (t * n, k), lambda: L.check(
        L.lib.rnnt_gemm_i8_i32_gate_blocked_ragged(L.ptr(x), L.ptr(w_ih), L.ptr(g_ws), t * n, 4 * H, k, n, gemm_ts, st)))
    # telemetry units of the recurrent kernel: (TRUE (frame, utterance) pairs the layer has to process — tile padding and
    # rows past an utterance's end are not algorithmic work —, (timestep, 128-row tile) items the kernel runs, bf16 output)
    items = t * ((n + 127) // 128) if tile_steps is None else int(tile_steps.sum())
    units = (int(true_rows) if true_rows is not None else t * n, items, bool(skip_quant_y))
    telemetry.bracket("lstm_recurrent", units, lambda: L.check(
        L.lib.rnnt_lstm_recurrent_i8_gate_blocked_ragged(L.ptr(g_ws), L.ptr(w_hh), L.ptr(bias), float(rb_scale), float(in_scale),
                                            float(out_scale), int(bool(skip_quant_y)), L.ptr(hx), L.ptr(cx), L.ptr(y), t, n, H,
                                            ts_ptr, L.ptr(ws), st)))
    telemetry.launches += 2
    return y


def lstm_recurrent_i8(g, hx, cx, w_hh, bias, rb_scale, in_scale, out_scale, skip_quant_y):
    """Phase B only (lstm_amx.cpp:122-152) on a precomputed int32 projection g [T,N,4H]."""
    L.require_cuda(g, hx, cx, w_hh, bias)
    t, n, _ = g.shape
    y = torch.empty((t, n, H), dtype=torch.bfloat16 if skip_quant_y else torch.int8, device=g.device)
    ws = _ws.get("lstm_rec", L.lib.rnnt_lstm_recurrent_workspace_bytes(n), g.device)
    L.check(L.lib.rnnt_lstm_recurrent_i8(L.ptr(g), L.ptr(w_hh), L.ptr(bias), float(rb_scale), float(in_scale),
                                         float(out_scale), int(bool(skip_quant_y)), L.ptr(hx), L.ptr(cx), L.ptr(y), t, n, H,
                                         L.ptr(ws), L.stream_ptr()))
    return y


def stack_time(x, lens, factor):
    """stack_time.cpp:27-59 — int8 {T,N,C} -> {ceil(T/f),N,C*f} with the per-sample zero tail."""
    L.require_cuda(x, lens)
    if x.dtype != torch.int8:
        raise RuntimeError("stack_time: int8 only")
    if lens.dtype != torch.int32:
        lens = lens.to(torch.int32)
    t, n, c = x.shape
    t2 = (t + factor - 1) // factor
    out = torch.empty((t2, n, c * factor), dtype=torch.int8, device=x.device)
    telemetry.launches += 1
    L.check(L.lib.rnnt_stack_time_i8(L.ptr(x), L.ptr(lens), int(factor), L.ptr(out), t, n, c, L.stream_ptr()))
    return out


def linear_bf16_bf16(x, w, tile_steps=None):
    """y = bf16(x . w^T) with the oracle's accumulation order (amx_linear.cpp:137-150, i16o16b0).  tile_steps (host int32
    per 128-row tile of a time-major {T, N, K} input): blocks of rows past a tile's last step are not computed."""
    L.require_cuda(x, w)
    if x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16:
        raise RuntimeError("linear_bf16_bf16 expects bfloat16")
    k = x.shape[-1]
    m = x.numel() // k
    o = w.shape[0]
    y = torch.empty(x.shape[:-1] + (o,), dtype=torch.bfloat16, device=x.device)
    telemetry.launches += 1
    ts_ptr, rows = None, 0
    if tile_steps is not None and x.dim() == 3:
        ts_ptr, rows = ctypes.c_void_p(tile_steps.ctypes.data), x.shape[1]
    telemetry.bracket("linear_bf16", (m, k, o), lambda: L.check(
        L.lib.rnnt_linear_bf16_bf16_ragged(L.ptr(x), L.ptr(w), L.ptr(y), m, k, o, rows, ts_ptr, L.stream_ptr())))
    return y


def gemm_bf16_bf16(x, w, tile_steps=None):
    """y = bf16(x . w^T) on the tensor cores (tcgen05 kind::f16, fp32 accumulation in TMEM): the hoisted transcription
    half of the joint fc1 (amx_linear.cpp:137-150).  Agrees with the order-defined linear_bf16_bf16 to <= 1 bf16 ulp.
    tile_steps: as in linear_bf16_bf16 (time-major {T, N, K} input)."""
    L.require_cuda(x, w)
    if x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16:
        raise RuntimeError("gemm_bf16_bf16 expects bfloat16")
    k = x.shape[-1]
    m = x.numel() // k
    o = w.shape[0]
    if w.shape[1] != k:
        raise RuntimeError(f"gemm_bf16_bf16: K mismatch {k} vs {w.shape[1]}")
    y = torch.empty(x.shape[:-1] + (o,), dtype=torch.bfloat16, device=x.device)
    telemetry.launches += 1
    ts_ptr, rows = None, 0
    if tile_steps is not None and x.dim() == 3:
        ts_ptr, rows = ctypes.c_void_p(tile_steps.ctypes.data), x.shape[1]
    telemetry.bracket("gemm_bf16", (m, k, o), lambda: L.check(
        L.lib.rnnt_gemm_bf16_bf16_ragged(L.ptr(x), L.ptr(w), L.ptr(y), m, o, k, rows, ts_ptr, L.stream_ptr())))
    return y


def lstm_postop(it, ft, gt, ot, ct, in_scale, out_scale, skip_quant_y):
    """lstm_postop (init.cpp:126-128; lstm_postop.cpp:14-72): fp32 gate chunks [N, H] (may be the four column chunks of
    one [N, 4H] tensor), ct fp16 [N, H] updated IN PLACE.  Returns [y_bf16, y_q, h_q, ct] like the reference (the y
    output the flag does not select is zero-filled; the reference leaves it uninitialised)."""
    L.require_cuda(ct)
    for g in (it, ft, gt, ot):
        if not g.is_cuda or g.dtype != torch.float32 or g.dim() != 2 or g.stride(1) != 1 or g.stride(0) != it.stride(0):
            raise RuntimeError("lstm_postop: gates must be fp32 CUDA [N, H] tensors with unit column stride and equal row pitch")
    if ct.dtype != torch.float16:
        raise RuntimeError("lstm_postop: ct must be float16")
    n, h = it.shape
    y_bf16 = torch.zeros((n, h), dtype=torch.bfloat16, device=it.device)
    y_q = torch.zeros((n, h), dtype=torch.int8, device=it.device)
    h_q = torch.empty((n, h), dtype=torch.int8, device=it.device)
    telemetry.launches += 1
    L.check(L.lib.rnnt_lstm_postop_fp16(L.ptr(it), L.ptr(ft), L.ptr(gt), L.ptr(ot), it.stride(0), L.ptr(ct), float(in_scale),
                                        float(out_scale), int(bool(skip_quant_y)), L.ptr(y_bf16), L.ptr(y_q), L.ptr(h_q), n, h,
                                        L.stream_ptr()))
    return [y_bf16, y_q, h_q, ct]


def make_decode_weights(embed, w_ih, w_hh, bias, w_pred, b1, w2, b2):
    """Pack device pointers into rnnt_decode_weights; keeps the tensors alive on the returned object."""
    tensors = [embed, *w_ih, *w_hh, *bias, w_pred, b1, w2, b2]
    L.require_cuda(*tensors)
    if tuple(w2.shape) != (V_PAD, J) or tuple(b2.shape) != (V_PAD,):
        raise RuntimeError("decode weights: fc2 must be padded to 32 rows (modeling_rnnt.py:241-250)")
    dw = L.DecodeWeights()
    dw.embed = embed.data_ptr()
    for i in range(2):
        dw.w_ih[i] = w_ih[i].data_ptr()
        dw.w_hh[i] = w_hh[i].data_ptr()
        dw.bias[i] = bias[i].data_ptr()
    dw.w_pred, dw.b1, dw.w2, dw.b2 = w_pred.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr()
    dw._keep = tensors
    return dw


def greedy_decode(y1, f_lens, dw, res, res_idx, symbols_added, time_idx, pre_g, pre_hg, pre_cg, iters_out=None,
                  exact_order=False):
    """Device-resident decode loop (decoder.py:188-211). State tensors are updated in place.  exact_order=False: the
    contractions run on the tensor cores; True: CUDA-core chains in the CPU oracle's accumulation order (bit-exact
    cross-checks)."""
    L.require_cuda(y1, f_lens, res, res_idx, symbols_added, time_idx, pre_g, *pre_hg, *pre_cg)
    t2, n, j = y1.shape
    assert j == J and y1.dtype == torch.bfloat16
    for t in (f_lens, res, res_idx, symbols_added, time_idx, pre_g):
        if t.dtype != torch.int32:
            raise RuntimeError("greedy_decode: integer state must be int32")
    ws = _ws.get("decode", L.lib.rnnt_greedy_decode_workspace_bytes(n), y1.device)
    st = L.stream_ptr()
    telemetry.launches += 1
    fn = L.lib.rnnt_greedy_decode_exact_order if exact_order else L.lib.rnnt_greedy_decode
    telemetry.bracket("greedy_decode", (t2, n), lambda: L.check(fn(
        L.ptr(y1), L.ptr(f_lens), t2, n, ctypes.byref(dw), L.ptr(res), res.shape[1], L.ptr(res_idx), L.ptr(symbols_added),
        L.ptr(time_idx), L.ptr(pre_g), L.ptr(pre_hg[0]), L.ptr(pre_hg[1]), L.ptr(pre_cg[0]), L.ptr(pre_cg[1]), L.ptr(iters_out),
        L.ptr(ws), st)))

# ---------------------------------------------------------------------------------------------------------
# per-iteration operators of the reference decode loop (same arithmetic as the fused kernel)
def linear_bf16_f32(x, w, bias=None):
    """amx_linear_i16o32 (amx_linear.cpp:113-117): fp32 [..., O] = x . w^T + bias."""
    L.require_cuda(x, w, bias)
    if x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16:
        raise RuntimeError("linear_bf16_f32 expects bfloat16 x / w")
    k = x.shape[-1]
    m = x.numel() // k
    o = w.shape[0]
    y = torch.empty(x.shape[:-1] + (o,), dtype=torch.float32, device=x.device)
    L.check(L.lib.rnnt_linear_bf16_f32(L.ptr(x), L.ptr(w), L.ptr(bias), L.ptr(y), m, k, o, L.stream_ptr()))
    return y


def linear_bf16_accum_relu(y1, x1, w1, bias):
    """second half of amx_linear_bf16_accum_relu (amx_linear.cpp:152-168); y1 = linear_bf16_bf16(x0, w0)."""
    L.require_cuda(y1, x1, w1, bias)
    for t in (y1, x1, w1):
        if t.dtype != torch.bfloat16:
            raise RuntimeError("linear_bf16_accum_relu expects bfloat16")
    k = x1.shape[-1]
    m = x1.numel() // k
    o = w1.shape[0]
    if y1.numel() != m * o:
        raise RuntimeError("linear_bf16_accum_relu: y1 shape mismatch")
    y = torch.empty(x1.shape[:-1] + (o,), dtype=torch.bfloat16, device=x1.device)
    L.check(L.lib.rnnt_linear_bf16_accum_relu(L.ptr(y1), L.ptr(x1), L.ptr(w1), L.ptr(bias), L.ptr(y), m, k, o, L.stream_ptr()))
    return y


def dequant_i32_f32(acc, bias, scale):
    """(float(acc) + bias) * scale — the i8o32b32 epilogue (amx_linear.cpp:18-92)."""
    L.require_cuda(acc, bias)
    o = acc.shape[-1]
    out = torch.empty(acc.shape, dtype=torch.float32, device=acc.device)
    L.check(L.lib.rnnt_dequant_i32_f32(L.ptr(acc), L.ptr(bias), float(scale), L.ptr(out), acc.numel() // o, o, L.stream_ptr()))
    return out


def lstm_cell_bf16(x, h, c, w_ih, w_hh, bias):
    """lstm_layer_amx_bf16 at seq_len 1 (lstm_amx.cpp:187-264). Returns NEW (hy bf16, cy fp32)."""
    L.require_cuda(x, h, c, w_ih, w_hh, bias)
    n, k = x.shape
    p = h.shape[1]
    hy = torch.empty((n, p), dtype=torch.bfloat16, device=x.device)
    cy = torch.empty((n, p), dtype=torch.float32, device=x.device)
    ws = _ws.get("cell_gates", 2 * n * 4 * p * 4, x.device)
    L.check(L.lib.rnnt_lstm_cell_bf16(L.ptr(x), L.ptr(h), L.ptr(c), L.ptr(w_ih), L.ptr(w_hh), L.ptr(bias), L.ptr(hy), L.ptr(cy),
                                      n, k, p, L.ptr(ws), L.stream_ptr()))
    return hy, cy


def greedy_decode_update(symbols, symbols_added, res, res_idx, f, f_lens, time_idx, fi, pre_g, pre_hg, pre_cg, hg, cg):
    """greedy_decode_update.cpp:13-96; in place; returns True when all lanes finished."""
    L.require_cuda(symbols, symbols_added, res, res_idx, f, f_lens, time_idx, fi, pre_g, *pre_hg, *pre_cg, *hg, *cg)
    if symbols.dtype != torch.int64:
        raise RuntimeError("greedy_decode_update: symbols must be int64 (greedy_decode_update.cpp:25)")
    for t in (symbols_added, res, res_idx, f_lens, time_idx, pre_g):
        if t.dtype != torch.int32:
            raise RuntimeError("greedy_decode_update: counters must be int32 (greedy_decode_update.cpp:26-43)")
    n, l = res.shape
    c = f.shape[2]
    p = hg[0].shape[1]
    flag = _ws.get("update_flag", 256, res.device)
    L.check(L.lib.rnnt_greedy_decode_update(L.ptr(symbols), L.ptr(symbols_added), L.ptr(res), L.ptr(res_idx), L.ptr(f),
                                            L.ptr(f_lens), L.ptr(time_idx), L.ptr(fi), L.ptr(pre_g), L.ptr(pre_hg[0]),
                                            L.ptr(pre_hg[1]), L.ptr(pre_cg[0]), L.ptr(pre_cg[1]), L.ptr(hg[0]), L.ptr(hg[1]),
                                            L.ptr(cg[0]), L.ptr(cg[1]), n, l, c, p, L.ptr(flag), L.stream_ptr()))
    return int(flag[:4].view(torch.int32).item()) == 0


# ------------------------------------------------------------------------------------------------ audio front-end
def _i32(lens, device):
    return lens.to(device=device, dtype=torch.int32).contiguous()


def preemphasis(x, lens, coeff=0.97, pad_size=0):
    """preemphasis.cpp:10-46: x fp32 [N, L] -> [N, L + 2 pad] (y = x[i] - coeff x[i-1], reflect padding, zeros behind)."""
    L.require_cuda(x)
    x = x.contiguous()
    n, l = x.shape
    out = torch.empty((n, l + 2 * int(pad_size)), dtype=torch.float32, device=x.device)
    L.check(L.lib.rnnt_preemphasis_f32(L.ptr(x), L.ptr(_i32(lens, x.device)), float(coeff), int(pad_size), L.ptr(out), n, l,
                                       L.stream_ptr()))
    return out


def power_spectrum(x, lens):
    """power_spectrum.cpp:10-33: x fp32 [N, T, F, 2] -> [N, T, F]."""
    L.require_cuda(x)
    x = x.contiguous()
    n, t, f, _ = x.shape
    out = torch.empty((n, t, f), dtype=torch.float32, device=x.device)
    L.check(L.lib.rnnt_power_spectrum_f32(L.ptr(x), L.ptr(_i32(lens, x.device)), L.ptr(out), n, t, f, L.stream_ptr()))
    return out


def frame_splicing(x, lens, factor):
    """frame_splicing.cpp:12-45: x fp32 [N, C, T] -> [N, C * factor, ceil(T / factor)]."""
    L.require_cuda(x)
    x = x.contiguous()
    n, c, t = x.shape
    factor = int(factor)
    out = torch.empty((n, c * factor, (t + factor - 1) // factor), dtype=torch.float32, device=x.device)
    L.check(L.lib.rnnt_frame_splicing_f32(L.ptr(x), L.ptr(_i32(lens, x.device)), L.ptr(out), n, c, t, factor, L.stream_ptr()))
    return out


def layernorm_pad(x, lens, eps=1e-12, unbiased=True, n_pad=None, c_out=None, t_out=None, weight=None, bias=None):
    """normalization.cpp:10-64 on fp32 [N, C, T]: out [n_pad, c_out, t_out] (t_out defaults to max(lens))."""
    L.require_cuda(x)
    x = x.contiguous()
    n, c, t = x.shape
    lens = _i32(lens, x.device)
    n_pad = n if n_pad is None else int(n_pad)
    c_out = c if c_out is None else int(c_out)
    t_out = int(lens.max().item()) if t_out is None else int(t_out)
    out = torch.empty((n_pad, c_out, max(t_out, 1)), dtype=torch.float32, device=x.device)
    L.check(L.lib.rnnt_layernorm_pad_f32(L.ptr(x), L.ptr(lens), L.ptr(weight) if weight is not None else None,
                                         L.ptr(bias) if bias is not None else None, L.ptr(out), n, c, t, n_pad, c_out, max(t_out, 1),
                                         float(eps), int(bool(unbiased)), L.stream_ptr()))
    return out


def logmel(y, frame_lens, window, fb, n_fft, hop, dither2, log_bias):
    """STFT + power spectrum + dither^2 + mel + log of features.py:201-227 in one kernel.  y fp32 [N, Lp] (pre-emphasised and
    padded), frame_lens [N], window [win_length], fb [n_mels, n_fft / 2 + 1] -> [N, n_mels, T], T = 1 + (Lp - n_fft) // hop."""
    L.require_cuda(y, window, fb)
    y = y.contiguous()
    n, lp = y.shape
    t = 1 + (lp - n_fft) // hop
    n_mels = fb.shape[0]
    out = torch.empty((n, n_mels, t), dtype=torch.float32, device=y.device)
    L.check(L.lib.rnnt_logmel_f32(L.ptr(y), L.ptr(_i32(frame_lens, y.device)), L.ptr(window.contiguous()), L.ptr(fb.contiguous()),
                                  L.ptr(out), n, lp, t, int(n_fft), int(hop), int(window.numel()), int(n_mels), float(dither2),
                                  float(log_bias), L.stream_ptr()))
    telemetry.launches += 1
    return out
