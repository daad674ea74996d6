"""ctypes front-end of the CPU oracle (oracle/rnnt_oracle.cpp).  TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module; the product package rnnt_inference_b200 never does.

All functions take/return numpy arrays.  binary16 / bfloat16 values travel as uint16 bit patterns
(helpers `f32_to_bf16_bits`, `bf16_bits_to_f32` below).  Weights are plain row-major
[out_features, in_features] like torch.nn.LSTM / Linear.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "librnnt_oracle.so")


def build(force=False):
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(
            os.path.join(_HERE, "rnnt_oracle.cpp")):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.orc_greedy_decode.restype = ctypes.c_int
        _lib.orc_greedy_decode_trace.restype = ctypes.c_int
        _lib.orc_greedy_update.restype = ctypes.c_int
        _lib.orc_selfcheck.restype = ctypes.c_int
        assert _lib.orc_selfcheck() == 0
    return _lib


def set_num_threads(n):
    """OpenMP threads used by the oracle's batch-parallel loops; returns the team size now in effect."""
    lib().orc_set_num_threads.restype = ctypes.c_int
    return int(lib().orc_set_num_threads(int(n)))


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _c(a, dtype):
    a = np.ascontiguousarray(a, dtype=dtype)
    return a


def f32_to_bf16_bits(x):
    """RNE fp32 -> bf16 bit patterns (same rule as the oracle's f2bf)."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    u = u + 0x7FFF + ((u >> 16) & 1)
    return (u >> 16).astype(np.uint16)


def bf16_bits_to_f32(b):
    return (np.ascontiguousarray(b, dtype=np.uint16).astype(np.uint32) << 16).view(np.float32)


def f16_bits_to_f32(b):
    return np.ascontiguousarray(b, dtype=np.uint16).view(np.float16).astype(np.float32)


def quantize_pad(x, scale, k_out):
    x = _c(x, np.float32)
    rows, k_in = int(np.prod(x.shape[:-1])), x.shape[-1]
    out = np.empty(x.shape[:-1] + (k_out,), dtype=np.int8)
    lib().orc_quantize_pad(_p(x), _p(out), ctypes.c_float(scale), ctypes.c_int64(rows), k_in, k_out)
    return out


def gemm_i8_i32(a, w):
    a, w = _c(a, np.int8), _c(w, np.int8)
    m, k = int(np.prod(a.shape[:-1])), a.shape[-1]
    n = w.shape[0]
    assert w.shape[1] == k
    c = np.empty(a.shape[:-1] + (n,), dtype=np.int32)
    lib().orc_gemm_i8_i32(_p(a), _p(w), _p(c), ctypes.c_int64(m), n, k)
    return c


def lstm_postop_fp16(gates, c, in_scale, out_scale, last):
    """gates [N,4H] f32, c [N,H] uint16 (binary16 bits, copied). Returns (y, hq, c_new)."""
    gates = _c(gates, np.float32)
    c = np.array(c, dtype=np.uint16, copy=True)
    n, h = c.shape
    hq = np.empty((n, h), np.int8)
    yq = np.empty((n, h), np.int8)
    ybf = np.empty((n, h), np.uint16)
    lib().orc_lstm_postop_fp16(_p(gates), _p(c), _p(hq), _p(yq), _p(ybf), ctypes.c_float(in_scale),
                               ctypes.c_float(out_scale), n, h, int(bool(last)))
    return (ybf if last else yq), hq, c


def lstm_layer_i8(x, hx, cx, w_ih, w_hh, bias, rb_scale, in_scale, out_scale, skip_quant_y, dump_gates=False):
    """x [T,N,K] int8; hx [N,H] int8, cx [N,H] uint16 — both UPDATED IN PLACE (like the reference)."""
    x = _c(x, np.int8)
    t, n, k = x.shape
    h = hx.shape[1]
    assert hx.dtype == np.int8 and cx.dtype == np.uint16 and hx.flags.c_contiguous and cx.flags.c_contiguous
    w_ih, w_hh, bias = _c(w_ih, np.int8), _c(w_hh, np.int8), _c(bias, np.float32)
    assert w_ih.shape == (4 * h, k) and w_hh.shape == (4 * h, h)
    y = np.empty((t, n, h), np.uint16 if skip_quant_y else np.int8)
    g = np.empty((t, n, 4 * h), np.float32) if dump_gates else None
    lib().orc_lstm_layer_i8(_p(x), _p(hx), _p(cx), _p(w_ih), _p(w_hh), _p(bias), ctypes.c_float(rb_scale),
                            ctypes.c_float(in_scale), ctypes.c_float(out_scale), int(bool(skip_quant_y)), _p(y),
                            t, n, k, h, _p(g))
    return (y, g) if dump_gates else y


def stack_time(x, lens, factor=2):
    x = _c(x, np.int8)
    lens = _c(lens, np.int32)
    t, n, c = x.shape
    t2 = (t + factor - 1) // factor
    out = np.empty((t2, n, c * factor), np.int8)
    lib().orc_stack_time(_p(x), _p(lens), factor, _p(out), t, n, c)
    return out


def lstm_cell_bf16(x, h, c, w_ih, w_hh, bias):
    x, h, c = _c(x, np.uint16), _c(h, np.uint16), _c(c, np.float32)
    w_ih, w_hh, bias = _c(w_ih, np.uint16), _c(w_hh, np.uint16), _c(bias, np.float32)
    n, k = x.shape
    hh = h.shape[1]
    hy, cy = np.empty((n, hh), np.uint16), np.empty((n, hh), np.float32)
    lib().orc_lstm_cell_bf16(_p(x), _p(h), _p(c), _p(w_ih), _p(w_hh), _p(bias), _p(hy), _p(cy), n, k, hh)
    return hy, cy


def linear_bf16_bf16(x, w):
    x, w = _c(x, np.uint16), _c(w, np.uint16)
    n, k = int(np.prod(x.shape[:-1])), x.shape[-1]
    o = w.shape[0]
    y = np.empty(x.shape[:-1] + (o,), np.uint16)
    lib().orc_linear_bf16_bf16(_p(x), _p(w), _p(y), ctypes.c_int64(n), k, o)
    return y


def linear_bf16_accum_relu(x0, w0, x1, w1, bias, y1_pre=None):
    x1, w1, bias = _c(x1, np.uint16), _c(w1, np.uint16), _c(bias, np.float32)
    n, k1 = x1.shape
    o = w1.shape[0]
    if y1_pre is None:
        x0, w0 = _c(x0, np.uint16), _c(w0, np.uint16)
        k0 = x0.shape[1]
    else:
        y1_pre = _c(y1_pre, np.uint16)
        k0 = 0
    y = np.empty((n, o), np.uint16)
    lib().orc_linear_bf16_accum_relu(_p(x0) if y1_pre is None else None, _p(w0) if y1_pre is None else None,
                                     _p(y1_pre), _p(x1), _p(w1), _p(bias), _p(y), n, k0, k1, o)
    return y


def linear_bf16_f32(x, w, bias):
    x, w, bias = _c(x, np.uint16), _c(w, np.uint16), _c(bias, np.float32)
    n, k = x.shape
    o = w.shape[0]
    y = np.empty((n, o), np.float32)
    lib().orc_linear_bf16_f32(_p(x), _p(w), _p(bias), _p(y), n, k, o)
    return y


def greedy_update(symbols, symbols_added, res, res_idx, f, f_lens, time_idx, fi, pre_g, pre_hg, pre_cg, hg, cg):
    """In-place twin of P.greedy_decode_update.  All arrays numpy, contiguous, reference dtypes."""
    n, l = res.shape
    c = f.shape[2]
    p = hg[0].shape[1]
    return bool(lib().orc_greedy_update(_p(_c(symbols, np.int64)), _p(symbols_added), _p(res), _p(res_idx), _p(f),
                                        _p(f_lens), _p(time_idx), _p(fi), _p(pre_g), _p(pre_hg[0]), _p(pre_hg[1]),
                                        _p(pre_cg[0]), _p(pre_cg[1]), _p(hg[0]), _p(hg[1]), _p(cg[0]), _p(cg[1]),
                                        n, l, c, p))


def greedy_decode(f, f_lens, params, res, res_idx, symbols_added, time_idx, pre_g, pre_hg, pre_cg, y1_pre=None,
                  max_iters=1 << 30, trace=None):
    """Batch-lockstep decode loop on an encoded batch; state arrays are updated in place.

    params: dict with embed [28,P], pred_w_ih/pred_w_hh (lists of 2 [4P,P]), pred_bias (list of 2 [4P]),
    w_trans [J,C], w_pred [J,P], b1 [J], w2 [32,J], b2 [32]  (bf16 bit patterns / fp32).
    """
    f = _c(f, np.uint16)
    t2, n, c = f.shape
    p = params["embed"].shape[1]
    j = params["w_pred"].shape[0]
    keep = [_c(params["embed"], np.uint16)]
    arr2 = ctypes.c_void_p * 2
    wih = [_c(w, np.uint16) for w in params["pred_w_ih"]]
    whh = [_c(w, np.uint16) for w in params["pred_w_hh"]]
    bs = [_c(b, np.float32) for b in params["pred_bias"]]
    wt, wp, b1 = _c(params["w_trans"], np.uint16), _c(params["w_pred"], np.uint16), _c(params["b1"], np.float32)
    w2, b2 = _c(params["w2"], np.uint16), _c(params["b2"], np.float32)
    assert w2.shape == (32, j) and b2.shape == (32,)
    y1 = None if y1_pre is None else _c(y1_pre, np.uint16)
    for a, dt in ((res, np.int32), (res_idx, np.int32), (symbols_added, np.int32), (time_idx, np.int32),
                  (pre_g, np.int32), (pre_hg[0], np.uint16), (pre_hg[1], np.uint16), (pre_cg[0], np.float32),
                  (pre_cg[1], np.float32)):
        assert a.dtype == dt and a.flags.c_contiguous
    f_lens = _c(f_lens, np.int32)
    args = [_p(f), _p(f_lens), t2, n, c, p, j, _p(keep[0]), arr2(*[w.ctypes.data for w in wih]),
            arr2(*[w.ctypes.data for w in whh]), arr2(*[b.ctypes.data for b in bs]), _p(wt), _p(wp), _p(b1), _p(w2), _p(b2),
            _p(y1), _p(res), res.shape[1], _p(res_idx), _p(symbols_added), _p(time_idx), _p(pre_g), _p(pre_hg[0]),
            _p(pre_hg[1]), _p(pre_cg[0]), _p(pre_cg[1]), int(max_iters)]
    if trace is None:
        return lib().orc_greedy_decode(*args)
    # trace: dict(sym int8 [N, M], margin fp32 [N, M], count int32 [N]) filled with every lane's decisions (argmax and
    # top-2 logit margin) while the lane is unfinished
    m = trace["sym"].shape[1]
    assert trace["sym"].dtype == np.int8 and trace["margin"].dtype == np.float32 and trace["count"].dtype == np.int32
    return lib().orc_greedy_decode_trace(*args, _p(trace["sym"]), _p(trace["margin"]), _p(trace["count"]), m)


def new_decision_trace(n, m):
    return dict(sym=np.full((n, m), -1, np.int8), margin=np.zeros((n, m), np.float32), count=np.zeros(n, np.int32))


def elementwise(name, *arrays, out_dtype, scale=None):
    """Thin access to orc_{sigmoid,tanh,rcp,fma,quant}_* for the hardware-golden tests."""
    arrays = [np.ascontiguousarray(a) for a in arrays]
    out = np.empty(arrays[0].shape, out_dtype)
    fn = getattr(lib(), "orc_" + name)
    if scale is not None:
        fn(_p(arrays[0]), ctypes.c_float(scale), _p(out), ctypes.c_int64(out.size))
    else:
        fn(*[_p(a) for a in arrays], _p(out), ctypes.c_int64(out.size))
    return out
