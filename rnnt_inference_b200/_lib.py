"""ctypes binding of librnnt_b200.so (C ABI in include/rnnt_b200.h).

The library is the product's only compute path: if it is missing this module raises at import time —
there is no CPU or PyTorch fallback.  Tensors are passed as raw device pointers; PyTorch only owns the
memory and the stream.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# RNNT_B200_LIB selects another build of the same ABI (profiles/trace_*.py load lib/librnnt_b200_dbg.so, `make debug`)
LIB_PATH = os.environ.get("RNNT_B200_LIB") or os.path.join(_HERE, "lib", "librnnt_b200.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(nvcc, sm_100a). rnnt_inference_b200 has no fallback path.")

lib = ctypes.CDLL(LIB_PATH)

_vp, _i, _i64, _f = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float


class DecodeWeights(ctypes.Structure):
    """rnnt_decode_weights in include/rnnt_b200.h"""
    _fields_ = [("embed", _vp), ("w_ih", _vp * 2), ("w_hh", _vp * 2), ("bias", _vp * 2), ("w_pred", _vp), ("b1", _vp),
                ("w2", _vp), ("b2", _vp)]


_SIGS = {
    "rnnt_last_error": (ctypes.c_char_p, []),
    "rnnt_abi_version": (_i, []),
    "rnnt_device_sm_count": (_i, []),
    "rnnt_check_device_status": (_i, [_vp]),
    "rnnt_peek_device_status": (_i, [_vp, _vp]),
    "rnnt_clear_device_status": (_i, [_vp]),
    "rnnt_quantize_pad_f32_i8": (_i, [_vp, _vp, _f, _i64, _i, _i, _vp]),
    "rnnt_quantize_ragged_f32_i8": (_i, [_vp, _vp, _vp, _vp, _f, _i, _i, _i, _i, _vp]),
    "rnnt_gemm_i8_i32": (_i, [_vp, _vp, _vp, _i64, _i, _i, _vp]),
    "rnnt_gemm_i8_i32_ragged": (_i, [_vp, _vp, _vp, _i64, _i, _i, _i, _vp, _vp]),
    "rnnt_gemm_i8_i32_gate_blocked_ragged": (_i, [_vp, _vp, _vp, _i64, _i, _i, _i, _vp, _vp]),
    "rnnt_linear_bf16_bf16_ragged": (_i, [_vp, _vp, _vp, _i64, _i, _i, _i, _vp, _vp]),
    "rnnt_gemm_i8_i32_simt": (_i, [_vp, _vp, _vp, _i64, _i, _i, _vp]),
    "rnnt_lstm_recurrent_workspace_bytes": (_i64, [_i]),
    "rnnt_lstm_recurrent_i8": (_i, [_vp, _vp, _vp, _f, _f, _f, _i, _vp, _vp, _vp, _i, _i, _i, _vp, _vp]),
    "rnnt_lstm_recurrent_i8_static_scratch": (_i, [_vp, _vp, _vp, _f, _f, _f, _i, _vp, _vp, _vp, _i, _i, _i, _vp, _vp]),
    "rnnt_lstm_recurrent_i8_ragged": (_i, [_vp, _vp, _vp, _f, _f, _f, _i, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp]),
    "rnnt_lstm_recurrent_i8_gate_blocked_ragged": (_i, [_vp, _vp, _vp, _f, _f, _f, _i, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp]),
    "rnnt_tune_recurrent_cluster": (_i, [_i]),
    "rnnt_preemphasis_f32": (_i, [_vp, _vp, _f, _i, _vp, _i, _i, _vp]),
    "rnnt_power_spectrum_f32": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "rnnt_frame_splicing_f32": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "rnnt_layernorm_pad_f32": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _f, _i, _vp]),
    "rnnt_logmel_f32": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _f, _f, _vp]),
    "rnnt_tune_gemm_cluster": (_i, [_i]),
    "rnnt_tune_gemm_pair": (_i, [_i]),
    "rnnt_lstm_layer_i8": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _f, _f, _f, _i, _vp, _i, _i, _i, _i, _vp, _vp, _vp]),
    "rnnt_stack_time_i8": (_i, [_vp, _vp, _i, _vp, _i, _i, _i, _vp]),
    "rnnt_linear_bf16_bf16": (_i, [_vp, _vp, _vp, _i64, _i, _i, _vp]),
    "rnnt_gemm_bf16_bf16": (_i, [_vp, _vp, _vp, _i64, _i, _i, _vp]),
    "rnnt_gemm_bf16_bf16_ragged": (_i, [_vp, _vp, _vp, _i64, _i, _i, _i, _vp, _vp]),
    "rnnt_lstm_postop_fp16": (_i, [_vp, _vp, _vp, _vp, _i64, _vp, _f, _f, _i, _vp, _vp, _vp, _i, _i, _vp]),
    "rnnt_linear_bf16_f32": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _i, _vp]),
    "rnnt_linear_bf16_accum_relu": (_i, [_vp, _vp, _vp, _vp, _vp, _i64, _i, _i, _vp]),
    "rnnt_dequant_i32_f32": (_i, [_vp, _vp, _f, _vp, _i64, _i, _vp]),
    "rnnt_lstm_cell_bf16": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp]),
    "rnnt_greedy_decode_update": (_i, [_vp] * 17 + [_i, _i, _i, _i, _vp, _vp]),
    "rnnt_greedy_decode_workspace_bytes": (_i64, [_i]),
    "rnnt_greedy_decode": (_i, [_vp, _vp, _i, _i, ctypes.POINTER(DecodeWeights), _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp,
                                _vp, _vp, _vp, _vp, _vp]),
    "rnnt_greedy_decode_exact_order": (_i, [_vp, _vp, _i, _i, ctypes.POINTER(DecodeWeights), _vp, _i, _vp, _vp, _vp, _vp, _vp,
                                            _vp, _vp, _vp, _vp, _vp, _vp]),
}
EXPORTS = tuple(_SIGS)
# profiling hooks of the debug build (include/rnnt_b200_debug.h); absent from the product library
_DEBUG_SIGS = {
    "rnnt_debug_set_recurrent_trace": (_i, [_vp]),
    "rnnt_debug_set_recurrent_flags": (_i, [_i]),
    "rnnt_debug_set_decode_trace": (_i, [_vp, _i]),
    "rnnt_debug_set_decode_trace_block": (_i, [_i]),
    "rnnt_debug_set_decode_skew": (_i, [_i]),
    "rnnt_debug_set_gemm_flags": (_i, [_i]),
}
for _name, (_res, _args) in _SIGS.items():
    _fn = getattr(lib, _name)      # AttributeError here = the .so does not match the header
    _fn.restype = _res
    _fn.argtypes = _args
HAS_DEBUG_HOOKS = hasattr(lib, "rnnt_debug_set_recurrent_trace")
if HAS_DEBUG_HOOKS:
    for _name, (_res, _args) in _DEBUG_SIGS.items():
        _fn = getattr(lib, _name)
        _fn.restype = _res
        _fn.argtypes = _args


def check(rc):
    if rc != 0:
        raise RuntimeError(f"librnnt_b200: {lib.rnnt_last_error().decode()} (code {rc})")


def ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def stream_ptr(device=None):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("librnnt_b200 operators need CUDA tensors (no CPU path exists)")
        if t is not None and not t.is_contiguous():
            raise RuntimeError("Input should be contiguous.")


def check_device_status():
    """Synchronise the current stream and raise if a device-side watchdog fired."""
    check(lib.rnnt_check_device_status(stream_ptr()))
