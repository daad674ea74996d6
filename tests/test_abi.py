"""The C-ABI boundary without a GPU: librnnt_b200.so loads, exports every function include/rnnt_b200.h declares, binds
with the signatures of _lib.py, rejects bad arguments with an error code + message (no exception, no crash), and the
product build carries none of the profiling hooks.  No compute call is made."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rnnt_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from rnnt_inference_b200 import _lib
    names = _declared("rnnt_b200.h")
    assert len(names) >= 30
    for name in names:
        assert hasattr(_lib.lib, name), f"{name} is declared in include/rnnt_b200.h but not exported"
    assert set(_lib.EXPORTS) <= set(names), set(_lib.EXPORTS) - set(names)
    assert _lib.lib.rnnt_abi_version() == 1


def test_product_library_has_no_debug_hooks():
    from rnnt_inference_b200 import _lib
    for name in _declared("rnnt_b200_debug.h"):
        assert name.startswith("rnnt_debug_")
        assert not hasattr(_lib.lib, name), f"{name} must only exist in the debug build (make debug)"
    assert not _lib.HAS_DEBUG_HOOKS


def test_bad_arguments_return_codes_not_crashes():
    from rnnt_inference_b200 import _lib
    lib = _lib.lib
    assert lib.rnnt_gemm_i8_i32(None, None, None, 4, 64, 64, None) == -1
    assert b"null pointer" in lib.rnnt_last_error()
    one = ctypes.c_void_p(16)
    assert lib.rnnt_gemm_i8_i32(one, one, one, 4, 63, 64, None) == -1            # N % 64
    assert lib.rnnt_gemm_bf16_bf16(one, one, one, 4, 64, 20, None) == -1         # K % 8
    assert lib.rnnt_lstm_recurrent_i8(one, one, one, 1.0, 1.0, 1.0, 0, one, one, one, 3, 32, 512, one, None) == -1   # H != 1024
    assert b"1024" in lib.rnnt_last_error()
    assert lib.rnnt_lstm_postop_fp16(one, one, one, one, 7, one, 1.0, 1.0, 0, None, one, one, 2, 8, None) == -1   # ld < H
    assert lib.rnnt_tune_recurrent_cluster(3) == -1
    assert lib.rnnt_lstm_recurrent_workspace_bytes(0) == 0 and lib.rnnt_lstm_recurrent_workspace_bytes(256) > 2 * 256 * 1024
    assert lib.rnnt_greedy_decode_workspace_bytes(2048) > 0


def test_operators_reject_cpu_tensors():
    import torch
    from rnnt_inference_b200 import ops
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.quantize_pad(torch.zeros(2, 240), 1.0, 256)
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.gemm_bf16_bf16(torch.zeros(2, 64, dtype=torch.bfloat16), torch.zeros(64, 64, dtype=torch.bfloat16))


def test_qsl_uses_dataset_x_lens(tmp_path):
    """Lengths come from x_lens like the reference (pytorch_sut.py:88): a sample tensor longer than its x_lens is used up
    to x_lens only; a shorter one is an error.  Also the on-disk format: a torch-saved {"x": [...], "x_lens": [...]}."""
    import torch
    from rnnt_inference_b200.rnnt_qsl import RNNTQSL
    xs = [torch.randn(9, 240), torch.randn(5, 240), torch.randn(7, 240)]
    data = {"x": xs, "x_lens": [torch.tensor(9), torch.tensor(4), torch.tensor(7)]}
    path = tmp_path / "dev-clean-npy.pt"
    torch.save(data, path)
    qsl = RNNTQSL(str(path), pin_memory=False)
    assert qsl.count == 3 and qsl.rows == [9, 4, 7] and qsl.x_lens_host == [9, 4, 7]
    assert sorted(range(3), key=qsl.sort_key) == [0, 2, 1]
    x, n = qsl[1]
    assert int(n) == 4 and torch.equal(x[:4], xs[1][:4])
    with pytest.raises(RuntimeError, match="x_lens"):
        RNNTQSL({"x": xs, "x_lens": [torch.tensor(9), torch.tensor(6), torch.tensor(7)]}, pin_memory=False)
