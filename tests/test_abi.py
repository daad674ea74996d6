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


def test_quantizer_scale_is_one_correctly_rounded_division():
    """scale = 127 / amax as ONE fp32 division (the reference divides two tensors): compared with numpy's float32
    division over many amax values.  torch evaluates `python_float / tensor` as reciprocal * scalar, which is one ulp
    off for some inputs (e.g. amax = 0.09375) and flips weights that sit on a rounding tie."""
    import numpy as np
    import torch
    from rnnt_inference_b200.quant_modules import QuantDescriptor, TensorQuantizer
    q = TensorQuantizer(QuantDescriptor(mode="quant"))
    rng = np.random.default_rng(0)
    for amax in [0.09375, 5.8358, 1.0, 0.0937499] + rng.uniform(0.01, 8.0, 2000).astype(np.float32).tolist():
        q.amax = torch.tensor(amax, dtype=torch.float32)
        want = np.float32(127.0) / np.float32(amax)
        assert np.float32(q.scale.item()) == want, amax


def test_module_quantised_parameters_match_oracle_on_cpu():
    """Scale algebra / weight quantisation of the modules (quant_lstm.py:193-215, modeling_rnnt.py:62-77) against the
    oracle's independent restatement, for the BENCH checkpoint (seed 7) and the test checkpoint (seed 11): int8
    weights, folded biases and every scale bit-equal.  Loading needs no GPU."""
    import numpy as np
    import torch
    from oracle import pipeline as OP
    from rnnt_inference_b200 import synthetic as S
    from rnnt_inference_b200.modeling_rnnt import RNNT
    for seed in (7, 11):
        sd = S.make_checkpoint(seed=seed)
        rnnt = RNNT(sd, "quant", True, device=torch.device("cpu")).eval()
        params = OP.prepare_params(sd)
        i = 0
        for rnn in (rnnt.transcription.pre_rnn, rnnt.transcription.post_rnn):
            for layer in range(rnn.num_layers):
                w_ih, w_hh, _b, bias = rnn.weights[layer]
                p = params["enc"][i]
                np.testing.assert_array_equal(w_ih.numpy(), p["w_ih"], err_msg=f"seed {seed} layer {i} w_ih")
                np.testing.assert_array_equal(w_hh.numpy(), p["w_hh"], err_msg=f"seed {seed} layer {i} w_hh")
                np.testing.assert_array_equal(bias.numpy().view(np.uint32), p["bias"].view(np.uint32))
                assert float(rnn.rb_scale[layer]) == p["rb_scale"] and float(rnn.in_scale[layer]) == p["in_scale"]
                i += 1
        d = params["dec"]
        bits = lambda t: t.detach().contiguous().view(torch.int16).numpy().view(np.uint16)
        np.testing.assert_array_equal(bits(rnnt.joint.linear1_trans.weight), d["w_trans"])
        np.testing.assert_array_equal(bits(rnnt.joint.linear1_pred.weight), d["w_pred"])
        np.testing.assert_array_equal(bits(rnnt.joint.linear2.weight), d["w2"])
        np.testing.assert_array_equal(rnnt.joint.linear1_bias.detach().numpy(), d["b1"])
