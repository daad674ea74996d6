"""fp32 tier (SURVEY 8(c) tier 1): this package's run_mode "f32" / "calib" modules against outputs of the REFERENCE's
own Python (models/*.py imported verbatim behind a stub `_C`; tests/golden/make_golden_f32.py made the fixture in the
build container).  Covers a6 GreedyDecoder.greedy_decode_f32, a17 QuantLSTM / QuantLSTMLayer, and the calibration pass
(amax collection, 8(f)-4).  Plain torch on the CPU: these modes never touch the CUDA library.
Tolerance: tokens identical; fp32 tensors within 2e-5 (summation order of torch's CPU GEMM may differ by thread count).
"""
import os

import numpy as np
import pytest
import torch

from rnnt_inference_b200 import synthetic as S
from rnnt_inference_b200.decoder import GreedyDecoder
from rnnt_inference_b200.modeling_rnnt import RNNT
from rnnt_inference_b200.quant_lstm import QuantLSTM

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_py_f32.npz"))


def _batch():
    lens = GOLD["lens"].tolist()
    feats = [S.make_features(l, int(i)) for l, i in zip(lens, GOLD["feature_ids"])]
    return torch.nn.utils.rnn.pad_sequence(feats), torch.tensor(lens, dtype=torch.int64)


@pytest.mark.parametrize("mode", ["f32", "calib"])
def test_greedy_decode_f32_vs_reference_python(mode):
    torch.set_num_threads(1)
    x, x_lens = _batch()
    rnnt = RNNT(S.make_checkpoint(with_amax=False), mode, False, device=torch.device("cpu")).eval()
    dec = GreedyDecoder(rnnt, mode)
    with torch.no_grad():
        res, res_len = dec(x.clone(), x_lens.clone())
    assert res.dtype == torch.int64                                   # decoder.py:30: int64 outside quant mode
    assert res_len.tolist() == GOLD[f"{mode}_res_len"].tolist()
    want = GOLD[f"{mode}_res"]
    assert np.array_equal(res[:, :want.shape[1]].numpy(), want)
    with torch.no_grad():
        z = lambda: [torch.zeros((2, 1024)) for _ in range(5)]
        hx, cx = z(), z()
        f, *_ = rnnt.transcription(x.clone(), x_lens.clone(), hx[:2], cx[:2], hx[2:], cx[2:])
    assert np.abs(f[::4].numpy() - GOLD[f"{mode}_f"]).max() <= 2e-5
    if mode == "calib":
        tr = rnnt.transcription
        cells = [getattr(tr.pre_rnn, f"lstm{i}") for i in range(2)] + [getattr(tr.post_rnn, f"lstm{i}") for i in range(3)]
        amax = np.array([float(c.input_quantizer.amax) for c in cells], np.float32)
        assert np.allclose(amax, GOLD["calib_input_amax"], rtol=1e-5)


@pytest.mark.parametrize("mode", ["f32", "calib"])
def test_quant_lstm_vs_reference_python(mode):
    torch.set_num_threads(1)
    lstm = QuantLSTM(24, 32, 2, skip_quant_y=False)
    prefix = f"qlstm_{mode}_sd_"
    lstm.load_state_dict({k[len(prefix):]: torch.from_numpy(GOLD[k]) for k in GOLD.files if k.startswith(prefix)})
    lstm._init_layers(mode)
    lstm._process_parameters(mode)
    with torch.no_grad():
        y, hy, cy = lstm(torch.from_numpy(GOLD[f"qlstm_{mode}_x"]).clone(), None, None)
    assert np.abs(y.numpy() - GOLD[f"qlstm_{mode}_y"]).max() <= 2e-6
    assert np.abs(torch.stack(hy).numpy() - GOLD[f"qlstm_{mode}_hy"]).max() <= 2e-6
    assert np.abs(torch.stack(cy).numpy() - GOLD[f"qlstm_{mode}_cy"]).max() <= 2e-6
    if mode == "calib":
        amax = np.array([float(getattr(lstm, f"lstm{i}").input_quantizer.amax) for i in range(2)], np.float32)
        assert np.allclose(amax, GOLD["qlstm_calib_amax"], rtol=1e-6)
