#!/usr/bin/env python
"""Generates tests/golden/ref_py_f32.npz by running the REFERENCE's own Python hot path (models/modeling_rnnt.py,
decoder.py, quant_lstm.py, quant_modules.py, utils.py, config.py imported verbatim from /root/reference/models) in
the modes that need no native plugin: run_mode "f32" and "calib" (SURVEY 8(c) tier 1; "fake_quant" cannot run at the
reference's HEAD: QuantLSTM.__init__ skips iLSTM.__init__, so `_process_parameters` fails on the missing
`self.weights`, quant_lstm.py:57-60,106-107).  The reference's
`_C.py` loads libmlperf_plugins.so unconditionally, so a stub module named `_C` with the three functions these modes
call (`lstm`, `prepack_lstm_weights`, `prepack_linear_weight`; textbook LSTM math, tests/rnn_f32.py:54-78 of the
reference) is placed earlier on sys.path.  Runs only in the build container (the reference is not on the GPU box);
the vectors are committed.  Test infrastructure.

    python tests/golden/make_golden_f32.py
"""
import os
import sys
import tempfile
import textwrap

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference/models"

STUB = textwrap.dedent('''
    import torch

    def prepack_lstm_weights(w_ih, w_hh):
        return [w_ih, w_hh]

    def prepack_linear_weight(w):
        return w

    def lstm(x, hx, cx, all_weights):
        """x {T, N, K}; per layer weights [w_ih, w_hh, b_ih, b_hh]; returns (y, [hy], [cy]) (new tensors)."""
        hy, cy = [], []
        for layer, (w_ih, w_hh, b_ih, b_hh) in enumerate(all_weights):
            h, c = hx[layer], cx[layer]
            ys = []
            for t in range(x.shape[0]):
                gates = torch.nn.functional.linear(x[t], w_ih, b_ih) + torch.nn.functional.linear(h, w_hh, b_hh)
                i, f, g, o = gates.chunk(4, 1)
                c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(g)
                h = torch.sigmoid(o) * torch.tanh(c)
                ys.append(h)
            x = torch.stack(ys, 0)
            hy.append(h)
            cy.append(c)
        return x, hy, cy
''')


def main():
    tmp = tempfile.mkdtemp()
    open(os.path.join(tmp, "_C.py"), "w").write(STUB)
    sys.path.insert(0, REF)
    sys.path.insert(0, tmp)            # the stub shadows models/_C.py
    sys.path.insert(1, ROOT)
    torch.set_num_threads(1)           # bit-reproducible fp32 sums
    import _C
    assert _C.__file__.startswith(tmp)
    from decoder import GreedyDecoder            # the reference's
    from modeling_rnnt import RNNT                # the reference's
    import quant_lstm as ref_quant_lstm
    from rnnt_inference_b200 import synthetic as S

    lens = [80, 61]
    feats = [S.make_features(l, 9000 + i) for i, l in enumerate(lens)]
    x = torch.nn.utils.rnn.pad_sequence(feats)                    # {T, N, 240}
    x_lens = torch.tensor(lens, dtype=torch.int64)
    out = {"lens": np.array(lens, np.int64), "feature_ids": np.array([9000, 9001], np.int64)}

    for mode in ("f32", "calib"):
        sd = S.make_checkpoint(with_amax=False)
        path = os.path.join(tmp, f"ckpt_{mode}.pt")
        torch.save(sd, path)
        rnnt = RNNT(path, mode, False, False).eval()
        dec = GreedyDecoder(rnnt, mode)
        with torch.no_grad():
            res, res_len = dec(x.clone(), x_lens.clone())
        out[f"{mode}_res_len"] = res_len.numpy().astype(np.int64)
        out[f"{mode}_res"] = res[:, :int(res_len.max())].numpy().astype(np.int64)
        print(mode, "tokens per utterance", res_len.tolist())
        with torch.no_grad():
            z = lambda: [torch.zeros((2, 1024)) for _ in range(5)]
            hx, cx = z(), z()
            f, *_ = rnnt.transcription(x.clone(), x_lens.clone(), hx[:2], cx[:2], hx[2:], cx[2:])
        out[f"{mode}_f"] = f[::4].numpy().astype(np.float32)      # every 4th stacked frame of the encoder output
        if mode == "calib":
            tr = rnnt.transcription
            layers = [getattr(tr.pre_rnn, f"lstm{i}") for i in range(2)] + [getattr(tr.post_rnn, f"lstm{i}") for i in range(3)]
            out["calib_input_amax"] = np.array([float(l.input_quantizer.amax) for l in layers], np.float32)
            print("calibrated input amax", out["calib_input_amax"])

    # ---- QuantLSTM stand-alone (a17), small sizes: f32 and calib cells with their state dict
    torch.manual_seed(5)
    for mode in ("f32", "calib"):
        lstm = ref_quant_lstm.QuantLSTM(24, 32, 2, skip_quant_y=False)
        sd0 = {k: v.clone() for k, v in lstm.state_dict().items()}
        lstm._init_layers(mode)
        lstm._process_parameters(mode)
        xs = torch.randn(6, 3, 24)
        with torch.no_grad():
            y, hy, cy = lstm(xs.clone(), None, None)
        out[f"qlstm_{mode}_x"] = xs.numpy()
        out[f"qlstm_{mode}_y"] = y.numpy()
        out[f"qlstm_{mode}_hy"] = torch.stack(hy).numpy()
        out[f"qlstm_{mode}_cy"] = torch.stack(cy).numpy()
        if mode == "calib":
            out["qlstm_calib_amax"] = np.array([float(getattr(lstm, f"lstm{i}").input_quantizer.amax) for i in range(2)], np.float32)
        for k, v in sd0.items():
            out[f"qlstm_{mode}_sd_{k}"] = v.numpy()
    np.savez_compressed(os.path.join(HERE, "ref_py_f32.npz"), **out)
    print("wrote", os.path.join(HERE, "ref_py_f32.npz"), os.path.getsize(os.path.join(HERE, "ref_py_f32.npz")), "bytes")


if __name__ == "__main__":
    main()
