"""Calibration -> LoadGen flow of the launcher (reference: models/main.py:19-93)."""
import os

import numpy as np
import pytest
import torch

from rnnt_inference_b200 import synthetic as S


def test_arguments_and_conf_parsing(tmp_path):
    from rnnt_inference_b200 import arguments, loadgen_shim as lg
    args = arguments.build_parser().parse_args(["--calibration", "--run_mode", "quant", "--batch_size", "8", "--accuracy"])
    assert args.calibration and args.accuracy and not args.benchmark and args.run_mode == "quant" and args.split_len == -1
    conf = tmp_path / "user.conf"
    conf.write_text("# comment\n*.Server.target_qps = 4200.0\n*.Offline.min_query_count = 300\nbert.Offline.min_query_count = 7\n"
                    "rnnt.Offline.min_query_count = 48\n")
    st = lg.TestSettings()
    assert st.FromConfig(str(conf), "rnnt", "Offline") == 0 and st.min_query_count == 48
    st = lg.TestSettings()
    st.FromConfig(str(conf), "rnnt", "Server")
    assert st.server_target_qps == 4200.0 and st.min_query_count == 0
    assert lg.TestSettings().FromConfig(str(tmp_path / "missing.conf"), "rnnt", "Offline") == 1


@pytest.mark.gpu
def test_calibrate_then_accuracy_run(tmp_path):
    """An fp32 checkpoint without ranges -> --calibration writes rnnt_calib.pt whose quantiser ranges cover the calibration
    set -> the int8 model built from it answers an AccuracyOnly query; its transcripts stay close to the fp32 model's."""
    from rnnt_inference_b200 import eval_accuracy as E
    from rnnt_inference_b200 import main as M
    from rnnt_inference_b200.pytorch_sut import PytorchSUT
    sd = S.make_checkpoint(with_amax=False)
    assert not any(k.endswith("_amax") for k in sd)
    model_path = str(tmp_path / "rnnt.pt")
    torch.save(sd, model_path)
    lens = [60, 44, 52, 37, 25, 48, 33, 58, 41, 29, 55, 36]
    feats = [S.make_features(n, 40 + i) for i, n in enumerate(lens)]
    data_path = str(tmp_path / "dev.pt")
    torch.save({"x": feats, "x_lens": [torch.tensor(n) for n in lens]}, data_path)
    conf = tmp_path / "user.conf"
    conf.write_text(f"*.Offline.min_query_count = {len(lens)}\n")
    out = M.main(["--calibration", "--accuracy", "--run_mode", "quant", "--enable_bf16", "--batch_size", "8", "--model_path", model_path,
                  "--calib_dataset_dir", data_path, "--infer_dataset_dir", data_path, "--user_conf", str(conf),
                  "--mlperf_conf", str(tmp_path / "none.conf"), "--log_dir", str(tmp_path / "logs"), "--save_jit"])
    calib_path = os.path.join(str(tmp_path), "rnnt_calib.pt")
    assert os.path.exists(calib_path)
    csd = torch.load(calib_path)
    amax = {k: float(v.max()) for k, v in csd.items() if k.endswith("_amax")}
    # every layer's activation (input) quantiser has collected a range; weight ranges are derived from the weights when the
    # int8 model is built and output quantisers are chained to the next layer's input there (modeling_rnnt.py:62-75)
    inputs = {k: v for k, v in amax.items() if "input_quantizer" in k}
    assert len(inputs) == 5 and all(v > 0 for v in inputs.values()), amax
    # the first layer's input quantiser saw the features themselves
    first = [v for k, v in amax.items() if "pre_rnn" in k and "input" in k]
    assert first and abs(max(first) - max(float(f.abs().max()) for f in feats)) < 1e-5
    assert out["completed"] == len(lens)
    got = {e["qsl_idx"]: list(np.frombuffer(bytes.fromhex(e["data"]), np.int32)) for e in out["accuracy_log"]}
    # fp32 model on the same data: calibrated int8 transcripts stay close (a quantised model, not a bit-exact one)
    f32 = PytorchSUT(model_path, data_path, 8, "f32", None, scenario="Offline")
    order = sorted(range(len(lens)), key=lambda i: -lens[i])
    ref = {}
    for c in range(0, len(order), 8):
        idx = order[c:c + 8]
        res, res_len = f32.inference(idx)
        res, res_len = res.cpu().numpy(), res_len.cpu().numpy()
        for j, i in enumerate(idx):
            ref[i] = [int(t) for t in res[j, :res_len[j]]]
    assert sum(len(v) for v in ref.values()) > 20
    wer = E.word_error_rate([" ".join(map(str, got[i])) for i in range(len(lens))], [" ".join(map(str, ref[i])) for i in range(len(lens))])
    assert wer[0] < 0.35, wer
