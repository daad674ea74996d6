"""Launcher: calibration -> (export) -> LoadGen run (reference: models/main.py:19-93).

  1. --calibration   PytorchSUT(model, calib set, batch, "calib"): every batch of the calibration set goes through
                     inference() with the activation quantisers in "calib" mode (they widen their `_amax` to cover what they
                     see, quant_modules.py:110-115 / quant_lstm.py:166-171); the state dict, now carrying the ranges, is saved
                     next to the model as rnnt_calib.pt and becomes the model of the following steps (main.py:22-38).
  2. --save_jit      the reference freezes TorchScript modules for its C++ harness (main.py:40-58).  There is nothing to export
                     here: the B200 path is the C-ABI library plus the state dict, and RNNT refuses TorchScript archives.
                     The flag is accepted and reported as a no-op.
  3. --benchmark / --accuracy   PytorchSUT(model, inference set, batch, run_mode) under LoadGen (the real mlperf_loadgen if
                     it is installed, else loadgen_shim) with the Offline / Server settings of the config files
                     (main.py:60-89).  Returns the shim's result dict (None under the real LoadGen, which writes its logs).
"""
import argparse
import os

import torch

try:
    import mlperf_loadgen as lg
except ImportError:   # not installed offline
    from . import loadgen_shim as lg

from . import arguments
from .pytorch_sut import PytorchSUT
from .utils import logger, seq_to_sen

scenario_map = {"SingleStream": lg.TestScenario.SingleStream, "Offline": lg.TestScenario.Offline, "Server": lg.TestScenario.Server}


def calibrate(args):
    """main.py:22-38.  Returns the path of the calibrated state dict."""
    # ranges are collected on the fp32 graph (calib_model.sh runs this step without --enable_bf16): with bf16 the joint's
    # parameters would be the packed bf16 matrices, which are not a checkpoint format
    calib_args = argparse.Namespace(**{**vars(args), "enable_bf16": False})
    sut = PytorchSUT(args.model_path, args.calib_dataset_dir, args.batch_size, "calib", calib_args)
    print("==> Running calibration...")
    for i in range(0, sut.qsl.count, args.batch_size):
        results, results_idx = sut.inference(range(i, min(i + args.batch_size, sut.qsl.count)))
        if logger.isEnabledFor(10):
            for j in range(len(results)):
                logger.debug(f"{i + j}::{seq_to_sen(results[j], results_idx[j])}")
    out = os.path.join(os.path.dirname(args.model_path), "rnnt_calib.pt")
    torch.save(sut.model.rnnt.state_dict(), out)
    return out


def run_loadgen(args):
    """main.py:60-89."""
    sut = PytorchSUT(args.model_path, args.infer_dataset_dir, args.batch_size, args.run_mode, args)
    settings = lg.TestSettings()
    settings.scenario = scenario_map[args.scenario]
    for conf in (args.mlperf_conf, args.user_conf):
        if conf and os.path.exists(conf):
            settings.FromConfig(conf, "rnnt", args.scenario)
    settings.mode = lg.TestMode.AccuracyOnly if args.accuracy else lg.TestMode.PerformanceOnly
    log_settings = None
    if args.log_dir:
        os.makedirs(args.log_dir, exist_ok=True)
        log_output_settings = lg.LogOutputSettings()
        log_output_settings.outdir = args.log_dir
        log_output_settings.copy_summary_to_stdout = True
        log_settings = lg.LogSettings()
        log_settings.log_output = log_output_settings
    print("==> Running loadgen test")
    result = lg.StartTestWithLogSettings(sut.sut, sut.qsl.qsl, settings, log_settings)
    print("Done!")
    return result


def main(argv=None):
    args = arguments.parse_args(argv)
    result = None
    if args.calibration:
        args.model_path = calibrate(args)
    if args.save_jit:
        print("==> --save_jit: nothing to export on the B200 path (C-ABI library + state dict); skipped")
        args.save_jit = False
    if args.benchmark or args.accuracy:
        result = run_loadgen(args)
    return result


if __name__ == "__main__":
    main()
