"""Command-line options of the launcher (reference: models/arguments.py) — same names and defaults; `--manifest_path` is
optional here (the reference requires it but only its dataset scripts read it)."""
import argparse


def build_parser():
    p = argparse.ArgumentParser(description="RNN-T inference on B200: calibration, accuracy and benchmark runs")
    p.add_argument("--perf_count", type=int, default=None, help="number of samples")
    p.add_argument("--batch_size", type=int, default=32)
    p.add_argument("--split_len", type=int, default=-1, help="max splitted sequence length(value=even)")
    p.add_argument("--scenario", default="Offline", choices=["SingleStream", "Offline", "Server"], help="Scenario")
    p.add_argument("--mlperf_conf", default="../configs/mlperf.conf", help="mlperf rules config")
    p.add_argument("--user_conf", default="../configs/user.conf", help="user config for user LoadGen settings such as target QPS")
    p.add_argument("--toml_path", type=str, default="../configs/rnnt.toml", help="rnn-t model parameters config")
    p.add_argument("--model_path", type=str, default="work_dir/rnnt.pt")
    p.add_argument("--processor_path", type=str, default="work_dir/processpr.pt")
    p.add_argument("--manifest_path", type=str, default=None)
    p.add_argument("--infer_dataset_dir", type=str)
    p.add_argument("--calib_dataset_dir", type=str)
    p.add_argument("--log_dir", type=str)
    p.add_argument("--run_mode", default=None, choices=[None, "f32", "calib", "quant", "fake_quant"], help="run_mode, default None(fp32)")
    p.add_argument("--enable_bf16", action="store_true", help="enable bf16 for prediction & joint")
    p.add_argument("--load_jit", action="store_true", help="load jit model")
    p.add_argument("--save_jit", action="store_true", help="save jit model")
    p.add_argument("--enable_process", action="store_true", help="enable audio process")
    p.add_argument("--accuracy", action="store_true", help="enable accuracy evaluation")
    p.add_argument("--benchmark", action="store_true", help="enable benchmark evaluation")
    p.add_argument("--calibration", action="store_true", help="enable model calibration")
    return p


def parse_args(argv=None):
    args = build_parser().parse_args(argv)
    print(args)
    return args
