"""Synthetic checkpoint, calibration ranges and utterance set (datasets and checkpoints are unavailable offline).

Follows SURVEY.md §8(d): the checkpoint is written in the MLPerf key format so that utils.migrate_state_dict and
RNNT._load_model run exactly as for the real rnnt.pt; `_amax` buffers stand in for a calibration run; the utterance
lengths follow a fixed-seed gamma fit of LibriSpeech dev-clean (<= 15 s); features look like per-feature-normalised
stacked log-mel frames (240 = 80 mel x 3).
"""
import math

import numpy as np
import torch

from .config import RNNTParam

INPUT_AMAX = 5.8358          # "input" range of tests/calibration_result_nv.cache in the reference
FRAME_SECONDS = 0.03         # one stacked frame = 3 x 10 ms
DEFAULT_LSTM_GAIN = 3.0       # encoder LSTM weights U(-g/sqrt(H), g/sqrt(H)): g = 1 (torch default) gives a nearly constant encoder
DEFAULT_BLANK_BIAS = -2.0     # fc2 bias of BLANK relative to the labels: tuned once (with memory 0.9, penalty 3) so that ~0.8
                              # symbols are emitted per stacked frame, LibriSpeech-like (~15 characters/s at 60 ms per frame)


def make_checkpoint(seed=7, blank_bias=DEFAULT_BLANK_BIAS, with_amax=True, lstm_gain=DEFAULT_LSTM_GAIN, evidence_gain=3.0,
                    memory=0.9, penalty=3.0):
    """Random-init MLPerf RNN-T state dict (keys as in the MLPerf checkpoint).

    The encoder is plain random.  A plain random prediction/joint network is bistable under greedy decoding (either
    never emits or emits 30 symbols on every frame), which would make the decode load meaningless, so that part is
    random PLUS a little structure that reproduces what a trained transducer does: the prediction LSTM keeps a decaying
    trace of the recently emitted labels, joint hidden units 0..27 read that trace and push the logit of those labels
    down, and the remaining 484 hidden units are a random read-out of the encoder frame ("acoustic evidence")."""
    g = torch.Generator().manual_seed(seed)
    H, P, J, V = (RNNTParam.trans_hidden_size, RNNTParam.pred_hidden_size, RNNTParam.joint_hidden_size, RNNTParam.num_labels)
    NL = V - 1   # 28 real labels

    def uni(shape, bound):
        return (torch.rand(shape, generator=g) * 2 - 1) * bound

    def lstm(prefix, in_size, hidden, layers, sd, gain):
        bound = gain / math.sqrt(hidden)
        for layer in range(layers):
            k = in_size if layer == 0 else hidden
            sd[f"{prefix}.weight_ih_l{layer}"] = uni((4 * hidden, k), bound)
            sd[f"{prefix}.weight_hh_l{layer}"] = uni((4 * hidden, hidden), bound)
            b_ih, b_hh = uni((4 * hidden,), bound), uni((4 * hidden,), bound)
            b_ih[hidden:2 * hidden] += 1.0      # forget_gate_bias = 1.0 (configs/rnnt.toml)
            sd[f"{prefix}.bias_ih_l{layer}"] = b_ih
            sd[f"{prefix}.bias_hh_l{layer}"] = b_hh

    sd = {}
    lstm("encoder.pre_rnn.lstm", RNNTParam.trans_input_size, H, RNNTParam.pre_num_layers, sd, lstm_gain)
    lstm("encoder.post_rnn.lstm", H * RNNTParam.stack_time_factor, H, RNNTParam.post_num_layers, sd, lstm_gain)
    # ---- prediction: embedding = 2 * e_k + noise; LSTM = small random + pass-through of the cell candidate
    embed = torch.randn((NL, P), generator=g) * 0.05
    embed[torch.arange(NL), torch.arange(NL)] += 2.0
    sd["prediction.embed.weight"] = embed
    lstm("prediction.dec_rnn.lstm", P, P, RNNTParam.pred_num_layers, sd, 0.3)
    f_bias = math.log(memory / (1.0 - memory))
    for layer in range(RNNTParam.pred_num_layers):
        w_ih = sd[f"prediction.dec_rnn.lstm.weight_ih_l{layer}"]
        w_ih[2 * P + torch.arange(P), torch.arange(P)] += 3.0                 # candidate gate follows its own input coordinate
        b = sd[f"prediction.dec_rnn.lstm.bias_ih_l{layer}"]
        b[0:P] += 4.0                                                         # input gate open
        b[P:2 * P] += (f_bias if layer == 0 else -4.0) - 1.0                  # layer 0 remembers, layer 1 is memoryless
        b[3 * P:4 * P] += 4.0                                                 # output gate open
    # ---- joint: hidden 0..27 = label trace (prediction half only), hidden 28.. = random evidence (encoder half only)
    w1 = torch.zeros((J, H + P))
    w1[NL:, :H] = torch.randn((J - NL, H), generator=g) * (evidence_gain / math.sqrt(H))
    w1[NL:, H:] = torch.randn((J - NL, P), generator=g) * (0.5 / math.sqrt(P))
    w1[torch.arange(NL), H + torch.arange(NL)] = 8.0
    sd["joint_net.0.weight"] = w1
    sd["joint_net.0.bias"] = torch.zeros(J)
    w2 = torch.zeros((V, J))
    w2[:, NL:] = torch.randn((V, J - NL), generator=g) * (evidence_gain / math.sqrt(J - NL))
    w2[torch.arange(NL), torch.arange(NL)] = -penalty
    sd["joint_net.3.weight"] = w2
    b2 = torch.full((V,), 10.0)      # keeps every real logit above the three zero pad columns of the 32-wide argmax
    b2[RNNTParam.BLANK] += blank_bias
    sd["joint_net.3.bias"] = b2
    if with_amax:
        # synthetic calibration: first layer sees the features, every other layer sees |h| < 1
        for rnn, layers in (("pre_rnn", RNNTParam.pre_num_layers), ("post_rnn", RNNTParam.post_num_layers)):
            for layer in range(layers):
                first = rnn == "pre_rnn" and layer == 0
                base = f"transcription.{rnn}.lstm{layer}"
                sd[f"{base}.input_quantizer._amax"] = torch.tensor(INPUT_AMAX if first else 1.0)
                sd[f"{base}.output_quantizer._amax"] = torch.tensor(1.0)
                sd[f"{base}.weight_quantizer._amax"] = torch.tensor(0.0)   # recomputed from the weights at load
    return sd


def utterance_lengths(count, seed=2024, min_s=1.3, max_s=15.0):
    """Stacked-frame counts of `count` utterances: gamma(k=2.6, theta=2.6 s) clipped to [min_s, max_s]."""
    rng = np.random.default_rng(seed)
    dur = np.clip(rng.gamma(2.6, 2.6, size=count), min_s, max_s)
    return np.ceil(dur / FRAME_SECONDS).astype(np.int64)


def make_features(length, utt_id, dim=RNNTParam.trans_input_size):
    """fp32 {length, dim} ~ N(0,1) clipped to +-INPUT_AMAX, seeded per utterance."""
    g = torch.Generator().manual_seed(1234 + int(utt_id))
    return torch.randn((int(length), dim), generator=g).clamp_(-INPUT_AMAX, INPUT_AMAX)


def make_dataset(count, seed=2024, fixed_len=None, unique=None):
    """The on-disk sample format of the reference QSL: {"x": [T_i x 240 fp32], "x_lens": [int tensor]}
    (datasets/convert_librispeech.py:135-163, models/rnnt_qsl.py:7).  `unique` bounds the number of distinct feature
    tensors (the rest are replicas, like MLPerf replicating 2513 utterances to the min query count)."""
    lens = np.full(count, fixed_len, np.int64) if fixed_len else utterance_lengths(count, seed)
    unique = count if unique is None else min(unique, count)
    base = [make_features(lens[i], i) for i in range(unique)]
    xs, xl = [], []
    for i in range(count):
        src = base[i % unique]
        xs.append(src)
        xl.append(torch.tensor(src.shape[0], dtype=torch.int64))
    return {"x": xs, "x_lens": xl}
