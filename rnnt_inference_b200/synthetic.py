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
DEFAULT_LSTM_GAIN = 3.0      # encoder LSTM weights U(-g/sqrt(H), g/sqrt(H)): g = 1 (torch default) gives a nearly constant encoder
# Decode-side constants, tuned once against the CPU oracle on the bench set (profiles/r02/decode_workload.md has the
# histograms): greedy emissions are spread over the utterance at ~0.25 symbols per stacked frame, almost
# never more than a handful per frame, transcripts differ between utterances, and a rare "burst" unit makes a few
# frames run into the 30-symbol cap so that quirk (b) of SURVEY 8(a) is exercised at bench scale.
DEFAULT_BLANK_BIAS = 2.0     # fc2 bias of BLANK relative to the labels (the emission threshold on the label evidence)
DEFAULT_MEMORY = 0.8         # per-emitted-symbol decay of the label trace kept by the prediction LSTM
DEFAULT_PENALTY = 8.0        # how hard a label's trace pushes its own logit down (no A-B-A-B ping-pong inside a frame)
DEFAULT_BURST_THRESHOLD = 4.22   # pre-activation above which the burst unit fires.  Seed and threshold were picked over 600 candidate
DEFAULT_BURST_SEED = 496         # directions (oracle-encoded bench set): 5 of its 279 778 stacked frames fire, at most 3 in one
                                 # utterance, and 2 of them are an utterance's LAST frame (quirk b).  Cap frames are rare in
                                 # real speech; a burst costs 30 strictly sequential prediction-network evaluations
BURST_GAIN = 400.0               # logit lift per unit of excess pre-activation: any firing frame runs into the cap


def make_checkpoint(seed=7, blank_bias=DEFAULT_BLANK_BIAS, with_amax=True, lstm_gain=DEFAULT_LSTM_GAIN, evidence_gain=3.0,
                    memory=DEFAULT_MEMORY, penalty=DEFAULT_PENALTY, burst_threshold=DEFAULT_BURST_THRESHOLD,
                    burst_seed=DEFAULT_BURST_SEED):
    """Random-init MLPerf RNN-T state dict (keys as in the MLPerf checkpoint).

    The encoder is plain random; its output varies smoothly in time (correlation time ~5 stacked frames) and between
    utterances.  A plain random prediction/joint network is bistable under greedy decoding (it either never emits or
    emits 30 symbols on every frame), which would make the decode load meaningless, so that part is random PLUS the
    little structure a trained transducer has:
      * joint hidden units 29.. are a random read-out of the encoder frame ("acoustic evidence" for each label); BLANK has
        no evidence, only a bias, so a label is emitted when its evidence rises above that threshold;
      * the prediction LSTM keeps a trace of the recently emitted labels that decays by `memory` per emitted symbol; joint
        hidden units 0..27 read it and push the logit of those labels down by `penalty` x trace, so a label is emitted
        once per excursion of its evidence and the frame then falls back to BLANK;
      * joint hidden unit 28 ("burst") is a rare, encoder-only detector that lifts every label far above BLANK: such a
        frame emits until the 30-symbol cap (max_symbols_per_step) stops it."""
    g = torch.Generator().manual_seed(seed)
    H, P, J, V = (RNNTParam.trans_hidden_size, RNNTParam.pred_hidden_size, RNNTParam.joint_hidden_size, RNNTParam.num_labels)
    NL = V - 1   # 28 real labels
    E0 = NL + 1  # first evidence unit of the joint hidden layer

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
    g = torch.Generator().manual_seed(seed + 1000)     # the decode-side draws do not depend on the encoder's
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
    # ---- joint fc1: hidden 0..27 = label trace (prediction half only), 28 = burst detector (encoder half only),
    #      29.. = random evidence (encoder half, plus a little of the prediction state)
    w1 = torch.zeros((J, H + P))
    w1[E0:, :H] = torch.randn((J - E0, H), generator=g) * (evidence_gain / math.sqrt(H))
    w1[E0:, H:] = torch.randn((J - E0, P), generator=g) * (0.5 / math.sqrt(P))
    w1[torch.arange(NL), H + torch.arange(NL)] = 8.0
    b1 = torch.zeros(J)
    gb = torch.Generator().manual_seed(seed + 2000 + burst_seed)
    w1[NL, :H] = torch.randn(H, generator=gb) * (evidence_gain / math.sqrt(H))
    b1[NL] = -burst_threshold
    sd["joint_net.0.weight"] = w1
    sd["joint_net.0.bias"] = b1
    # ---- joint fc2: label logit = 10 + evidence - penalty * trace + BURST_GAIN * burst; BLANK logit = 10 + blank_bias
    w2 = torch.zeros((V, J))
    w2[:NL, E0:] = torch.randn((NL, J - E0), generator=g) * (evidence_gain / math.sqrt(J - E0))
    w2[torch.arange(NL), torch.arange(NL)] = -penalty
    w2[:NL, NL] = BURST_GAIN
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
