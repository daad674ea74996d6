"""Small helpers shared by the model and the SUT (reference: models/utils.py)."""
import logging
import os

import torch

LOG_LEVEL = int(os.environ.get("RNNT_LOG_LEVEL", logging.INFO))
logger = logging.getLogger("RNNTLogger")
logger.setLevel(LOG_LEVEL)
if not logger.handlers:
    _handler = logging.StreamHandler()
    _handler.setFormatter(logging.Formatter("[%(filename)s:%(lineno)d %(levelname)s] %(message)s"))
    logger.addHandler(_handler)

# index -> character; 28 symbols, BLANK = 28 is never emitted (models/utils.py:23-52)
labels = [" "] + [chr(ord("a") + i) for i in range(26)] + ["'"]


def seq_to_sen(seq, seq_lens):
    """Token ids -> text (models/utils.py:55-57)."""
    return "".join(labels[int(seq[i])] for i in range(int(seq_lens)))


_KEY_RENAMES = (
    ("encoder.pre_rnn.lstm", "transcription.pre_rnn"),
    ("encoder.post_rnn.lstm", "transcription.post_rnn"),
    ("dec_rnn.lstm", "pred_rnn"),
    ("joint_net.0", "joint.linear1"),
    ("joint_net.3", "joint.linear2"),
)
_DROPPED_KEYS = ("audio_preprocessor.featurizer.fb", "audio_preprocessor.featurizer.window")


def migrate_state_dict(model, split_fc1=True):
    """MLPerf checkpoint keys -> this package's module names (models/utils.py:60-81).  The first joint linear
    [512, 1344] is split column-wise into the transcription half [:, :1024] (bias := 0) and the prediction half
    [:, 1024:] (keeps the bias), so that the transcription half can be hoisted out of the decode loop."""
    state_dict = model["state_dict"] if "state_dict" in model else model
    out = {}
    for key, value in state_dict.items():
        if split_fc1 and key == "joint_net.0.weight":
            out["joint.linear1_trans.weight"] = value[:, :1024]
            out["joint.linear1_pred.weight"] = value[:, 1024:]
            continue
        if split_fc1 and key == "joint_net.0.bias":
            out["joint.linear1_trans.bias"] = torch.zeros(value.shape[0])
            out["joint.linear1_pred.bias"] = value
        for old, new in _KEY_RENAMES:
            key = key.replace(old, new)
        out[key] = value
    for key in _DROPPED_KEYS:
        out.pop(key, None)
    return out
