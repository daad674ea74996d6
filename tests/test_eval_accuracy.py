"""eval_accuracy: WER / token agreement / margin attribution (reference: eval_accuracy.py:28-101)."""
import json

import numpy as np

from rnnt_inference_b200 import eval_accuracy as E


def test_levenshtein_and_wer():
    assert E.levenshtein("kitten", "sitting") == 3
    assert E.levenshtein([], [1, 2]) == 2
    wer, edits, words = E.word_error_rate(["the cat sat", "a b"], ["the cat sat down", "a c"])
    assert (edits, words) == (2, 6) and abs(wer - 2 / 6) < 1e-12


def test_accuracy_log_roundtrip(tmp_path):
    toks = {0: [1, 2, 0, 27], 3: []}
    path = tmp_path / "mlperf_log_accuracy.json"
    path.write_text(json.dumps([{"qsl_idx": k, "data": np.array(v, np.int32).tobytes().hex()} for k, v in toks.items()]))
    assert E.load_accuracy_log(str(path)) == toks
    assert E.to_sentence([8, 9, 0, 1]) == "hi a"


def test_compare_transcripts_attributes_difference_to_margin():
    truth = [[1, 2, 3], [4, 5]]
    test = [[1, 2, 3], [4, 6, 7]]
    # oracle decisions of lane 1: emit 4 (margin 2.0), blank (0.5), emit 5 (0.001), blank (3.0)
    trace = dict(sym=np.array([[1, 2, 3, 28, 28], [4, 28, 5, 28, -1]], np.int8),
                 margin=np.array([[1, 1, 1, 1, 1], [2.0, 0.5, 0.001, 3.0, 0]], np.float32), count=np.array([5, 4], np.int32))
    rep = E.compare_transcripts(test, truth, trace, margin_tol=0.01)
    assert rep["identical_utterances"] == 1 and rep["token_edit_distance"] == 2
    d = rep["differences"][0]
    assert d["utterance"] == 1 and d["first_differing_token"] == 1 and abs(d["margin"] - 0.001) < 1e-9
    assert rep["unexplained"] == []
    assert E.compare_transcripts(test, truth, trace, margin_tol=1e-4)["unexplained"] == [d]
