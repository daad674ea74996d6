"""Accuracy tooling (reference: eval_accuracy.py:28-101).

  * `load_accuracy_log(path)`       LoadGen's mlperf_log_accuracy.json ([{"qsl_idx", "data": hex int32 tokens}]) ->
                                    {qsl_idx: [token ids]} (the reference reads the same bytes with array("I"))
  * `word_error_rate(hyp, ref)`     Levenshtein distance over words / reference words (eval_accuracy.py:28-78)
  * `compare_transcripts(...)`      the report SURVEY 7.2 / north_star ask for: token-sequence agreement and WER delta of
                                    one run against another (here: the CUDA path against the CPU oracle), and — given
                                    the oracle's decision trace — the top-2 logit margin at the decision every
                                    difference starts at, so that an argmax flip is attributable to a near-tie.
CLI:  python -m rnnt_inference_b200.eval_accuracy --log_path mlperf_log_accuracy.json --manifest_path manifest.json
"""
import argparse
import array
import json

from .utils import seq_to_sen


def load_accuracy_log(path):
    with open(path) as fh:
        entries = json.load(fh)
    return {int(e["qsl_idx"]): array.array("i", bytes.fromhex(e["data"])).tolist() for e in entries}


def levenshtein(a, b):
    """Edit distance between two sequences (insert / delete / substitute, unit cost)."""
    if len(a) < len(b):
        a, b = b, a
    prev = list(range(len(b) + 1))
    for i, x in enumerate(a, 1):
        cur = [i]
        for j, y in enumerate(b, 1):
            cur.append(min(prev[j] + 1, cur[j - 1] + 1, prev[j - 1] + (x != y)))
        prev = cur
    return prev[-1]


def word_error_rate(hypotheses, references):
    """Returns (wer, edit operations, reference words); wer = inf when the references hold no word."""
    if len(hypotheses) != len(references):
        raise ValueError(f"word_error_rate: {len(hypotheses)} hypotheses for {len(references)} references")
    edits = words = 0
    for hyp, ref in zip(hypotheses, references):
        r = ref.split()
        words += len(r)
        edits += levenshtein(hyp.split(), r)
    return (edits / words if words else float("inf")), edits, words


def to_sentence(tokens):
    return seq_to_sen(list(tokens), len(tokens))


def _first_difference(a, b):
    n = min(len(a), len(b))
    for i in range(n):
        if a[i] != b[i]:
            return i
    return n if len(a) != len(b) else None


def margin_at_divergence(decisions, margins, count, token_pos, cap=30, blank=28):
    """Smallest top-2 margin among the oracle's decisions taken while it had emitted exactly `token_pos` tokens (the
    blanks after token token_pos-1 and the decision that emitted token token_pos): the decisions at which a run that
    agreed on the first token_pos tokens can have departed."""
    emitted = per_frame = 0
    best = float("inf")
    for k in range(int(count)):
        sym = int(decisions[k])
        if emitted == token_pos:
            best = min(best, float(margins[k]))
        if sym != blank and per_frame != cap:
            emitted += 1
            per_frame += 1
            if emitted > token_pos:
                break
        else:
            per_frame = 0
    return best


def compare_transcripts(test, truth, trace=None, margin_tol=None):
    """test / truth: lists of token-id lists, same order.  trace: optional oracle decision trace
    (dict sym [N, M] int8, margin [N, M] fp32, count [N]) belonging to `truth`.
    Returns a dict: utterances, identical, token agreement, WER of test measured against truth (= the WER delta of the
    two runs when truth stands in for the labels), and for every differing utterance the position and the margin."""
    assert len(test) == len(truth)
    identical = sum(1 for a, b in zip(test, truth) if list(a) == list(b))
    tok_total = sum(len(b) for b in truth)
    tok_edits = sum(levenshtein(list(a), list(b)) for a, b in zip(test, truth) if list(a) != list(b))
    wer, edits, words = word_error_rate([to_sentence(a) for a in test], [to_sentence(b) for b in truth])
    diffs = []
    for n, (a, b) in enumerate(zip(test, truth)):
        pos = _first_difference(list(a), list(b))
        if pos is None:
            continue
        entry = {"utterance": n, "first_differing_token": pos, "len_test": len(a), "len_truth": len(b)}
        if trace is not None:
            entry["margin"] = margin_at_divergence(trace["sym"][n], trace["margin"][n], trace["count"][n], pos)
        diffs.append(entry)
    out = {"utterances": len(truth), "identical_utterances": identical, "tokens_truth": tok_total,
           "token_edit_distance": tok_edits, "token_agreement": 1.0 - tok_edits / max(1, tok_total),
           "wer_vs_truth": wer, "word_edits": edits, "words_truth": words, "differences": diffs}
    if trace is not None:
        import numpy as np
        m = np.concatenate([trace["margin"][n, :int(c)] for n, c in enumerate(trace["count"])]) if len(truth) else np.zeros(0)
        out["margin_quantiles"] = {q: float(np.quantile(m, q)) for q in (0.0001, 0.001, 0.01, 0.1, 0.5)} if m.size else {}
        out["decisions"] = int(m.size)
        if margin_tol is not None:
            out["unexplained"] = [d for d in diffs if not d.get("margin", float("inf")) <= margin_tol]
    return out


def eval_acc(log_path, manifest_path, max_duration=15.0):
    """eval_accuracy.py:85-101: WER of an accuracy log against a LibriSpeech manifest."""
    with open(manifest_path) as fh:
        manifest = json.load(fh)
    references = [s["transcript"] for s in manifest if s["original_duration"] <= max_duration]
    hyp = load_accuracy_log(log_path)
    hypotheses = [to_sentence(hyp.get(i, [])) for i in range(len(references))]
    wer, _, _ = word_error_rate(hypotheses, references)
    print("Word Error Rate: {:}%, accuracy={:}%".format(wer * 100, (1 - wer) * 100))
    return wer


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--log_path", required=True)
    ap.add_argument("--manifest_path", required=True)
    ap.add_argument("--max_duration", type=float, default=15.0)
    a = ap.parse_args()
    eval_acc(a.log_path, a.manifest_path, a.max_duration)
