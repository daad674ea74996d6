#!/usr/bin/env python
"""Statistics of the synthetic decode workload, produced by the CPU ORACLE on the bench set (2513 utterances):
tokens per stacked frame, tokens vs T2, where in the utterance tokens are emitted, distinct transcripts, 30-symbol-cap
events (and how many fall on an utterance's last frame = quirk (b) of SURVEY 8(a)), active lanes per lock-step
iteration.  Writes profiles/r02/decode_workload.md.  Test infrastructure (uses oracle/).

    python profiles/decode_workload.py [--cache DIR] [--chunk 128]

The int8 encoder of the oracle takes ~25 minutes for the whole set on 8 cores; --cache keeps its outputs (bf16 encoder
frames per chunk of length-sorted utterances) so that the decode-side constants of synthetic.py can be re-tuned
without re-encoding (the encoder weights do not depend on them)."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as O                      # noqa: E402
from oracle import pipeline as PL                   # noqa: E402
from rnnt_inference_b200 import synthetic as S      # noqa: E402


def encoded_chunks(params, cache, chunk):
    lens_all = S.utterance_lengths(2513)
    order = np.argsort(-lens_all, kind="stable")
    for c in range(0, len(order), chunk):
        idx = order[c:c + chunk]
        path = os.path.join(cache, f"chunk_{c:05d}.npz") if cache else None
        if path and os.path.exists(path):
            d = np.load(path)
            yield d["f"], d["lens"]
            continue
        lens = lens_all[idx]
        x, xl = PL.pad_batch([S.make_features(n, i) for i, n in zip(idx, lens)], pad_to=0)
        f = PL.encode(params, x, xl)
        if path:
            os.makedirs(cache, exist_ok=True)
            np.savez(path, f=f, lens=lens, idx=idx)
        yield f, lens


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cache", default=None)
    ap.add_argument("--chunk", type=int, default=128)
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r02", "decode_workload.md"))
    a = ap.parse_args()
    params = PL.prepare_params(S.make_checkpoint())
    per_frame_hist = np.zeros(32, np.int64)
    quart = np.zeros(4, np.int64)
    tok_counts, t2s, transcripts, active = [], [], set(), []
    caps = caps_last = utts = 0
    margins = []
    for f, lens in encoded_chunks(params, a.cache, a.chunk):
        n = f.shape[1]
        f_lens = ((lens + 1) // 2).astype(np.int32)
        ds = PL.new_decode_state(n, int(lens.max()))
        tr = O.new_decision_trace(n, 4 * int(f_lens.max()) + 256)
        iters = O.greedy_decode(f, f_lens, params["dec"], ds["res"], ds["res_idx"], ds["symbols_added"], ds["time_idx"],
                                ds["pre_g"], ds["pre_hg"], ds["pre_cg"], trace=tr)
        act = np.zeros(iters + 1, np.int64)
        for i in range(n):
            c = int(tr["count"][i])
            sym = tr["sym"][i, :c]
            margins.append(tr["margin"][i, :c])
            frame = emitted_here = 0
            per = np.zeros(int(f_lens[i]), np.int64)
            for k in range(c):          # decisions taken while the lane was unfinished, in lock-step order
                if sym[k] != 28 and emitted_here != 30:
                    per[frame] += 1
                    emitted_here += 1
                    act[k] += 1
                else:
                    frame += 1
                    emitted_here = 0
            per_frame_hist += np.bincount(np.minimum(per, 31), minlength=32)
            caps += int((per >= 30).sum())
            caps_last += int(per[-1] >= 30)
            pos = np.repeat(np.arange(len(per)), per) / len(per)
            quart += np.histogram(pos, bins=4, range=(0, 1))[0]
            ntok = int(ds["res_idx"][i]) + 1
            tok_counts.append(ntok)
            t2s.append(int(f_lens[i]))
            transcripts.add(tuple(ds["res"][i, :ntok]))
        active.append(act[:iters])
        utts += n
        print(f"{utts} utterances", file=sys.stderr, flush=True)
    tok_counts, t2s = np.array(tok_counts), np.array(t2s)
    active = np.concatenate(active)
    m = np.concatenate(margins)
    lines = [
        "# Synthetic decode workload of the bench set (CPU oracle, `profiles/decode_workload.py`)", "",
        f"Checkpoint `synthetic.make_checkpoint()`: blank bias {S.DEFAULT_BLANK_BIAS}, trace memory {S.DEFAULT_MEMORY}, penalty "
        f"{S.DEFAULT_PENALTY}, burst threshold {S.DEFAULT_BURST_THRESHOLD} (seed {S.DEFAULT_BURST_SEED}); 2513 utterances decoded in "
        f"length-sorted batches of {a.chunk}.", "",
        f"* utterances {utts}, stacked frames {int(t2s.sum())}, tokens {int(tok_counts.sum())}: **{tok_counts.sum() / t2s.sum():.3f} tokens per "
        f"stacked frame** (mean {tok_counts.mean():.1f} per utterance, min {tok_counts.min()}, max {tok_counts.max()})",
        f"* correlation(tokens, T2) = {np.corrcoef(tok_counts, t2s)[0, 1]:.3f}; distinct transcripts {len(transcripts)} of {utts}",
        f"* frames emitting at least one token: {1 - per_frame_hist[0] / per_frame_hist.sum():.3f}",
        f"* tokens by quarter of the utterance (first .. last): {quart.tolist()}",
        f"* frames that hit the 30-symbol cap: {caps} ({caps / per_frame_hist.sum():.2e} of the frames), of which on an utterance's LAST frame "
        f"(quirk b: the lane keeps emitting until its batch ends): {caps_last}",
        f"* accepted symbols (= prediction-network evaluations) per lock-step iteration, batches of {a.chunk}: mean {active.mean():.1f}, "
        f"median {int(np.median(active))}, p90 {int(np.percentile(active, 90))}, max {int(active.max())}",
        f"* top-2 logit margin over all {m.size} decisions: " + ", ".join(f"q{q} = {np.quantile(m, q):.4f}" for q in (0.0001, 0.001, 0.01, 0.1, 0.5)),
        "", "Tokens emitted per stacked frame (histogram, last bin = 31+):", "",
        "| tokens | " + " | ".join(str(i) for i in range(32)) + " |", "|---|" + "---|" * 32,
        "| frames | " + " | ".join(str(int(v)) for v in per_frame_hist) + " |", ""]
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    open(a.out, "w").write("\n".join(lines))
    print("\n".join(lines))


if __name__ == "__main__":
    main()
