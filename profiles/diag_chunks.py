"""Device time of the two chunks of the bench query for several cuts of its short end (sharding.plan_chunks): per chunk the
kernel times of one forward_ragged.  Usage: python profiles/diag_chunks.py k1 k2 ...   (k = utterances in the head chunk)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from rnnt_inference_b200 import ops, synthetic as S
from rnnt_inference_b200.decoder import GreedyDecoder
from rnnt_inference_b200.modeling_rnnt import RNNT

torch.cuda.set_device(0)
rnnt = RNNT(S.make_checkpoint(), "quant", True).eval()
dec = GreedyDecoder(rnnt, "quant", True, -1, 2048)
all_lens = S.utterance_lengths(2513)
order = np.argsort(-all_lens, kind="stable")
feats = {int(i): S.make_features(int(all_lens[i]), int(i)) for i in order}


def run(idx):
    lens = all_lens[idx]
    n_pad = (len(idx) + 31) // 32 * 32
    packed = torch.cat([feats[int(i)] for i in idx], 0).cuda()
    offs = np.zeros(n_pad, np.int64); offs[:len(idx)] = np.concatenate([[0], np.cumsum(lens[:-1])])
    ln = np.zeros(n_pad, np.int32); ln[:len(idx)] = lens
    offsets, lens_t = torch.tensor(offs).cuda(), torch.tensor(ln).cuda()
    host = [int(x) for x in ln]
    out = None
    with torch.no_grad():
        for rep in range(3):
            if rep == 2:
                ops.telemetry.events = {}
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            res, res_len = dec.forward_ragged(packed, offsets, lens_t, int(lens.max()), host)
        e1.record()
        torch.cuda.synchronize()
    ev = ops.telemetry.events
    ops.telemetry.events = None
    k = {name: round(sum(s.elapsed_time(e) for s, e, _ in v), 2) for name, v in ev.items()}
    return round(e0.elapsed_time(e1), 2), k, int(dec.last_iters.item()), int(res_len.sum())


for k in [int(a) for a in sys.argv[1:]]:
    head, tail = order[len(order) - k:], order[:len(order) - k]
    for name, idx in (("head", head), ("tail", tail)):
        if len(idx) > 2048:
            continue
        ms, kern, iters, toks = run(idx)
        print(f"cut {k:5d} {name}: N {len(idx):5d} Tmax {int(all_lens[idx].max()):4d} total {ms:7.2f} ms  kernels {kern}  decode iters {iters} tokens {toks}", flush=True)
