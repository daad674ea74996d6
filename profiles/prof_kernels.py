"""Profiling driver (run under ncu on the GPU box): one int8 LSTM layer at the C2 shape (N=256, K=1024) for a short
sequence, and one greedy decode of a ragged batch.  Usage: python profiles/prof_kernels.py [lstm|decode|all] [T]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from rnnt_inference_b200 import ops, synthetic as S
from rnnt_inference_b200.decoder import GreedyDecoder
from rnnt_inference_b200.modeling_rnnt import RNNT

what = sys.argv[1] if len(sys.argv) > 1 else "all"
T = int(sys.argv[2]) if len(sys.argv) > 2 else 100
torch.cuda.set_device(0)
rnnt = RNNT(S.make_checkpoint(), "quant", True).eval()
dec = GreedyDecoder(rnnt, "quant", True, -1, 1024)
n = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
with torch.no_grad():
    if what in ("lstm", "all"):
        post = rnnt.transcription.post_rnn
        w_ih, w_hh, _b, bias = post.weights[1]
        x = torch.randint(-60, 61, (T, n, 1024), dtype=torch.int8, device="cuda")
        for _ in range(3):
            hx = torch.zeros((n, 1024), dtype=torch.int8, device="cuda")
            cx = torch.zeros((n, 1024), dtype=torch.float16, device="cuda")
            ops.lstm_layer_i8(x, hx, cx, w_ih, w_hh, bias, float(post.rb_scale[1]), float(post.in_scale[1]),
                              float(post.out_scale[1]), False)
        torch.cuda.synchronize()
    if what in ("decode", "all"):
        lens = np.sort(S.utterance_lengths(2513))[::-1][1000:1000 + n]
        lens = np.minimum(lens, 2 * T)
        feats = [S.make_features(l, i) for i, l in enumerate(lens)]
        packed = torch.cat(feats, 0).cuda()
        offsets = torch.tensor(np.concatenate([[0], np.cumsum(lens[:-1])]), dtype=torch.int64).cuda()
        lens_t = torch.tensor(lens.copy(), dtype=torch.int32).cuda()
        for _ in range(3):
            res, res_len = dec.forward_ragged(packed, offsets, lens_t, int(lens.max()))
        torch.cuda.synchronize()
        print("decode tokens", int(res_len.sum()), "iters", int(dec.last_iters.item()), "T2", (int(lens.max()) + 1) // 2)
from rnnt_inference_b200 import _lib
_lib.check_device_status()
print("ok")
