"""The serial tail of the fused greedy decode (a lane capped on its LAST frame keeps emitting one symbol per lock-step
iteration until its batch's loop ends — the reference's quirk): per CTA of the lane group that holds such a lane, the
cycle breakdown of the trailing single-lane rounds.  Debug build.  Usage: python profiles/trace_decode_tail.py [N]"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("RNNT_B200_LIB", os.path.join(ROOT, "rnnt_inference_b200", "lib", "librnnt_b200_dbg.so"))
import numpy as np
import torch

from rnnt_inference_b200 import _lib, synthetic as S
from rnnt_inference_b200.decoder import GreedyDecoder
from rnnt_inference_b200.modeling_rnnt import RNNT

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
torch.cuda.set_device(0)
rnnt = RNNT(S.make_checkpoint(), "quant", True).eval()
dec = GreedyDecoder(rnnt, "quant", True, -1, n)
all_lens = S.utterance_lengths(2513)
order = np.argsort(-all_lens, kind="stable")[:n]
lens = all_lens[order]
packed = torch.cat([S.make_features(int(l), int(i)) for i, l in zip(order, lens)], 0).cuda()
offsets = torch.tensor(np.concatenate([[0], np.cumsum(lens[:-1])]), dtype=torch.int64).cuda()
lens_t = torch.tensor(lens.copy(), dtype=torch.int32).cuda()
host = [int(x) for x in lens]
CAP = 2000
trace = torch.zeros((CAP, 16), dtype=torch.int64, device="cuda")


def run(block):
    trace.zero_()
    _lib.lib.rnnt_debug_set_decode_trace_block(block)
    _lib.lib.rnnt_debug_set_decode_trace(ctypes.c_void_p(trace.data_ptr()), CAP)
    dec.forward_ragged(packed, offsets, lens_t, int(lens.max()), host)
    torch.cuda.synchronize()
    _lib.lib.rnnt_debug_set_decode_trace(None, 0)
    tr = trace.cpu().numpy().astype(np.float64)
    return tr[tr[:, 0] > 0]


with torch.no_grad():
    for _ in range(2):
        dec.forward_ragged(packed, offsets, lens_t, int(lens.max()), host)
    # the group with the longest trailing run of rounds with <= 2 active lanes
    best = (0, -1)
    for g in range(7):
        rows = run(20 * g)
        k = len(rows)
        while k > 0 and rows[k - 1, 7] <= 2:
            k -= 1
        best = max(best, (len(rows) - k, g))
    tail_len, g = best
    print(f"group {g}: {tail_len} trailing rounds with <= 2 active lanes")
    for s_ in range(20):
        rows = run(20 * g + s_)
        t = rows[len(rows) - tail_len:]
        m = lambda a, b: float(np.mean(t[:, a] - t[:, b]))
        dcs = float(np.mean(t[:, 8]))
        line = (f"  CTA {s_:2d}: round {float(np.mean(np.diff(t[:, 0]))):7.0f} cycles | lstm0 {m(1, 0):6.0f} bar {m(2, 1):6.0f} lstm1+bar {m(3, 2):6.0f} "
                f"fc1+bar {m(4, 3):6.0f} joint {m(5, 4):6.0f} bar {m(6, 5):6.0f} | decisions {dcs:.2f}")
        own = t[t[:, 9] > 0]
        if len(own):
            o = lambda a, b: float(np.mean(own[:, a] - own[:, b]))
            line += (f" | owner rounds {len(own)}: entry->loop {o(9, 4):6.0f} y staged {o(10, 9):6.0f} mma {o(11, 10):6.0f} "
                     f"update {o(14, 11):6.0f} exit {o(15, 14):6.0f} then barrier {o(6, 5):6.0f}")
        print(line)
_lib.lib.rnnt_debug_set_decode_trace_block(0)
