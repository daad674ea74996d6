"""Per-iteration timeline of CTA 0 of the fused greedy-decode kernel (clock64 stamps via rnnt_debug_set_decode_trace).
Usage: python profiles/trace_decode.py [N] [max_frames]"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
# the cycle traces exist in the debug build only (make -C rnnt_inference_b200/csrc debug)
os.environ.setdefault("RNNT_B200_LIB", os.path.join(ROOT, "rnnt_inference_b200", "lib", "librnnt_b200_dbg.so"))
import numpy as np
import torch

from rnnt_inference_b200 import _lib, synthetic as S
from rnnt_inference_b200.decoder import GreedyDecoder
from rnnt_inference_b200.modeling_rnnt import RNNT

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
cap_frames = int(sys.argv[2]) if len(sys.argv) > 2 else 500
offset = int(sys.argv[3]) if len(sys.argv) > 3 else 0          # skip the `offset` longest utterances (second bench batch: 465 2048)
torch.cuda.set_device(0)
rnnt = RNNT(S.make_checkpoint(), "quant", True).eval()
dec = GreedyDecoder(rnnt, "quant", True, -1, n)
lens = np.sort(S.utterance_lengths(2513))[::-1][offset:offset + n]
lens = np.minimum(lens, cap_frames)
order = np.argsort(-S.utterance_lengths(2513), kind="stable")[offset:offset + n]      # the bench's utterances, not look-alikes
feats = [S.make_features(int(l), int(i)) for i, l in zip(order, lens)]
packed = torch.cat(feats, 0).cuda()
offsets = torch.tensor(np.concatenate([[0], np.cumsum(lens[:-1])]), dtype=torch.int64).cuda()
lens_t = torch.tensor(lens.copy(), dtype=torch.int32).cuda()
CAP = 2000
trace = torch.zeros((CAP, 16), dtype=torch.int64, device="cuda")
from rnnt_inference_b200 import ops
with torch.no_grad():
    for rep in range(3):
        if rep == 2:
            _lib.lib.rnnt_debug_set_decode_trace(ctypes.c_void_p(trace.data_ptr()), CAP)
            ops.telemetry.events = {}
        res, res_len = dec.forward_ragged(packed, offsets, lens_t, int(lens.max()), [int(x) for x in lens])
        torch.cuda.synchronize()
_lib.lib.rnnt_debug_set_decode_trace(None, 0)
ev = ops.telemetry.events
for k, v in ev.items():
    print(k, "ms", round(sum(s.elapsed_time(e) for s, e, _ in v), 3))
iters = int(dec.last_iters.item())
print("N", n, "T2", (int(lens.max()) + 1) // 2, "iters", iters, "tokens", int(res_len.sum()))
tr_all = trace.cpu().numpy().astype(np.float64)
rows = tr_all[tr_all[:, 0] > 0]
if len(rows):
    span = rows[-1, 6] - rows[0, 0]
    ph = [np.sum(rows[:, b] - rows[:, a]) for a, b in ((0, 1), (1, 2), (2, 3), (3, 4), (4, 5), (5, 6))]
    work = rows[rows[:, 7] > 0]
    idle = rows[rows[:, 7] == 0]
    print("rounds of CTA 0 (group 0): %d (%d with prediction-network work, %d blank-skipping only); first round start -> last round end: "
          "%.0f cycles" % (len(rows), len(work), len(idle), span))
    print("  sum over rounds (cycles): lstm0 %.0f  barrier %.0f  lstm1+barrier %.0f  fc1+barrier %.0f  joint/update %.0f  barrier %.0f; "
          "between rounds (state refresh, lists) %.0f" % (*ph, span - sum(ph)))
    gap = np.concatenate([rows[1:, 0] - rows[:-1, 6], [0.0]])       # end of a round -> start of the next one
    for lo, hi in ((1, 4), (5, 32), (33, 64), (65, 400)):
        m = (rows[:, 7] >= lo) & (rows[:, 7] <= hi)
        if m.any():
            r = rows[m]
            print("  rounds with %3d..%3d active lanes: %3d; mean cycles: lstm0 %.0f  barrier %.0f  lstm1+barrier %.0f  fc1+barrier %.0f  "
                  "joint/update %.0f (%.1f decisions of CTA 0)  barrier %.0f  gap to next round %.0f" %
                  (lo, hi, m.sum(), *[np.mean(r[:, b2] - r[:, a2]) for a2, b2 in ((0, 1), (1, 2), (2, 3), (3, 4), (4, 5))],
                   np.mean(r[:, 8]), np.mean(r[:, 6] - r[:, 5]), np.mean(gap[m])))
    if len(idle):
        print("  blank-skipping rounds: mean joint %.0f  barrier %.0f cycles" %
              (np.mean(idle[:, 5] - idle[:, 0]), np.mean(idle[:, 6] - idle[:, 5])))
tr = tr_all[:min(iters, CAP)]
full = tr[tr[:, 7] > 0]
d = lambda a, b: np.median(full[:, a] - full[:, b])
print("iterations with LSTM work:", len(full), "median active lanes in group 0:", np.median(full[:, 7]))
print("cycles (median): lstm0 %.0f  barrier %.0f  lstm1+barrier %.0f  fc1+barrier %.0f  joint/update %.0f  barrier %.0f" %
      (d(1, 0), d(2, 1), d(3, 2), d(4, 3), d(5, 4), d(6, 5)))
j = tr[tr[:, 8] > 0]
if len(j):
    print("joint of lane 0 (cycles, median): loads+fc1-epilogue %.0f  fc2 %.0f  argmax+update %.0f" %
          (np.median(j[:, 9] - j[:, 8]), np.median(j[:, 10] - j[:, 9]), np.median(j[:, 11] - j[:, 10])))
l0 = tr[(tr[:, 12] > 0) & (tr[:, 7] <= 12)]
if len(l0):
    print("lstm layer 0, single chunk (cycles, median): staging %.0f  dot %.0f  cell update %.0f" %
          (np.median(l0[:, 12] - l0[:, 0]), np.median(l0[:, 13] - l0[:, 12]), np.median(l0[:, 1] - l0[:, 13])))
per = np.diff(tr[:, 0])
print("iteration period (cycles): median %.0f mean %.0f" % (np.median(per), per.mean()))
print("iteration period quantiles 10/25/50/75/90/99:", [int(x) for x in np.percentile(per, [10, 25, 50, 75, 90, 99])])
print("active lanes (group 0) quantiles 10/50/90/max:", [int(x) for x in np.percentile(tr[:, 7], [10, 50, 90, 100])])
for lo, hi in ((0, 0), (1, 4), (5, 12), (13, 24), (25, 48), (49, 200)):
    m = (tr[1:, 7] >= lo) & (tr[1:, 7] <= hi)
    if m.any():
        print("  lanes %3d..%3d: %4d iterations, mean period %.0f cycles" % (lo, hi, m.sum(), per[m[:len(per)]].mean() if len(per) == len(m) else per[m[:-1]].mean()))
hv = tr[tr[:, 7] >= 49]
if len(hv):
    m = lambda a, b: np.mean(hv[:, a] - hv[:, b])
    print("heavy iterations (>= 49 active lanes in group 0): %d, mean lanes %.0f; mean cycles: lstm0 %.0f  barrier %.0f  lstm1+barrier %.0f  "
          "fc1+barrier %.0f  joint/update %.0f  barrier %.0f" % (len(hv), hv[:, 7].mean(), m(1, 0), m(2, 1), m(3, 2), m(4, 3), m(5, 4), m(6, 5)))
    h12 = hv[hv[:, 12] > 0]
    if len(h12):
        print("  first (full 12-lane) chunk of lstm0 in heavy iterations: staging %.0f  dot %.0f" %
              (np.mean(h12[:, 12] - h12[:, 0]), np.mean(h12[:, 13] - h12[:, 12])))
