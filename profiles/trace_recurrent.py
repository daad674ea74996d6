"""Per-timestep timeline of CTA (0,0) of the recurrent kernel (clock64 stamps via rnnt_debug_set_recurrent_trace).
Usage: python profiles/trace_recurrent.py [N] [T]"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
# the cycle traces exist in the debug build only (make -C rnnt_inference_b200/csrc debug)
os.environ.setdefault("RNNT_B200_LIB", os.path.join(ROOT, "rnnt_inference_b200", "lib", "librnnt_b200_dbg.so"))
import numpy as np
import torch

from rnnt_inference_b200 import _lib, ops, synthetic as S
from rnnt_inference_b200.modeling_rnnt import RNNT

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
T = int(sys.argv[2]) if len(sys.argv) > 2 else 60
if len(sys.argv) > 4:
    _lib.check(_lib.lib.rnnt_debug_set_recurrent_flags(int(sys.argv[4])))
if len(sys.argv) > 3:
    _lib.check(_lib.lib.rnnt_tune_recurrent_cluster(int(sys.argv[3])))
torch.cuda.set_device(0)
rnnt = RNNT(S.make_checkpoint(), "quant", True).eval()
post = rnnt.transcription.post_rnn
w_ih, w_hh, _b, bias = post.weights[1]
x = torch.randint(-60, 61, (T, n, 1024), dtype=torch.int8, device="cuda")
trace = torch.zeros((T, 16), dtype=torch.int64, device="cuda")
names = ["h_ready", "chunk0_landed", "mma_issued", "acc_ready", "tmem_read", "math_stores_done", "publisher_saw_done", "published"]
for rep in range(3):
    hx = torch.zeros((n, 1024), dtype=torch.int8, device="cuda")
    cx = torch.zeros((n, 1024), dtype=torch.float16, device="cuda")
    if rep == 2:
        _lib.lib.rnnt_debug_set_recurrent_trace(ctypes.c_void_p(trace.data_ptr()))
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    ops.lstm_layer_i8(x, hx, cx, w_ih, w_hh, bias, float(post.rb_scale[1]), float(post.in_scale[1]), float(post.out_scale[1]), False)
    ev1.record()
    torch.cuda.synchronize()
    print(f"rep {rep}: layer {ev0.elapsed_time(ev1) * 1e3:.1f} us for T={T} N={n}")
_lib.lib.rnnt_debug_set_recurrent_trace(None)
tr = trace.cpu().numpy().astype(np.float64)
base = tr[:, 0:1]
rel = tr - base
for i, nm in enumerate(["mma: wait done_bar", "mma: wait full (sum of 8)", "epi: wait G", "epi: wait accumulator", "storer: wait staged"]):
    print(f"  [{nm}] {np.median(tr[5:, 8 + i]):9.0f} cycles")
print("cycles relative to h_ready of the same step (median over steps 5..):")
for i, nm in enumerate(names):
    print(f"  {nm:14s} {np.median(rel[5:, i]):9.0f}")
step = np.diff(tr[:, 0])
print("step period (cycles): median", np.median(step[5:]), "min", step[5:].min(), "max", step[5:].max())
print("published(t) -> h_ready(t+1):", np.median(tr[6:, 0] - tr[5:-1, 7]))
