"""Per-operator CUDA-event timing of the audio front-end on the bench's 2048-utterance batch as waveforms
(profiles/r02/frontend_bench.md).  Usage: python profiles/frontend_breakdown.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from rnnt_inference_b200 import synthetic as S, ops, _C as P
from rnnt_inference_b200.features import AudioProcessing, MLPERF_INPUT_EVAL
torch.cuda.set_device(0)
lens_f = np.sort(S.utterance_lengths(2513))[::-1]
samples = (lens_f.astype(np.int64) * 3 - 1) * 160
proc = AudioProcessing("quant", pad_batch_size=True, **MLPERF_INPUT_EVAL).cuda().eval()
f = proc.featurizer
n, lmax = 2048, int(samples[:2048].max())
wav = torch.randn((n, lmax), device="cuda") * 0.1
wl = torch.tensor(samples[:2048].copy(), dtype=torch.int32, device="cuda")
def timed(name, fn):
    for _ in range(2): r = fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): r = fn()
    e1.record(); torch.cuda.synchronize()
    print(f"{name:12s} {e0.elapsed_time(e1)/5:7.2f} ms  out {tuple(r.shape) if hasattr(r,'shape') else ''}")
    return r
y = timed("preemphasis", lambda: P.preemphasis(wav, wl, coeff=0.97, pad_size=256))
fl = torch.floor(wl / 160 + 1).to(torch.int32)
lm = timed("logmel", lambda: ops.logmel(y, fl, f.window, f.fb[0], 512, 160, 1e-10, 1e-20))
sp = timed("splice", lambda: P.frame_splicing(lm, fl, 3))
fl2 = torch.ceil(fl / 3).to(torch.int32)
shape = torch.tensor([2048, 256, 1680], dtype=torch.int32)
timed("layernorm", lambda: P.i_layernorm_pad(sp, None, None, fl2, 1e-12, unbiased=1, output_shape=shape)[0])
timed("whole", lambda: proc(wav, wl, True)[0])
