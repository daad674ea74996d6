"""Audio front-end (features.AudioProcessing, csrc/frontend.cu) on the bench's utterance set: waveforms of the same durations
as the 2513 feature sequences (frames x 3 x 160 samples), processed in the SUT's batches; CUDA-event time per batch, the CPU
oracle on a bounded sample beside it.  Usage: python profiles/frontend_bench.py [out.md]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from oracle import frontend as F
from rnnt_inference_b200 import synthetic as S
from rnnt_inference_b200.features import AudioProcessing, MLPERF_INPUT_EVAL

torch.cuda.set_device(0)
lens_f = np.sort(S.utterance_lengths(2513))[::-1]
samples = (lens_f.astype(np.int64) * 3 - 1) * 160                     # -> floor(n / 160 + 1) = 3 * frames - ... spliced back to `frames`
proc = AudioProcessing("quant", pad_batch_size=True, **MLPERF_INPUT_EVAL).cuda().eval()
g = torch.Generator(device="cuda").manual_seed(0)
rows = []
total_ms = 0.0
for lo, hi in ((0, 2048), (2048, 2513)):
    n, lmax = hi - lo, int(samples[lo:hi].max())
    wav = torch.randn((n, lmax), device="cuda", generator=g) * 0.1
    wl = torch.tensor(samples[lo:hi].copy(), dtype=torch.int32, device="cuda")
    wav = wav * (torch.arange(lmax, device="cuda")[None, :] < wl[:, None])
    for _ in range(2):
        feats, fl = proc(wav, wl, True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        feats, fl = proc(wav, wl, True)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    total_ms += ms
    frames = int((wl // 160 + 1).sum())
    rows.append((n, lmax, frames, ms, int(wl.sum()) * 4, feats.numel() * 4))
    assert int(fl[:n].sum()) == int(lens_f[lo:hi].sum()), (int(fl[:n].sum()), int(lens_f[lo:hi].sum()))
# CPU oracle on a bounded sample: 32 utterances from the middle of the set
mid = slice(1200, 1232)
lw = samples[mid]
xw = (np.random.default_rng(0).normal(size=(32, int(lw.max()))) * 0.1).astype(np.float32)
for i, n in enumerate(lw):
    xw[i, n:] = 0
t0 = time.perf_counter()
F.filterbank_features(xw, lw, True)
cpu_s = time.perf_counter() - t0
out = ["# Audio front-end on one B200 (`profiles/frontend_bench.py`)", "",
       "Waveforms with the durations of the bench's 2513 utterances (mean 10.7 s), MLPerf RNN-T `input_eval` configuration; "
       "pre-emphasis + reflect pad -> fused STFT / power / mel / log -> splice x3 -> per-feature normalisation, fp32.", "",
       "| batch | longest (samples) | STFT frames | ms | waveform bytes read | feature bytes written |", "|---|---|---|---|---|---|"]
for n, lmax, frames, ms, rb, wb in rows:
    out.append(f"| {n} | {lmax} | {frames} | {ms:.2f} | {rb / 1e6:.0f} MB | {wb / 1e6:.0f} MB |")
frames_all = sum(r[2] for r in rows)
out += ["", f"Whole set: {total_ms:.2f} ms = {2513 / total_ms * 1e3:,.0f} utterances/s, {frames_all / total_ms / 1e3:.1f} M frames/s "
        f"(the RNN-T step itself takes ~40 ms for the same utterances).",
        f"CPU oracle (numpy, 1 thread, 32 mid-length utterances): {cpu_s:.2f} s = {32 / cpu_s:.1f} utterances/s."]
text = "\n".join(out)
print(text)
if len(sys.argv) > 1:
    open(sys.argv[1], "w").write(text + "\n")
