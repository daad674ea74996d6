"""BASELINE config C5: long-utterance stress — 35 s utterances (T = 1167 frames) at batch 1 ... 2048, int8 encoder only:
the recurrent-step roofline map north_star asks for.  Per batch size: encoder ms, recurrent-kernel us per (layer,
timestep), algorithmic GB/s and fraction of the measured HBM bandwidth, int8 Top/s of the recurrent product and of the
projection GEMM.  Usage: python profiles/c5_sweep.py [out.md]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import bench
from rnnt_inference_b200 import _lib, ops

torch.cuda.set_device(0)
device = torch.device("cuda", 0)
_sd, decoder = bench.make_model(device)
trans = decoder.rnnt.transcription
peaks = bench.load_peaks()
T = 1167
rows = []
for n in (1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048):
    x = torch.randn((T, n, 240), device=device).clamp_(-5.8, 5.8)
    lens = torch.full((n,), T, dtype=torch.int32, device=device)

    def run():
        z = lambda dt: [torch.zeros((n, 1024), dtype=dt, device=device) for _ in range(5)]
        hx, cx = z(torch.int8), z(torch.float16)
        return trans(x, lens, hx[:2], cx[:2], hx[2:], cx[2:])

    with torch.no_grad():
        run()
        ops.telemetry.events = {}
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run()
        e1.record()
        torch.cuda.synchronize()
        ev = ops.telemetry.events
        ops.telemetry.events = None
    _lib.check_device_status()
    roof, item_us = bench.recurrent_roofline(ev, peaks)
    shares = bench.kernel_shares(ev)
    steps = 2 * T + 3 * ((T + 1) // 2)
    gemm_ops = 2.0 * n * (T * (4096 * 256 + 4096 * 1024) + ((T + 1) // 2) * (4096 * 2048 + 2 * 4096 * 1024))
    rows.append((n, e0.elapsed_time(e1), shares["lstm_recurrent"] * 1e3 / steps, roof["achieved"], roof["frac"], roof["tensor_top_s"],
                 shares["gemm_i8"], gemm_ops / (shares["gemm_i8"] * 1e-3) / 1e12))
    print(rows[-1], flush=True)
    del x
lines = ["# C5: 35 s utterances (T = 1167 frames), int8 encoder only, batch sweep (`profiles/c5_sweep.py`)", "",
         f"HBM peak {peaks.get('hbm_gbs')} GB/s ({peaks.get('_source')}); recurrent bytes = 23 552 B (24 576 B last layer) per (frame, utterance).", "",
         "| batch | encoder ms | recurrent us per (layer, timestep) | recurrent GB/s (algorithmic) | fraction of HBM | recurrent int8 Top/s | projection GEMM ms | GEMM Top/s |",
         "|---|---|---|---|---|---|---|---|"]
for r in rows:
    lines.append(f"| {r[0]} | {r[1]:.2f} | {r[2]:.2f} | {r[3]:.0f} | {r[4]:.3f} | {r[5]:.0f} | {r[6]:.2f} | {r[7]:.0f} |")
out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "c5_sweep.md")
open(out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
