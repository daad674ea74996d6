"""Host timeline of ONE Offline query through PytorchSUT (bench workload): where the end-to-end time beyond the device
time goes.  Usage: python profiles/diag_e2e_timeline.py [head_fraction]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import bench

hf = float(sys.argv[1]) if len(sys.argv) > 1 else None
torch.cuda.set_device(0)
device = torch.device("cuda", 0)
sd, decoder = bench.make_model(device)
batches, frames = bench.build_query(1, 0, bench.UTTS_PER_GPU, bench.BATCH)
feats = bench.features_for(batches)
sut, count = bench.make_sut(sd, batches, feats, device, bench.BATCH)
if hf is not None:
    sut.head_fraction = hf
for _ in range(2):
    bench.run_e2e_query(sut, count)
log = []
t_base = [0.0]


def wrap(name):
    fn = getattr(sut, name)

    def inner(*a, **k):
        t0 = time.perf_counter()
        r = fn(*a, **k)
        log.append((name, (t0 - t_base[0]) * 1e3, (time.perf_counter() - t0) * 1e3))
        return r
    setattr(sut, name, inner)


for m in ("_stage", "_launch", "_finish", "_respond"):
    wrap(m)
orig_issue = sut.issue_queries


def issue(samples):
    t0 = time.perf_counter()
    orig_issue(samples)
    log.append(("issue_queries", (t0 - t_base[0]) * 1e3, (time.perf_counter() - t0) * 1e3))


sut.sut.issue = issue
torch.cuda.synchronize()
t_base[0] = time.perf_counter()
r = bench.run_e2e_query(sut, count)
torch.cuda.synchronize()
total = (time.perf_counter() - t_base[0]) * 1e3
print(f"head_fraction {sut.head_fraction}: StartTest wall {total:.2f} ms, shim seconds {r['seconds'] * 1e3:.2f} ms, completed {r['completed']}")
for name, at, dur in sorted(log, key=lambda x: x[1]):
    print(f"  {at:8.2f} ms  {name:14s} {dur:7.2f} ms")
