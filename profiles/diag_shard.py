"""One GPU runs, one after the other, the shard every rank of a WORLD-rank job would run (bench.py's query), and reports the
device status word and the step time of each - finds shard-dependent failures of a multi-GPU run without the box.
Usage: python profiles/diag_shard.py [world] [steps]"""
import importlib.util
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.argv, argv = [sys.argv[0]], sys.argv
import torch

spec = importlib.util.spec_from_file_location("bench", os.path.join(ROOT, "bench.py"))
bench = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bench)
from rnnt_inference_b200 import _lib

world = int(argv[1]) if len(argv) > 1 else 8
steps = int(argv[2]) if len(argv) > 2 else 3
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
_sd, decoder = bench.make_model(dev)
for rank in range(world):
    batches, _ = bench.build_query(world, rank, bench.UTTS_PER_GPU, bench.BATCH)
    feats = bench.features_for(batches)
    staged = bench.stage_on_device(batches, feats, dev)
    with torch.no_grad():
        for it in range(steps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            tok = bench.run_step_device(decoder, staged)
            torch.cuda.synchronize()
            ms = (time.perf_counter() - t0) * 1e3
            try:
                _lib.check_device_status()
                status = "ok"
            except RuntimeError as e:
                status = str(e)
                _lib.check(_lib.lib.rnnt_clear_device_status(_lib.stream_ptr()))
            print(f"rank {rank}/{world} step {it}: {ms:8.2f} ms  tokens {sum(int(t.item()) for t in tok)}  {status}", flush=True)
    del staged, feats
