"""Write-only HBM bandwidth on this GPU (profiling aid for DESIGN.md 4.1: what can the projection GEMM's 1.6 GB of
int32 output cost at best?).  Times fill / memset / copy of a 1.6 GB buffer with CUDA events."""
import torch

torch.cuda.set_device(0)
n = 400 * 1024 * 1024     # int32 elements = 1.68 GB
a = torch.empty(n, dtype=torch.int32, device="cuda")
b = torch.empty(n, dtype=torch.int32, device="cuda")


def timed(fn, reps=10):
    best = 1e9
    for _ in range(3):
        fn()
    for _ in range(reps):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        best = min(best, s.elapsed_time(e))
    return best


gb = n * 4 / 1e9
for name, fn, factor in (("fill_(1)   write only", lambda: a.fill_(1), 1), ("zero_()    write only (memset)", lambda: a.zero_(), 1),
                         ("copy_      read + write", lambda: b.copy_(a), 2), ("sum        read only", lambda: a.sum(), 1)):
    ms = timed(fn)
    print(f"{name:34s} {ms:7.3f} ms  {factor * gb / ms:7.2f} TB/s")
