"""Diagnostics of the tcgen05 bf16 GEMM against a float64 product (error size, where, which K block)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from rnnt_inference_b200 import ops, _lib

torch.cuda.set_device(0)
rng = np.random.default_rng(0)
m, n, k = 300, 512, 1024
x = torch.from_numpy(rng.uniform(-1, 1, (m, k)).astype(np.float32)).to(torch.bfloat16).cuda()
w = torch.from_numpy(rng.normal(0, 0.05, (n, k)).astype(np.float32)).to(torch.bfloat16).cuda()
ref64 = x.double() @ w.double().T
got = ops.gemm_bf16_bf16(x, w).double()
_lib.check_device_status()
simt = ops.linear_bf16_bf16(x, w).double()
err = (got - ref64).abs()
print("tcgen05 vs f64: max abs", err.max().item(), "mean abs", err.mean().item(), "| simt vs f64: max abs", (simt - ref64).abs().max().item())
rel = err / ref64.abs().clamp_min(1e-3)
bad = (rel > 2 ** -7).nonzero()
print("elements with rel err > 2^-7:", len(bad), "of", m * n)
if len(bad):
    rows = bad[:, 0].cpu().numpy(); cols = bad[:, 1].cpu().numpy()
    print(" rows%128 histogram (16 bins):", np.histogram(rows % 128, bins=16, range=(0, 128))[0].tolist())
    print(" cols%256 histogram (16 bins):", np.histogram(cols % 256, bins=16, range=(0, 256))[0].tolist())
    for r, c in bad[:10].tolist():
        print(f"  ({r},{c}) got {got[r, c].item():.6f} ref {ref64[r, c].item():.6f}")
# which K block is wrong: only block j of x non-zero
for j in range(0, 16):
    xj = torch.zeros_like(x)
    xj[:, 64 * j:64 * j + 64] = x[:, 64 * j:64 * j + 64]
    e = (ops.gemm_bf16_bf16(xj, w).double() - xj.double() @ w.double().T).abs().max().item()
    print(f" k-block {j}: max abs err {e:.3e}", end=";")
print()
# within a k block: 16-element sub-blocks
for s in range(4):
    xj = torch.zeros_like(x)
    xj[:, 16 * s:16 * s + 16] = x[:, 16 * s:16 * s + 16]
    e = (ops.gemm_bf16_bf16(xj, w).double() - xj.double() @ w.double().T).abs().max().item()
    print(f" k sub-block {s} (of block 0): max abs err {e:.3e}")
