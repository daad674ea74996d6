"""Gate-blocked projection layout check: rnnt_gemm_i8_i32_gate_blocked_ragged against the CUDA-core GEMM, element by
element, with the mismatch pattern by slice / gate / octet / row / unit.  Usage: python profiles/diag_gate_blocked.py"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rnnt_inference_b200 import _lib, ops
torch.cuda.set_device(0)
for m in (96, 128, 256, 4096):
    k, n = 256, 4096
    g = torch.Generator(device="cuda").manual_seed(1)
    a = torch.randint(-128, 128, (m, k), dtype=torch.int8, device="cuda", generator=g)
    w = torch.randint(-128, 128, (n, k), dtype=torch.int8, device="cuda", generator=g)
    ref = ops.gemm_i8_i32(a, w, simt=True)           # [m][4096], col = gate*1024 + slice*32 + octet*8 + u
    c = torch.full((m * n,), 123456789, dtype=torch.int32, device="cuda")
    _lib.check(_lib.lib.rnnt_gemm_i8_i32_gate_blocked_ragged(_lib.ptr(a), _lib.ptr(w), _lib.ptr(c), m, n, k, 0, None, _lib.stream_ptr()))
    _lib.check_device_status()
    got = c.view(32, 4, 4, m, 8)                        # [slice][gate][octet][row][u]
    exp = ref.view(m, 4, 32, 4, 8).permute(2, 1, 3, 0, 4)   # [slice][gate][octet][row][u]
    bad = (got != exp)
    print("M", m, "mismatch", int(bad.sum()), "of", bad.numel(), "untouched", int((got == 123456789).sum()))
    if bad.any():
        idx = bad.nonzero()[:8].tolist()
        print(" first bad [slice,gate,octet,row,u]:", idx)
        for ax, name in enumerate(["slice", "gate", "octet", "row", "u"]):
            dims = [d for d in range(5) if d != ax]
            print("  bad by", name, bad.sum(dim=dims).tolist()[:40])
