"""Attribution of the projection GEMM's time to its pipelines: the same launch with operand loads / MMAs / result stores
switched off (debug build), for the two epilogues and K = 256, 1024, 2048.  Usage: python profiles/gemm_attrib.py"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("RNNT_B200_LIB", os.path.join(ROOT, "rnnt_inference_b200", "lib", "librnnt_b200_dbg.so"))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    import torch
    from rnnt_inference_b200 import _lib
    torch.cuda.set_device(0)
    m, n = 102400, 4096
    _lib.lib.rnnt_tune_gemm_pair(0)            # the switch-off flags exist in the single-CTA kernel
    for k in (256, 1024, 2048):
        a = torch.randint(-128, 128, (m, k), dtype=torch.int8, device="cuda")
        w = torch.randint(-128, 128, (n, k), dtype=torch.int8, device="cuda")
        c = torch.empty((m, n), dtype=torch.int32, device="cuda")
        row = []
        for flags in (0, 1, 2, 4, 5, 3, 6, 8, 9, 11):
            _lib.lib.rnnt_debug_set_gemm_flags(flags)
            ts = []
            for rep in range(4):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                _lib.check(_lib.lib.rnnt_gemm_i8_i32_gate_blocked_ragged(_lib.ptr(a), _lib.ptr(w), _lib.ptr(c), m, n, k, 0, None, _lib.stream_ptr()))
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            row.append((flags, round(min(ts[1:]), 3)))
        _lib.lib.rnnt_debug_set_gemm_flags(0)
        _lib.lib.rnnt_tune_gemm_pair(1)           # the CTA-pair kernel (gemm_i8_pair.cu), full pipeline only
        ts = []
        for rep in range(4):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            _lib.check(_lib.lib.rnnt_gemm_i8_i32_gate_blocked_ragged(_lib.ptr(a), _lib.ptr(w), _lib.ptr(c), m, n, k, 0, None, _lib.stream_ptr()))
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        _lib.lib.rnnt_tune_gemm_pair(0)
        row.append(("pair", round(min(ts[1:]), 3)))
        _lib.lib.rnnt_check_device_status(None)
        print(f"  K={k}: ms by flags (1 no loads, 2 no MMAs, 4 no stores, 8 every CTA stores onto one L2-resident tile of its own): {row}; algorithmic: tensor {2.0 * m * n * k / 4.74e15 * 1e3:.3f} ms, "
              f"HBM write {m * n * 4 / 6.5e12 * 1e3:.3f} ms")
else:
    for ep in ("tma",) if "--tma-only" in sys.argv else ("direct", "tma"):
        print("epilogue", ep, flush=True)
        env = dict(os.environ, RNNT_GEMM_EPILOGUE=ep)
        subprocess.run([sys.executable, os.path.abspath(__file__), "child"], env=env, check=False)
