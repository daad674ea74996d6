"""rnnt_lstm_recurrent_i8_static_scratch — the recurrent kernel with its scratch in a file-scope device array of the library
(no device-memory allocation inside the call; include/rnnt_b200.h) — against the existing entry point
rnnt_lstm_recurrent_i8 with a caller-owned workspace.  Integer / fp16 work: bit-exact.  Two calls in a row (the array keeps
the first call's contents), then the size limit.  Collected last on purpose (pytest -x)."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

H = 1024


def _inputs(seed, t, n, skip_quant_y):
    g = torch.Generator(device="cuda").manual_seed(seed)
    G = torch.randint(-60000, 60001, (t, n, 4 * H), dtype=torch.int32, device="cuda", generator=g)
    w_hh = torch.randint(-127, 128, (4 * H, H), dtype=torch.int8, device="cuda", generator=g)
    bias = (torch.rand((4 * H,), device="cuda", generator=g) - 0.5) * 2.0
    hx = torch.randint(-127, 128, (n, H), dtype=torch.int8, device="cuda", generator=g)
    cx = ((torch.rand((n, H), device="cuda", generator=g) - 0.5) * 4.0).to(torch.float16)
    y_dt = torch.bfloat16 if skip_quant_y else torch.int8
    return G, w_hh, bias, hx, cx, y_dt


def _run_pair(L, seed, t, n, skip_quant_y, status_host):
    """Both entry points on the same inputs; returns ((y, hx, cx) existing, (y, hx, cx) static-scratch) as byte tensors."""
    G, w_hh, bias, hx, cx, y_dt = _inputs(seed, t, n, skip_quant_y)
    rb, ins, outs = 1.0 / (60.0 * 127.0), 2.0e-4, 127.0
    ws = torch.empty(L.lib.rnnt_lstm_recurrent_workspace_bytes(n) + 256, dtype=torch.uint8, device="cuda")
    ws = ws[(-ws.data_ptr()) % 256:]
    hx_a, cx_a, hx_b, cx_b = hx.clone(), cx.clone(), hx.clone(), cx.clone()
    y_a = torch.zeros((t, n, H), dtype=y_dt, device="cuda")
    y_b = torch.zeros((t, n, H), dtype=y_dt, device="cuda")
    torch.cuda.synchronize()
    L.check(L.lib.rnnt_lstm_recurrent_i8(L.ptr(G), L.ptr(w_hh), L.ptr(bias), rb, ins, outs, int(skip_quant_y), L.ptr(hx_a), L.ptr(cx_a),
                                         L.ptr(y_a), t, n, H, L.ptr(ws), L.stream_ptr()))
    L.check_device_status()
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    status_host.fill_(-1)
    L.check(L.lib.rnnt_lstm_recurrent_i8_static_scratch(L.ptr(G), L.ptr(w_hh), L.ptr(bias), rb, ins, outs, int(skip_quant_y), L.ptr(hx_b),
                                                        L.ptr(cx_b), L.ptr(y_b), t, n, H, ctypes.c_void_p(status_host.data_ptr()),
                                                        L.stream_ptr()))
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    assert int(status_host.item()) == 0, "device-side status of the static-scratch call"
    as_bytes = lambda x: x.contiguous().view(torch.uint8)
    return (as_bytes(y_a), as_bytes(hx_a), as_bytes(cx_a)), (as_bytes(y_b), as_bytes(hx_b), as_bytes(cx_b)), free_before - free_after


def test_static_scratch_matches_workspace_entry_point_twice():
    from rnnt_inference_b200 import _lib as L
    torch.cuda.set_device(0)
    status_host = torch.zeros(1, dtype=torch.int32).pin_memory()
    # call 1: two batch tiles, one of them partial, int8 y; call 2 straight after it on the array the first call left dirty:
    # the largest N the scratch admits, bf16 y, another T
    grew = []
    for seed, t, n, skip in ((11, 9, 200, False), (12, 6, 512, True)):
        ref, got, d = _run_pair(L, seed, t, n, skip, status_host)
        for name, a, b in zip(("y", "hx", "cx"), ref, got):
            assert torch.equal(a, b), f"{name} differs (T={t}, N={n}, skip_quant_y={skip})"
        assert int(ref[0].to(torch.int32).abs().sum().item()) > 0
        grew.append(d)
    # the second call runs with every module already loaded: the driver's free-memory figure must not move
    assert grew[1] == 0, f"device memory in use changed by {grew[1]} bytes across the static-scratch call"


def test_static_scratch_rejects_sizes_beyond_the_array():
    from rnnt_inference_b200 import _lib as L
    torch.cuda.set_device(0)
    t, n = 2, 640
    G, w_hh, bias, hx, cx, y_dt = _inputs(3, t, n, False)
    y = torch.zeros((t, n, H), dtype=y_dt, device="cuda")
    rc = L.lib.rnnt_lstm_recurrent_i8_static_scratch(L.ptr(G), L.ptr(w_hh), L.ptr(bias), 1.0, 1.0, 1.0, 0, L.ptr(hx), L.ptr(cx), L.ptr(y),
                                                     t, n, H, None, L.stream_ptr())
    assert rc == -1                                       # RNNT_ERR_INVALID
    msg = L.lib.rnnt_last_error().decode()
    assert "512" in msg and "640" in msg, msg
    torch.cuda.synchronize()
    assert int(y.to(torch.int32).abs().sum().item()) == 0     # nothing ran
