"""Greedy decode with ONE CTA of every lane group scanning the lane states late (debug build of the library, hook
rnnt_debug_set_decode_skew): rounds without prediction work have no group barrier between the scan and the decisions, so
the late CTA scans while its group is already taking the round's decisions.  The lane states are double-buffered by round
parity, hence tokens, counters and prediction state stay bit-exact against the oracle.  Runs in a subprocess because the
debug build is selected when the package loads (RNNT_B200_LIB)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DBG = os.path.join(ROOT, "rnnt_inference_b200", "lib", "librnnt_b200_dbg.so")

SCRIPT = r'''
import sys
import numpy as np
import torch
sys.path.insert(0, sys.argv[1])
from oracle import oracle as O
from tests import helpers as Hh
from rnnt_inference_b200 import _lib, ops
assert _lib.HAS_DEBUG_HOOKS
dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
bf = lambda a: dev(a).view(torch.bfloat16)
bits = lambda t: t.view(torch.int16).cpu().numpy().view(np.uint16)
for skew in (0, 30000):
    _lib.check(_lib.lib.rnnt_debug_set_decode_skew(skew))
    for n, t2, blank_bias, seed in [(150, 12, 2.5, 3), (40, 5, -3.0, 4), (300, 16, 1.0, 5)]:
        prm = Hh.make_decode_params(seed, blank_bias=blank_bias, scale=2.0)
        rng = np.random.default_rng(seed)
        f = O.f32_to_bf16_bits(rng.normal(0, 1.0, (t2, n, 1024)).astype(np.float32))
        f_lens = rng.integers(0, t2 + 1, n).astype(np.int32)
        f_lens[0] = t2
        L = 30 * t2 * 2
        y1 = O.linear_bf16_bf16(f, prm["w_trans"])
        so = Hh.fresh_decode_state(n, L)
        iters = O.greedy_decode(f, f_lens, prm, so["res"], so["res_idx"], so["symbols_added"], so["time_idx"], so["pre_g"],
                                so["pre_hg"], so["pre_cg"], y1_pre=y1)
        sg = Hh.fresh_decode_state(n, L)
        d = {k: (dev(v) if not isinstance(v, list) else None) for k, v in sg.items()}
        pre_hg = [bf(v) for v in sg["pre_hg"]]
        pre_cg = [dev(v) for v in sg["pre_cg"]]
        dw = ops.make_decode_weights(bf(prm["embed"]), [bf(w) for w in prm["pred_w_ih"]], [bf(w) for w in prm["pred_w_hh"]],
                                     [dev(b) for b in prm["pred_bias"]], bf(prm["w_pred"]), dev(prm["b1"]), bf(prm["w2"]),
                                     dev(prm["b2"]))
        iters_out = torch.zeros(1, dtype=torch.int32, device="cuda")
        ops.greedy_decode(bf(y1), dev(f_lens), dw, d["res"], d["res_idx"], d["symbols_added"], d["time_idx"], d["pre_g"],
                          pre_hg, pre_cg, iters_out, exact_order=True)
        _lib.check_device_status()
        assert int(iters_out.item()) == iters, (skew, n)
        np.testing.assert_array_equal(d["res_idx"].cpu().numpy(), so["res_idx"])
        np.testing.assert_array_equal(d["res"].cpu().numpy(), so["res"])
        np.testing.assert_array_equal(d["pre_g"].cpu().numpy(), so["pre_g"])
        np.testing.assert_array_equal(d["symbols_added"].cpu().numpy(), so["symbols_added"])
        np.testing.assert_array_equal(d["time_idx"].cpu().numpy(), so["time_idx"])
        for i in range(2):
            np.testing.assert_array_equal(bits(pre_hg[i]), so["pre_hg"][i])
        print("ok", skew, n, iters, flush=True)
print("DONE")
'''


def test_decode_is_exact_when_one_cta_scans_late():
    if not os.path.exists(DBG):
        pytest.skip("debug build of the library not present (make -C rnnt_inference_b200/csrc debug)")
    env = dict(os.environ, RNNT_B200_LIB=DBG)
    out = subprocess.run([sys.executable, "-c", SCRIPT, ROOT], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "DONE" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]
