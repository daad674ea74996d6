"""bf16 (prediction / joint) tier against golden vectors produced by REAL AMX-BF16 / VRCP14PS hardware
(tests/golden/amx_bf16_hw.npz, generator oracle/hwprobe/gen_amx_bf16_golden.cpp, recipe make_golden_amx.sh).

TDPBF16PS's internal accumulation and VRCP14PS's table cannot be reproduced bit for bit off the chip, so
this tier is pinned by MEASURED tolerances, the same for the CPU oracle and for the CUDA kernels:
  tile products (fp32)          |d| <= 4e-6 * (1 + |x|)        measured: oracle 8.4e-7 abs at |acc| ~ 0.9
  bf16-rounded outputs          <= 1 bf16 ulp, <= 0.1 % of the elements differ   measured: oracle 1 / 16384
  LSTM cell c (VRCP14PS gates)  |d| <= 2e-4                    measured: oracle 5.5e-5
  LSTM cell h (bf16)            |d| <= 2e-4 + 1 bf16 ulp       measured: oracle 120 / 10240 elements differ
  fc2 logits                    |d| <= 1e-5, same argmax       measured: oracle 2.4e-6
"""
import os

import numpy as np
import pytest

from oracle import oracle as O
from tests import amx_inputs as A

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "amx_bf16_hw.npz"))


def bf(b):
    return O.bf16_bits_to_f32(b)


def check_bf16(name, got, want, max_frac=1e-3):
    got, want = np.asarray(got, np.uint16), np.asarray(want, np.uint16)
    diff = got != want
    ulp = np.abs(got.astype(np.int32) - want.astype(np.int32))
    near_zero = np.abs(bf(got) - bf(want)) <= 2e-5          # a sum that cancels to ~0: order differences are many ulps of it
    assert (ulp[~near_zero] <= 1).all(), f"{name}: {ulp[~near_zero].max()} bf16 ulps"
    assert diff.mean() <= max_frac, f"{name}: {diff.sum()} of {diff.size} elements differ"


def check_cell(hy, cy):
    assert np.abs(cy - GOLD["lstm_cy"]).max() <= 2e-4
    want = bf(GOLD["lstm_hy"])
    tol = 2e-4 + np.abs(want) * 2.0 ** -7
    assert (np.abs(bf(hy) - want) <= tol).all()


def test_generator_twin_spot_values():
    i = A.inputs()     # the values gen_amx_bf16_golden.cpp prints
    assert i["x"].ravel()[7] == 0xBF58 and i["w_ih"].ravel()[12345] == 0x3D2C
    assert float(i["bias"][100]).hex() == "0x1.20496c0000000p-2" and float(i["c"].ravel()[999]).hex() == "0x1.cc17040000000p+0"


def test_oracle_tile_product_vs_amx():
    i = A.inputs()
    # G_ih of the LSTM cell through the oracle's fp32 linear (no bias)
    g = O.linear_bf16_f32(i["x"], i["w_ih"], np.zeros(4 * A.P, np.float32))
    want = GOLD["lstm_g_ih"]
    assert (np.abs(g - want) <= 4e-6 * (1 + np.abs(want))).all()


def test_oracle_lstm_cell_vs_amx():
    i = A.inputs()
    check_cell(*O.lstm_cell_bf16(i["x"], i["h"], i["c"], i["w_ih"], i["w_hh"], i["bias"]))


def test_oracle_joint_vs_amx():
    i = A.inputs()
    check_bf16("y1", O.linear_bf16_bf16(i["f"], i["w_trans"]), GOLD["fc1_y1"])
    check_bf16("y", O.linear_bf16_accum_relu(i["f"], i["w_trans"], i["g"], i["w_pred"], i["b1"]), GOLD["fc1_y"])
    check_bf16("y | hw y1", O.linear_bf16_accum_relu(None, None, i["g"], i["w_pred"], i["b1"], y1_pre=GOLD["fc1_y1"]), GOLD["fc1_y"])
    lg = O.linear_bf16_f32(GOLD["fc1_y"], i["w2"], i["b2"])
    assert np.abs(lg - GOLD["fc2_logits"]).max() <= 1e-5
    assert (lg.argmax(1) == GOLD["fc2_logits"].argmax(1)).all()


# ---------------------------------------------------------------------------------------------- the CUDA kernels
@pytest.mark.gpu
def test_gpu_bf16_ops_vs_amx():
    import torch
    from rnnt_inference_b200 import ops
    i = A.inputs()
    dev = torch.device("cuda", 0)
    tb = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.int16)).to(dev).view(torch.bfloat16)
    tf = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    nb = lambda t: t.view(torch.int16).cpu().numpy().view(np.uint16)
    hy, cy = ops.lstm_cell_bf16(tb(i["x"]), tb(i["h"]), tf(i["c"]), tb(i["w_ih"]), tb(i["w_hh"]), tf(i["bias"]))
    check_cell(nb(hy), cy.cpu().numpy())
    # hoisted f . W_trans on the tensor cores (tcgen05 kind::f16) and the exact-order CUDA-core operator
    check_bf16("y1 tcgen05", nb(ops.gemm_bf16_bf16(tb(i["f"]), tb(i["w_trans"]))), GOLD["fc1_y1"])
    check_bf16("y1 simt", nb(ops.linear_bf16_bf16(tb(i["f"]), tb(i["w_trans"]))), GOLD["fc1_y1"])
    y = ops.linear_bf16_accum_relu(tb(GOLD["fc1_y1"]), tb(i["g"]), tb(i["w_pred"]), tf(i["b1"]))
    check_bf16("y", nb(y), GOLD["fc1_y"])
    lg = ops.linear_bf16_f32(tb(GOLD["fc1_y"]), tb(i["w2"]), tf(i["b2"])).cpu().numpy()
    assert np.abs(lg - GOLD["fc2_logits"]).max() <= 1e-5
    assert (lg.argmax(1) == GOLD["fc2_logits"].argmax(1)).all()
