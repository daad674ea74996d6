"""CPU tests that PIN the oracle: (1) against vectors produced by real AVX512-FP16 silicon
(tests/golden/postop_fp16_hw.npz, generator oracle/hwprobe/gen_postop_golden.cpp + tests/golden/make_golden_hw.sh),
(2) against restatements of the reference's own known-answer tests (mlperf_plugins/test/pyfiles/test_amx_linear.py)."""
import os

import numpy as np
import pytest

from oracle import oracle as O

GOLD = os.path.join(os.path.dirname(__file__), "golden", "postop_fp16_hw.npz")


@pytest.fixture(scope="module")
def hw():
    return np.load(GOLD)


def _half_inputs():
    bits = np.arange(65536, dtype=np.uint16)
    f = bits.view(np.float16)
    return bits, ~np.isnan(f)


def test_tanh_ph_every_binary16_input(hw):
    """_mm512_tanh_ph (el_common_intrin.hpp:95-123) over all 63488 non-NaN binary16 inputs == hardware."""
    bits, ok = _half_inputs()
    got = O.elementwise("tanh_fp16_h", bits, out_dtype=np.uint16)
    np.testing.assert_array_equal(got[ok], hw["tanh_all"][ok])


def test_sigmoid_ph_every_binary16_input(hw):
    """_mm512_sigmoid_ph (el_common_intrin.hpp:126-152) over all non-NaN binary16 inputs == hardware."""
    bits, ok = _half_inputs()
    x = bits.view(np.float16).astype(np.float32)
    got = O.elementwise("sigmoid_fp16", x, out_dtype=np.uint16)
    np.testing.assert_array_equal(got[ok], hw["sigmoid_all"][ok])


def test_fma_ph_random_triples(hw):
    """VFMADD132PH single rounding (used for c = fma(i, g, f*c), lstm_postop_tpp.cpp:54)."""
    a, b, c, want = hw["fma_a"], hw["fma_b"], hw["fma_c"], hw["fma_out"]
    got = O.elementwise("fma_fp16", a, b, c, out_dtype=np.uint16)
    nan = np.isnan(want.view(np.float16))
    np.testing.assert_array_equal(got[~nan], want[~nan])
    assert np.isnan(got[nan].view(np.float16)).all()


def test_lstm_postop_rows(hw):
    """lstm_postop_tpp::ref (lstm_postop_tpp.cpp:12-98) on 48 rows x 64 units with gate spreads 0.5 .. 14:
    new cell state, int8 h (in_scale), int8 y (out_scale) and bf16 y all bit-identical to the hardware run."""
    in_scale, out_scale = [float(v) for v in hw["scales"]]
    yq, hq, c_out = O.lstm_postop_fp16(hw["gates"], hw["c_in"], in_scale, out_scale, last=False)
    np.testing.assert_array_equal(c_out, hw["c_out"])
    np.testing.assert_array_equal(hq, hw["hq"])
    np.testing.assert_array_equal(yq, hw["yq"])
    ybf, hq2, _ = O.lstm_postop_fp16(hw["gates"], hw["c_in"], in_scale, out_scale, last=True)
    np.testing.assert_array_equal(ybf, hw["ybf"])
    np.testing.assert_array_equal(hq2, hw["hq"])


# ------------------------------------------------------------------------------------- reference KATs, restated
@pytest.mark.parametrize("bs", [32, 30, 33, 64])
def test_kat_amx_linear_i8o32(bs):
    """test_amx_linear.py:103-126 (2048 -> 4096, deterministic arange data): int8 GEMM == exact integer matmul."""
    k, n = 2048, 4096
    x = (np.arange(bs * k) % 256 - 128).astype(np.int8).reshape(bs, k)
    w = (np.arange(n * k) % 251 - 125).astype(np.int8).reshape(n, k)
    got = O.gemm_i8_i32(x, w)
    want = x.astype(np.int64) @ w.astype(np.int64).T
    np.testing.assert_array_equal(got, want.astype(np.int32))


def test_kat_amx_linear_padding():
    """test_amx_linear.py:129-152: K = 240 zero-padded to 256 gives the same accumulators."""
    rng = np.random.default_rng(0)
    x = rng.integers(-128, 128, (32, 240), dtype=np.int8)
    w = rng.integers(-128, 128, (1024, 240), dtype=np.int8)
    xp, wp = np.zeros((32, 256), np.int8), np.zeros((1024, 256), np.int8)
    xp[:, :240], wp[:, :240] = x, w
    np.testing.assert_array_equal(O.gemm_i8_i32(xp, wp), O.gemm_i8_i32(x, w))


def test_kat_linear_bf16_exact_small_integers():
    """test_amx_linear.py:155-211: with small-integer data every product and partial sum is exactly representable, so
    any accumulation order (AMX's or the oracle's) must give torch.matmul's fp32 result exactly; K = 121, out 29 -> 32."""
    rng = np.random.default_rng(1)
    x = rng.integers(-2, 3, (33, 121)).astype(np.float32)
    w = np.zeros((32, 121), np.float32)
    w[:29] = rng.integers(-2, 3, (29, 121))
    b = np.zeros(32, np.float32)
    b[:29] = rng.integers(-3, 4, 29)
    got = O.linear_bf16_f32(O.f32_to_bf16_bits(x), O.f32_to_bf16_bits(w), b)
    np.testing.assert_array_equal(got, x @ w.T + b)
    x1 = rng.integers(-2, 3, (33, 64)).astype(np.float32)
    w1 = rng.integers(-2, 3, (32, 64)).astype(np.float32)
    y = O.linear_bf16_accum_relu(O.f32_to_bf16_bits(x), O.f32_to_bf16_bits(w), O.f32_to_bf16_bits(x1), O.f32_to_bf16_bits(w1), b)
    want = np.maximum(x @ w.T + x1 @ w1.T + b, 0)
    assert np.abs(want).max() < 256 and np.abs(x @ w.T).max() < 256          # exactly representable in bf16
    np.testing.assert_array_equal(O.bf16_bits_to_f32(y), want)


@pytest.mark.parametrize("bs", [32, 64])
def test_kat_lstm_amx_int8_composition(bs):
    """test_amx_linear.py:214-295 (the hot-path pin): the fused layer == amx_linear_i8o32 x2 + lstm_postop per step,
    bit-equal y / h / c; 3 layers, T = 2, K = H = 1024, scales 0.01-like, last layer bf16."""
    h = 1024
    rng = np.random.default_rng(bs)
    x = rng.integers(-20, 21, (2, bs, h), dtype=np.int8)
    for layer in range(3):
        last = layer == 2
        w_ih = rng.integers(-30, 31, (4 * h, h), dtype=np.int8)
        w_hh = rng.integers(-30, 31, (4 * h, h), dtype=np.int8)
        bias = rng.normal(0, 50, 4 * h).astype(np.float32)
        rb, in_s, out_s = 1e-4, 100.0, 90.0
        hx = np.zeros((bs, h), np.int8)
        cx = np.zeros((bs, h), np.uint16)
        y = O.lstm_layer_i8(x, hx, cx, w_ih, w_hh, bias, rb, in_s, out_s, last)
        h2 = np.zeros((bs, h), np.int8)
        c2 = np.zeros((bs, h), np.uint16)
        ys = []
        for t in range(2):
            acc = O.gemm_i8_i32(x[t], w_ih) + O.gemm_i8_i32(h2, w_hh)        # int32 add first (tile_product_tpp.hpp:336-342)
            gates = (acc.astype(np.float32) + bias) * np.float32(rb)
            yt, h2, c2 = O.lstm_postop_fp16(gates, c2, in_s, out_s, last)
            ys.append(yt)
        np.testing.assert_array_equal(y, np.stack(ys))
        np.testing.assert_array_equal(hx, h2)
        np.testing.assert_array_equal(cx, c2)
        if not last:
            x = y
