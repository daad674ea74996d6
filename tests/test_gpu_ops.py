"""GPU parity tests of the C-ABI operators against the CPU oracle (bit-exact unless stated)."""
import numpy as np
import pytest
import torch

from oracle import oracle as O
from tests import helpers as Hh

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from rnnt_inference_b200 import ops as _ops
    return _ops


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def bf(a_bits):
    return torch.from_numpy(np.ascontiguousarray(a_bits).view(np.int16)).cuda().view(torch.bfloat16)


def bits(t):
    return t.contiguous().view(torch.int16).cpu().numpy().view(np.uint16)


def test_quantize_pad(ops):
    rng = np.random.default_rng(0)
    x = (rng.normal(0, 2.5, (7, 33, 240))).astype(np.float32)
    x.flat[:8] = [0.5 / 21.7, 1.5 / 21.7, 2.5 / 21.7, -0.5 / 21.7, 1e9, -1e9, 5.8358, -5.8358]
    scale = 127.0 / 5.8358
    got = ops.quantize_pad(dev(x), scale, 256).cpu().numpy()
    np.testing.assert_array_equal(got, O.quantize_pad(x, scale, 256))


@pytest.mark.parametrize("m,n,k", [(1, 4096, 256), (300, 4096, 256), (1000, 4096, 1024), (257, 4096, 2048), (130, 192, 64),
                                   (64, 128, 1040)])
def test_gemm_vs_oracle(ops, m, n, k):
    rng = np.random.default_rng(m + n + k)
    a = rng.integers(-128, 128, (m, k), dtype=np.int8)
    w = rng.integers(-128, 128, (n, k), dtype=np.int8)
    got = ops.gemm_i8_i32(dev(a), dev(w))
    from rnnt_inference_b200 import _lib
    _lib.check_device_status()
    np.testing.assert_array_equal(got.cpu().numpy(), O.gemm_i8_i32(a, w))


def test_gemm_large_vs_simt(ops):
    g = torch.Generator(device="cuda").manual_seed(1)
    m, n, k = 40000, 4096, 2048
    a = torch.randint(-128, 128, (m, k), dtype=torch.int8, device="cuda", generator=g)
    w = torch.randint(-128, 128, (n, k), dtype=torch.int8, device="cuda", generator=g)
    got = ops.gemm_i8_i32(a, w)
    ref = ops.gemm_i8_i32(a, w, simt=True)
    from rnnt_inference_b200 import _lib
    _lib.check_device_status()
    assert torch.equal(got, ref)


@pytest.mark.parametrize("t,n,c", [(5, 3, 1024), (6, 40, 1024), (1, 2, 64)])
def test_stack_time(ops, t, n, c):
    rng = np.random.default_rng(t * n)
    x = rng.integers(-128, 128, (t, n, c), dtype=np.int8)
    lens = rng.integers(0, t + 1, n).astype(np.int32)
    lens[0] = t
    got = ops.stack_time(dev(x), dev(lens), 2).cpu().numpy()
    np.testing.assert_array_equal(got, O.stack_time(x, lens, 2))


@pytest.mark.parametrize("t,n,k,last", [(3, 32, 256, False), (4, 17, 1024, False), (3, 160, 1024, True), (2, 300, 2048, False),
                                        (5, 128, 1024, True), (6, 640, 256, False), (5, 1000, 256, True),
                                        (40, 256, 256, False)])
def test_lstm_layer_i8(ops, t, n, k, last):
    """The hot-path pin restated from test_amx_linear.py:214-295 (fused layer vs composed ops), here GPU vs oracle:
    y / hx / cx must be bit-identical."""
    lp = Hh.make_lstm_layer(seed=k + n, k_in=k)
    rng = np.random.default_rng(t + n + k)
    x = rng.integers(-100, 101, (t, n, k), dtype=np.int8)
    hx0 = rng.integers(-127, 128, (n, 1024), dtype=np.int8)
    cx0 = rng.normal(0, 0.5, (n, 1024)).astype(np.float16)
    hx_o, cx_o = hx0.copy(), cx0.view(np.uint16).copy()
    y_o = O.lstm_layer_i8(x, hx_o, cx_o, lp["w_ih"], lp["w_hh"], lp["bias"], lp["rb_scale"], lp["in_scale"], lp["out_scale"], last)
    hx, cx = dev(hx0), dev(cx0)
    y = ops.lstm_layer_i8(dev(x), hx, cx, dev(lp["w_ih"]), dev(lp["w_hh"]), dev(lp["bias"]), lp["rb_scale"], lp["in_scale"],
                          lp["out_scale"], last)
    from rnnt_inference_b200 import _lib
    _lib.check_device_status()
    y_np = bits(y) if last else y.cpu().numpy()
    np.testing.assert_array_equal(y_np, y_o)
    np.testing.assert_array_equal(hx.cpu().numpy(), hx_o)
    np.testing.assert_array_equal(cx.cpu().numpy().view(np.uint16), cx_o)


@pytest.mark.parametrize("cluster", [1, 2, 4, 8])
def test_lstm_layer_i8_ragged_tiles(ops, cluster):
    """Ragged batch sorted by length: tile b runs tile_steps[b] steps.  Each 128-row tile must equal the oracle run on that
    tile alone for its own number of steps (y beyond it stays zero), for every cluster size of the recurrent kernel."""
    from rnnt_inference_b200 import _lib
    t, n, k = 7, 520, 256
    steps = np.array([7, 6, 4, 4, 1], np.int32)
    lp = Hh.make_lstm_layer(seed=77, k_in=k)
    rng = np.random.default_rng(cluster + 5)
    x = rng.integers(-100, 101, (t, n, k), dtype=np.int8)
    hx0 = rng.integers(-127, 128, (n, 1024), dtype=np.int8)
    cx0 = rng.normal(0, 0.5, (n, 1024)).astype(np.float16)
    y_o = np.zeros((t, n, 1024), np.int8)
    hx_o, cx_o = hx0.copy(), cx0.view(np.uint16).copy()
    for b, s in enumerate(steps):
        r = slice(128 * b, min(n, 128 * b + 128))
        h_t, c_t = np.ascontiguousarray(hx_o[r]), np.ascontiguousarray(cx_o[r])
        y_o[:s, r] = O.lstm_layer_i8(np.ascontiguousarray(x[:s, r]), h_t, c_t, lp["w_ih"], lp["w_hh"], lp["bias"], lp["rb_scale"],
                                     lp["in_scale"], lp["out_scale"], False)
        hx_o[r], cx_o[r] = h_t, c_t
    _lib.check(_lib.lib.rnnt_debug_set_recurrent_cluster(cluster))
    try:
        hx, cx = dev(hx0), dev(cx0)
        y = ops.lstm_layer_i8(dev(x), hx, cx, dev(lp["w_ih"]), dev(lp["w_hh"]), dev(lp["bias"]), lp["rb_scale"], lp["in_scale"],
                              lp["out_scale"], False, tile_steps=steps)
        _lib.check_device_status()
    finally:
        _lib.lib.rnnt_debug_set_recurrent_cluster(1)
    np.testing.assert_array_equal(y.cpu().numpy(), y_o)
    np.testing.assert_array_equal(hx.cpu().numpy(), hx_o)
    np.testing.assert_array_equal(cx.cpu().numpy().view(np.uint16), cx_o)


def test_linear_bf16_bf16(ops):
    rng = np.random.default_rng(5)
    x = O.f32_to_bf16_bits(rng.normal(0, 1, (3, 37, 1024)).astype(np.float32))
    w = O.f32_to_bf16_bits(rng.normal(0, 0.05, (512, 1024)).astype(np.float32))
    got = bits(ops.linear_bf16_bf16(bf(x), bf(w)))
    np.testing.assert_array_equal(got, O.linear_bf16_bf16(x, w))


@pytest.mark.parametrize("t,n,o", [(19, 500, 512), (40, 384, 500)])
def test_linear_bf16_bf16_large_tiles(ops, t, n, o):
    """Shapes big enough for the 128 x 128-tile kernel (>= 296 tiles), ragged edges in M and O: same bits as the oracle."""
    rng = np.random.default_rng(t)
    x = O.f32_to_bf16_bits(rng.normal(0, 1, (t, n, 1024)).astype(np.float32))
    w = O.f32_to_bf16_bits(rng.normal(0, 0.05, (o, 1024)).astype(np.float32))
    got = bits(ops.linear_bf16_bf16(bf(x), bf(w)))
    np.testing.assert_array_equal(got, O.linear_bf16_bf16(x, w))


def test_linear_bf16_bf16_large_tiles_ragged(ops):
    """tile_steps: rows of batch tile b are computed for t < tile_steps[b] only, the rest is never written."""
    t, n, o = 24, 512, 512
    steps = np.array([24, 17, 5, 1], np.int32)
    rng = np.random.default_rng(9)
    x = O.f32_to_bf16_bits(rng.normal(0, 1, (t, n, 1024)).astype(np.float32))
    w = O.f32_to_bf16_bits(rng.normal(0, 0.05, (o, 1024)).astype(np.float32))
    ref = O.linear_bf16_bf16(x, w)
    got = bits(ops.linear_bf16_bf16(bf(x), bf(w), steps))
    for b, s in enumerate(steps):
        np.testing.assert_array_equal(got[:s, 128 * b:128 * b + 128], ref[:s, 128 * b:128 * b + 128])


@pytest.mark.parametrize("n,t2,blank_bias,seed", [(4, 6, 0.0, 1), (33, 9, 1.5, 2), (150, 12, 2.5, 3), (40, 5, -3.0, 4)])
def test_greedy_decode(ops, n, t2, blank_bias, seed):
    """Tokens, counters and prediction state bit-exact vs the oracle's batch-lockstep loop (incl. the reference's
    finished-lanes-keep-running behaviour, exercised by blank_bias <= 0)."""
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
    ops.greedy_decode(bf(y1), dev(f_lens), dw, d["res"], d["res_idx"], d["symbols_added"], d["time_idx"], d["pre_g"], pre_hg,
                      pre_cg, iters_out)
    from rnnt_inference_b200 import _lib
    _lib.check_device_status()
    assert int(iters_out.item()) == iters
    np.testing.assert_array_equal(d["res_idx"].cpu().numpy(), so["res_idx"])
    np.testing.assert_array_equal(d["res"].cpu().numpy(), so["res"])
    np.testing.assert_array_equal(d["pre_g"].cpu().numpy(), so["pre_g"])
    np.testing.assert_array_equal(d["symbols_added"].cpu().numpy(), so["symbols_added"])
    np.testing.assert_array_equal(d["time_idx"].cpu().numpy(), so["time_idx"])
    for i in range(2):
        np.testing.assert_array_equal(bits(pre_hg[i]), so["pre_hg"][i])
        np.testing.assert_array_equal(pre_cg[i].cpu().numpy().view(np.uint32), so["pre_cg"][i].view(np.uint32))
