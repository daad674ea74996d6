"""GPU parity tests of the C-ABI operators against the CPU oracle (bit-exact unless stated)."""
import numpy as np
import pytest
import torch

from oracle import oracle as O
from tests import helpers as Hh

pytestmark = pytest.mark.gpu

# see tests/test_gpu_pipeline.py: one bf16 ulp of the prediction state moves a logit of the synthetic joint by up to 0.25
DECODE_MARGIN_TOL = 0.3


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
    _lib.check(_lib.lib.rnnt_tune_recurrent_cluster(cluster))
    try:
        hx, cx = dev(hx0), dev(cx0)
        y = ops.lstm_layer_i8(dev(x), hx, cx, dev(lp["w_ih"]), dev(lp["w_hh"]), dev(lp["bias"]), lp["rb_scale"], lp["in_scale"],
                              lp["out_scale"], False, tile_steps=steps)
        _lib.check_device_status()
    finally:
        _lib.lib.rnnt_tune_recurrent_cluster(1)
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
                      pre_cg, iters_out, exact_order=True)
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


# ------------------------------------------------------------------------------------------------------- round 2
@pytest.mark.parametrize("n", [128, 520])
def test_lstm_layer_single_step(ops, n):
    """T == 1: every tile's only step is also its last one.  The final h goes through the ping-pong buffer and a copy
    kernel instead of being written to hx while other slice CTAs may still read h_0 from it.  Repeated so
    that a timing-dependent overwrite would show."""
    k = 256
    lp = Hh.make_lstm_layer(seed=3 + n, k_in=k)
    rng = np.random.default_rng(n)
    x = rng.integers(-100, 101, (1, n, k), dtype=np.int8)
    hx0 = rng.integers(-127, 128, (n, 1024), dtype=np.int8)
    cx0 = rng.normal(0, 0.5, (n, 1024)).astype(np.float16)
    hx_o, cx_o = hx0.copy(), cx0.view(np.uint16).copy()
    y_o = O.lstm_layer_i8(x, hx_o, cx_o, lp["w_ih"], lp["w_hh"], lp["bias"], lp["rb_scale"], lp["in_scale"], lp["out_scale"], False)
    w_ih, w_hh, bias, xd = dev(lp["w_ih"]), dev(lp["w_hh"]), dev(lp["bias"]), dev(x)
    from rnnt_inference_b200 import _lib
    for _ in range(20):
        hx, cx = dev(hx0), dev(cx0)
        y = ops.lstm_layer_i8(xd, hx, cx, w_ih, w_hh, bias, lp["rb_scale"], lp["in_scale"], lp["out_scale"], False)
        _lib.check_device_status()
        np.testing.assert_array_equal(y.cpu().numpy(), y_o)
        np.testing.assert_array_equal(hx.cpu().numpy(), hx_o)
        np.testing.assert_array_equal(cx.cpu().numpy().view(np.uint16), cx_o)


def _check_bf16_close(got, want, max_frac=1e-3, abs_tol=2e-5):
    """<= 1 bf16 ulp, except that a sum that cancels to (almost) zero may differ by abs_tol: there an fp32 accumulation-
    order difference of ~1e-6 is many ulps of the tiny result."""
    ulp = np.abs(got.astype(np.int32) - want.astype(np.int32))
    absd = np.abs(O.bf16_bits_to_f32(got) - O.bf16_bits_to_f32(want))
    bad = (ulp > 1) & (absd > abs_tol)
    assert not bad.any(), f"{bad.sum()} elements off by more than 1 bf16 ulp and {abs_tol}: max ulp {ulp[bad].max()}, max abs {absd[bad].max()}"
    assert (ulp != 0).mean() <= max_frac, f"{(ulp != 0).sum()} of {ulp.size} elements differ"


@pytest.mark.parametrize("m,n,k", [(1, 512, 1024), (130, 512, 1024), (1000, 512, 1024), (300, 192, 320), (257, 64, 72)])
def test_gemm_bf16_tcgen05_vs_oracle(ops, m, n, k):
    """tcgen05 kind::f16 GEMM (the hoisted f.W_trans): <= 1 bf16 ulp from the oracle's order-defined product, and only on
    a small fraction of the elements (the fp32 sums differ in the last bits; 1e-3 is the measured AMX-vs-oracle scale)."""
    rng = np.random.default_rng(m + n + k)
    x = O.f32_to_bf16_bits(rng.uniform(-1, 1, (m, k)).astype(np.float32))
    w = O.f32_to_bf16_bits(rng.normal(0, 0.05, (n, k)).astype(np.float32))
    got = bits(ops.gemm_bf16_bf16(bf(x), bf(w)))
    from rnnt_inference_b200 import _lib
    _lib.check_device_status()
    _check_bf16_close(got, O.linear_bf16_bf16(x, w))


def test_gemm_bf16_tcgen05_exact_on_small_integers(ops):
    """With small-integer data every product and partial sum is exactly representable, so ANY accumulation order must
    give the exact result (the reference's own KAT for its bf16 kernels, test_amx_linear.py:155-211)."""
    rng = np.random.default_rng(7)
    x = rng.integers(-4, 5, (700, 1024)).astype(np.float32)
    w = rng.integers(-4, 5, (512, 1024)).astype(np.float32)
    got = bits(ops.gemm_bf16_bf16(bf(O.f32_to_bf16_bits(x)), bf(O.f32_to_bf16_bits(w))))
    np.testing.assert_array_equal(got, O.f32_to_bf16_bits(x @ w.T))


def test_gemm_bf16_tcgen05_ragged(ops):
    t, n = 24, 512
    steps = np.array([24, 17, 5, 1], np.int32)
    rng = np.random.default_rng(9)
    x = O.f32_to_bf16_bits(rng.uniform(-1, 1, (t, n, 1024)).astype(np.float32))
    w = O.f32_to_bf16_bits(rng.normal(0, 0.05, (512, 1024)).astype(np.float32))
    ref = O.linear_bf16_bf16(x, w)
    got = bits(ops.gemm_bf16_bf16(bf(x), bf(w), steps))
    for b, s in enumerate(steps):
        _check_bf16_close(got[:s, 128 * b:128 * b + 128], ref[:s, 128 * b:128 * b + 128], max_frac=2e-3)


@pytest.mark.parametrize("batch_size", [32, 64, 128])
def test_kat_lstm_amx_int8_through_C(batch_size):
    """The reference's hot-path KAT (mlperf_plugins/test/pyfiles/test_amx_linear.py:214-295) replayed through the
    reference-named operator surface on the GPU: lstm_layer_amx_int8 == amx_linear_i8o32 x2 + lstm_postop per step,
    same deterministic arange / ones data (int8 wrap-around included), scales 0.01, 3 layers, T = 2, exact equality."""
    from rnnt_inference_b200 import _C as P
    isz = hsz = 1024
    osz = 4 * hsz
    T, layers = 2, 3

    def arange_i8(r, c):        # torch.tensor(np.arange(r * c).reshape(r, c), dtype=int8): wraps modulo 256
        return torch.from_numpy((np.arange(r * c) % 256).astype(np.uint8).view(np.int8).reshape(r, c).copy()).cuda()

    def arange_half(r, c):
        return torch.from_numpy((np.arange(r * c).reshape(r, c) / c / r).astype(np.float16)).cuda()

    ww_i = [arange_i8(osz, isz) for _ in range(layers)]          # K-major [4H, K] = get_weight(...) before the x86 tiling
    ww_h = [arange_i8(osz, hsz) for _ in range(layers)]
    b_i = [torch.from_numpy((np.arange(osz) / osz).astype(np.float32)).cuda() for _ in range(layers)]
    b_h = [b.clone() for b in b_i]
    x = torch.ones((T, batch_size, isz), dtype=torch.int8, device="cuda")
    xx = x.clone()
    h = [arange_i8(batch_size, hsz) for _ in range(layers)]
    c = [arange_half(batch_size, hsz) for _ in range(layers)]
    hh = [t.clone() for t in h]
    cc = [t.clone() for t in c]
    for i in range(layers):
        last = i == layers - 1
        y = [P.amx_linear_i8o32(x[j], ww_i[i], b_i[i], 0.01) for j in range(T)]
        outs = []
        for j in range(T):
            y[j] += P.amx_linear_i8o32(h[i], ww_h[i], b_h[i], 0.01)
            it, ft, gt, ot = y[j].chunk(4, 1)
            y_bf16, y_q, h_q, _ = P.lstm_postop(it, ft, gt, ot, c[i], 0.01, 0.01, last)
            h[i] = h_q
            outs.append(y_bf16 if last else y_q)
        x = torch.stack(outs)
        xx, hh[i], cc[i] = P.lstm_layer_amx_int8(xx, hh[i], cc[i], ww_i[i], ww_h[i], b_i[i], b_h[i] + b_i[i], 0.01, 0.01, 0.01, last)
        assert torch.equal(xx, x), f"layer{i}"
        assert torch.equal(hh[i], h[i]), f"layer{i}"
        assert torch.equal(cc[i], c[i]), f"layer{i}"
    from rnnt_inference_b200 import _lib
    _lib.check_device_status()


def test_lstm_postop_vs_oracle(ops):
    """lstm_postop as an operator == the oracle's (hardware-pinned) post-op: c in place, int8 h / y, bf16 y."""
    rng = np.random.default_rng(4)
    n, h = 37, 1024
    gates = (rng.normal(0, 3.0, (n, 4 * h))).astype(np.float32)
    c0 = rng.normal(0, 1.0, (n, h)).astype(np.float16)
    for last in (False, True):
        y_o, hq_o, c_o = O.lstm_postop_fp16(gates, c0.view(np.uint16), 101.5, 87.25, last)
        g = dev(gates)
        ct = dev(c0)
        y_bf16, y_q, h_q, ct2 = ops.lstm_postop(*g.chunk(4, 1), ct, 101.5, 87.25, last)
        assert ct2 is ct
        np.testing.assert_array_equal(ct.cpu().numpy().view(np.uint16), c_o)
        np.testing.assert_array_equal(h_q.cpu().numpy(), hq_o)
        np.testing.assert_array_equal(bits(y_bf16) if last else y_q.cpu().numpy(), y_o)


def test_ilinear_and_amx_linear_i8o32(ops):
    """a18: iLinear.forward = P.linear (int8 x int8 -> fp32 with the bias folded and one output scale) against the
    oracle's int32 GEMM + the two separately rounded fp32 ops of the i8o32b32 epilogue; also K = 240 -> 256 padding
    (test_amx_linear.py:103-152)."""
    from rnnt_inference_b200 import _C as P
    from rnnt_inference_b200.quant_linear import iLinear
    rng = np.random.default_rng(12)
    for m, k, o in ((32, 2048, 4096), (50, 256, 192)):
        x = rng.integers(-128, 128, (m, k), dtype=np.int8)
        w = rng.integers(-128, 128, (o, k), dtype=np.int8)
        b = rng.normal(0, 100, o).astype(np.float32)
        want = (O.gemm_i8_i32(x, w).astype(np.float32) + b) * np.float32(0.0123)
        got = P.amx_linear_i8o32(dev(x), dev(w), dev(b), 0.0123).cpu().numpy()
        np.testing.assert_array_equal(got.view(np.uint32), want.view(np.uint32))
        assert torch.equal(P.linear(dev(x), dev(w), dev(b), 0.0123, None).cpu(), torch.from_numpy(want))
    # the module: calibrated ranges -> quantised weight, folded bias, o_scale
    lin = iLinear(240, 128)
    lin._init_quantizers("quant")
    lin.input_quantizer.amax = torch.tensor(4.0)
    lin._quant_parameters("quant")
    lin = lin.cuda()
    xf = rng.normal(0, 1.5, (20, 240)).astype(np.float32)
    in_scale = np.float32(127.0) / np.float32(4.0)
    xq = np.clip(np.rint(xf * in_scale), -128, 127).astype(np.int8)
    xq_pad = np.zeros((20, 256), np.int8)
    xq_pad[:, :240] = xq
    wq = lin.weight.detach().cpu().numpy()
    assert wq.dtype == np.int8
    wq_pad = np.zeros((128, 256), np.int8)
    wq_pad[:, :wq.shape[1]] = wq
    y = lin(dev(xq_pad) if wq.shape[1] == 256 else dev(xq)).cpu().numpy()
    acc = O.gemm_i8_i32(xq_pad, wq_pad)
    want = (acc.astype(np.float32) + lin.bias.detach().cpu().numpy()) * np.float32(lin.o_scale.item())
    np.testing.assert_array_equal(y.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("ragged", [False, True])
def test_lstm_layer_i8_at_benched_shape(ops, ragged):
    """The benched shape: N = 2048 (16 batch tiles, 4 per CTA), T = 500, K = 1024.  Batch lanes are independent, so the
    oracle runs a sample of lanes (three per tile, over all 500 steps) and those lanes must be bit-identical."""
    t, n, k = 500, 2048, 1024
    lp = Hh.make_lstm_layer(seed=21, k_in=k)
    g = torch.Generator(device="cuda").manual_seed(5 + int(ragged))
    x = torch.randint(-100, 101, (t, n, k), dtype=torch.int8, device="cuda", generator=g)
    hx = torch.randint(-127, 128, (n, 1024), dtype=torch.int8, device="cuda", generator=g)
    cx = (torch.randn((n, 1024), device="cuda", generator=g) * 0.5).to(torch.float16)
    steps = np.linspace(500, 60, 16).astype(np.int32) if ragged else None
    lanes = np.array(sorted({128 * b + o for b in range(16) for o in (0, 77, 127)}))
    x_s = np.ascontiguousarray(x[:, lanes].cpu().numpy())
    hx_o = np.ascontiguousarray(hx[lanes].cpu().numpy())
    cx_o = np.ascontiguousarray(cx[lanes].cpu().numpy().view(np.uint16))
    y = ops.lstm_layer_i8(x, hx, cx, dev(lp["w_ih"]), dev(lp["w_hh"]), dev(lp["bias"]), lp["rb_scale"], lp["in_scale"],
                          lp["out_scale"], False, tile_steps=steps)
    from rnnt_inference_b200 import _lib
    _lib.check_device_status()
    y_s, hx_s, cx_s = y[:, lanes].cpu().numpy(), hx[lanes].cpu().numpy(), cx[lanes].cpu().numpy().view(np.uint16)
    if not ragged:
        y_o = O.lstm_layer_i8(x_s, hx_o, cx_o, lp["w_ih"], lp["w_hh"], lp["bias"], lp["rb_scale"], lp["in_scale"], lp["out_scale"], False)
        np.testing.assert_array_equal(y_s, y_o)
        np.testing.assert_array_equal(hx_s, hx_o)
        np.testing.assert_array_equal(cx_s, cx_o)
        return
    for b in range(16):
        sel = np.nonzero(lanes // 128 == b)[0]
        s = int(steps[b])
        h_t, c_t = np.ascontiguousarray(hx_o[sel]), np.ascontiguousarray(cx_o[sel])
        y_o = O.lstm_layer_i8(np.ascontiguousarray(x_s[:s, sel]), h_t, c_t, lp["w_ih"], lp["w_hh"], lp["bias"], lp["rb_scale"],
                              lp["in_scale"], lp["out_scale"], False)
        np.testing.assert_array_equal(y_s[:s, sel], y_o)
        np.testing.assert_array_equal(y_s[s:, sel], 0)
        np.testing.assert_array_equal(hx_s[sel], h_t)
        np.testing.assert_array_equal(cx_s[sel], c_t)


@pytest.mark.parametrize("exact_order", [True, False])
def test_greedy_decode_at_benched_shape(ops, exact_order):
    """2048 decode lanes (7 groups x 293), T2 = 250, the synthetic checkpoint's prediction / joint weights (emissions
    spread over the utterance).  Lanes are independent unless one hits the 30-symbol cap on its last frame (quirk b),
    so the oracle decodes a sample of lanes that provably do not (no burst on their last frame).
    exact_order=True (CUDA-core chains in the oracle's order): tokens, counters and prediction state bit for bit.
    exact_order=False (tensor cores, the product path): every transcript difference must start at a near-tie of the
    top-2 logits, >= 90 % of the sampled lanes identical, and on those the prediction state agrees to bf16 rounding."""
    from oracle import pipeline as OP
    from rnnt_inference_b200 import eval_accuracy as E
    from rnnt_inference_b200 import synthetic as S
    prm = OP.prepare_params(S.make_checkpoint())["dec"]
    t2, n = 250, 2048
    rng = np.random.default_rng(31)
    # smooth random encoder frames: AR(1) in time, like the synthetic encoder's output (correlation time ~5 frames)
    e = rng.normal(0, 0.3, (t2, n, 1024)).astype(np.float32)
    for t in range(1, t2):
        e[t] = 0.8 * e[t - 1] + 0.6 * e[t]
    f = O.f32_to_bf16_bits(np.clip(e, -1, 1))
    del e
    f_lens = np.sort(rng.integers(20, t2 + 1, n))[::-1].astype(np.int32).copy()
    f_lens[0] = t2
    y1 = bits(ops.gemm_bf16_bf16(bf(f), bf(prm["w_trans"])))           # the pipeline's own hoisted product
    L = 30 * t2
    sg = Hh.fresh_decode_state(n, L)
    d = {k: (dev(v) if not isinstance(v, list) else None) for k, v in sg.items()}
    pre_hg = [bf(v) for v in sg["pre_hg"]]
    pre_cg = [dev(v) for v in sg["pre_cg"]]
    dw = ops.make_decode_weights(bf(prm["embed"]), [bf(w) for w in prm["pred_w_ih"]], [bf(w) for w in prm["pred_w_hh"]],
                                 [dev(b) for b in prm["pred_bias"]], bf(prm["w_pred"]), dev(prm["b1"]), bf(prm["w2"]),
                                 dev(prm["b2"]))
    iters_out = torch.zeros(1, dtype=torch.int32, device="cuda")
    ops.greedy_decode(bf(y1), dev(f_lens), dw, d["res"], d["res_idx"], d["symbols_added"], d["time_idx"], d["pre_g"], pre_hg,
                      pre_cg, iters_out, exact_order=exact_order)
    from rnnt_inference_b200 import _lib
    _lib.check_device_status()
    res_idx = d["res_idx"].cpu().numpy()
    assert res_idx.max() >= 10 and (res_idx + 1).sum() > n, "degenerate: hardly any token emitted"
    # lanes whose last frame cannot burst: the burst unit (joint hidden 28) is off there
    last_y1 = O.bf16_bits_to_f32(y1[f_lens - 1, np.arange(n), 28])
    safe = np.nonzero(last_y1 + prm["b1"][28] < -0.5)[0]
    lanes = safe[np.linspace(0, len(safe) - 1, 48).astype(int)]
    so = Hh.fresh_decode_state(len(lanes), L)
    trace = O.new_decision_trace(len(lanes), 2 * t2 + 256)
    O.greedy_decode(np.ascontiguousarray(f[:, lanes]), f_lens[lanes], prm, so["res"], so["res_idx"], so["symbols_added"],
                    so["time_idx"], so["pre_g"], so["pre_hg"], so["pre_cg"], y1_pre=np.ascontiguousarray(y1[:, lanes]), trace=trace)
    res = d["res"].cpu().numpy()[lanes]
    if exact_order:
        np.testing.assert_array_equal(res_idx[lanes], so["res_idx"])
        for i in range(len(lanes)):
            np.testing.assert_array_equal(res[i, :so["res_idx"][i] + 1], so["res"][i, :so["res_idx"][i] + 1])
        np.testing.assert_array_equal(d["pre_g"].cpu().numpy()[lanes], so["pre_g"])
        for i in range(2):
            np.testing.assert_array_equal(bits(pre_hg[i])[lanes], so["pre_hg"][i])
        return
    got = [list(res[i, :res_idx[lanes[i]] + 1]) for i in range(len(lanes))]
    truth = [list(so["res"][i, :so["res_idx"][i] + 1]) for i in range(len(lanes))]
    rep = E.compare_transcripts(got, truth, trace, margin_tol=DECODE_MARGIN_TOL)
    assert rep["unexplained"] == [], rep["differences"]
    assert rep["identical_utterances"] >= 0.9 * len(lanes), rep["differences"]
    same = [i for i in range(len(lanes)) if got[i] == truth[i]]
    for layer in range(2):
        h_g = O.bf16_bits_to_f32(bits(pre_hg[layer])[lanes][same])
        h_o = O.bf16_bits_to_f32(so["pre_hg"][layer][same])
        assert np.abs(h_g - h_o).max() <= 2.0 ** -6, np.abs(h_g - h_o).max()      # a few bf16 ulps of |h| <= 1
        c_g = pre_cg[layer].cpu().numpy()[lanes][same]
        assert np.abs(c_g - so["pre_cg"][layer][same]).max() <= 2e-2


@pytest.mark.parametrize("n,t2", [(40, 30), (300, 40)])
def test_greedy_decode_tensor_cores_small(ops, n, t2):
    """Tensor-core decode on small batches (one lane group partially filled / several groups): against the oracle on ALL
    lanes, same y1: differences only at near-ties, counters consistent."""
    from oracle import pipeline as OP
    from rnnt_inference_b200 import eval_accuracy as E
    from rnnt_inference_b200 import synthetic as S
    prm = OP.prepare_params(S.make_checkpoint(seed=11))["dec"]
    rng = np.random.default_rng(n)
    e = rng.normal(0, 0.3, (t2, n, 1024)).astype(np.float32)
    for t in range(1, t2):
        e[t] = 0.8 * e[t - 1] + 0.6 * e[t]
    f = O.f32_to_bf16_bits(np.clip(e, -1, 1))
    f_lens = rng.integers(0, t2 + 1, n).astype(np.int32)
    f_lens[0] = t2
    y1 = O.linear_bf16_bf16(f, prm["w_trans"])
    L = 30 * t2
    so = Hh.fresh_decode_state(n, L)
    trace = O.new_decision_trace(n, 32 * t2)
    iters = O.greedy_decode(f, f_lens, prm, so["res"], so["res_idx"], so["symbols_added"], so["time_idx"], so["pre_g"],
                            so["pre_hg"], so["pre_cg"], y1_pre=y1, trace=trace)
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
    res, res_idx = d["res"].cpu().numpy(), d["res_idx"].cpu().numpy()
    got = [list(res[i, :res_idx[i] + 1]) for i in range(n)]
    truth = [list(so["res"][i, :so["res_idx"][i] + 1]) for i in range(n)]
    assert sum(len(t) for t in truth) > n // 2
    rep = E.compare_transcripts(got, truth, trace, margin_tol=DECODE_MARGIN_TOL)
    assert rep["unexplained"] == [], rep["differences"]
    assert rep["identical_utterances"] >= 0.9 * n, rep["differences"]
    if rep["identical_utterances"] == n:
        assert int(iters_out.item()) == iters
    assert (res_idx[f_lens == 0] == -1).all()


@pytest.mark.parametrize("cm", [1, 2, 4, "pair"])
def test_gemm_i8_clusters(ops, cm):
    """Projection GEMM with CM CTAs per cluster sharing the W operand by TMA multicast, or on CTA pairs (cta_group::2
    MMAs, 256 x 256 tiles): same int32 results as the CUDA-core GEMM for a large problem (several tiles per CTA, an odd
    number of m tiles with an edge tile) and, for a ragged batch whose super tiles straddle timesteps and mix live and
    dead m tiles, live rows exact and dead rows untouched."""
    import ctypes
    from rnnt_inference_b200 import _lib
    _lib.check(_lib.lib.rnnt_tune_gemm_pair(1 if cm == "pair" else 0))
    if cm != "pair":
        _lib.check(_lib.lib.rnnt_tune_gemm_cluster(cm))
    try:
        g = torch.Generator(device="cuda").manual_seed(7 if cm == "pair" else cm)
        m, n, k = 20000 + 77, 4096, 1024
        a = torch.randint(-128, 128, (m, k), dtype=torch.int8, device="cuda", generator=g)
        w = torch.randint(-128, 128, (n, k), dtype=torch.int8, device="cuda", generator=g)
        got = ops.gemm_i8_i32(a, w)
        ref = ops.gemm_i8_i32(a, w, simt=True)
        _lib.check_device_status()
        assert torch.equal(got, ref)
        # ragged: 6 timesteps x 1280 rows (10 batch tiles), tile b alive for t < steps[b]
        t, rows, k = 6, 1280, 256
        steps = np.array([6, 6, 5, 5, 4, 3, 3, 2, 1, 0], np.int32)
        a = torch.randint(-128, 128, (t * rows, k), dtype=torch.int8, device="cuda", generator=g)
        c = torch.full((t * rows, n), 123456789, dtype=torch.int32, device="cuda")
        _lib.check(_lib.lib.rnnt_gemm_i8_i32_ragged(_lib.ptr(a), _lib.ptr(w[:, :k].contiguous()), _lib.ptr(c), t * rows, n, k, rows,
                                                    ctypes.c_void_p(steps.ctypes.data), _lib.stream_ptr()))
        ref = ops.gemm_i8_i32(a, w[:, :k].contiguous(), simt=True)
        _lib.check_device_status()
        c, ref = c.view(t, 10, 128, n), ref.view(t, 10, 128, n)
        for b, s in enumerate(steps):
            assert torch.equal(c[:s, b], ref[:s, b]), f"tile {b}"
            assert bool((c[s:, b] == 123456789).all()), f"dead rows of tile {b} were written"
    finally:
        _lib.lib.rnnt_tune_gemm_cluster(1)
        _lib.lib.rnnt_tune_gemm_pair(1)



@pytest.mark.parametrize("rows", [1000, 96, 1696])
def test_gemm_i8_ragged_any_batch_width(ops, rows):
    """Ragged row skipping when the batch width is not a multiple of the 128-row GEMM tile (m tiles straddle batch tiles
    and timesteps): every live row exact, and whole blocks of dead rows really are skipped."""
    import ctypes
    from rnnt_inference_b200 import _lib
    g = torch.Generator(device="cuda").manual_seed(rows)
    t, n, k = 9, 4096, 256
    nb = (rows + 127) // 128
    steps = np.maximum(0, t - np.arange(nb) * 2).astype(np.int32)          # 9, 7, 5, ... sorted like a ragged batch
    a = torch.randint(-128, 128, (t * rows, k), dtype=torch.int8, device="cuda", generator=g)
    w = torch.randint(-128, 128, (n, k), dtype=torch.int8, device="cuda", generator=g)
    c = torch.full((t * rows, n), 123456789, dtype=torch.int32, device="cuda")
    _lib.check(_lib.lib.rnnt_gemm_i8_i32_ragged(_lib.ptr(a), _lib.ptr(w), _lib.ptr(c), t * rows, n, k, rows,
                                                ctypes.c_void_p(steps.ctypes.data), _lib.stream_ptr()))
    ref = ops.gemm_i8_i32(a, w, simt=True)
    _lib.check_device_status()
    live = torch.from_numpy((np.arange(t)[:, None] < steps[np.arange(rows) // 128][None, :]).reshape(-1)).cuda()
    assert torch.equal(c[live], ref[live])
    untouched = (c == 123456789).all(dim=1)
    assert not bool((untouched & live).any())
    if nb > 1:
        assert int(untouched.sum()) >= 128                               # some all-dead block was skipped
    # the hoisted joint GEMM (bf16) shares the skip logic
    x = torch.randn((t * rows, 64), device="cuda", generator=g).to(torch.bfloat16)
    wb = torch.randn((512, 64), device="cuda", generator=g).to(torch.bfloat16)
    full = ops.gemm_bf16_bf16(x, wb)
    part = ops.gemm_bf16_bf16(x.view(t, rows, 64), wb, tile_steps=steps).view(t * rows, 512)
    _lib.check_device_status()
    assert torch.equal(part[live], full[live])
