"""Shared test data builders (seeded, numpy).  Test infrastructure."""
import numpy as np

from oracle import oracle as O

H, P, J = 1024, 320, 512


def make_lstm_layer(seed, k_in, k_pad=None, in_amax=1.0, out_amax=1.0, hidden=H):
    """Random fp32 LSTM layer quantised with the reference's scale algebra (quant_lstm.py:193-215):
    one weight amax over cat[w_ih, w_hh], bias folded as (b_ih+b_hh)*in_scale*w_scale, rb = 1/(in*w)."""
    rng = np.random.default_rng(seed)
    k_pad = k_in if k_pad is None else k_pad
    bound = 1.0 / np.sqrt(hidden)
    w_ih = rng.uniform(-bound, bound, (4 * hidden, k_in)).astype(np.float32)
    w_hh = rng.uniform(-bound, bound, (4 * hidden, hidden)).astype(np.float32)
    b = (rng.uniform(-bound, bound, 4 * hidden) + rng.uniform(-bound, bound, 4 * hidden)).astype(np.float32)
    b[hidden:2 * hidden] += 1.0  # forget gate bias
    w_amax = np.float32(max(np.abs(w_ih).max(), np.abs(w_hh).max()))
    w_scale = np.float32(127.0) / w_amax
    in_scale = np.float32(127.0) / np.float32(in_amax)
    out_scale = np.float32(127.0) / np.float32(out_amax)

    def q(w):
        return np.clip(np.rint(w * w_scale), -128, 127).astype(np.int8)

    w_ih_q = np.zeros((4 * hidden, k_pad), np.int8)
    w_ih_q[:, :k_in] = q(w_ih)
    b_scale = np.float32(in_scale * w_scale)
    return dict(w_ih=w_ih_q, w_hh=q(w_hh), bias=(b * b_scale).astype(np.float32), rb_scale=float(np.float32(1.0) / b_scale),
                in_scale=float(in_scale), out_scale=float(out_scale))


def make_decode_params(seed, blank_bias=0.0, scale=1.0):
    """Random prediction/joint weights as bf16 bit patterns in the oracle's layout."""
    rng = np.random.default_rng(seed)
    bf = O.f32_to_bf16_bits

    def u(shape, bound):
        return rng.uniform(-bound, bound, shape).astype(np.float32)

    bp = scale / np.sqrt(P)
    prm = dict(
        embed=bf(rng.normal(0, 0.3, (28, P)).astype(np.float32)),
        pred_w_ih=[bf(u((4 * P, P), bp)) for _ in range(2)],
        pred_w_hh=[bf(u((4 * P, P), bp)) for _ in range(2)],
        pred_bias=[(u(4 * P, bp) + u(4 * P, bp)).astype(np.float32) for _ in range(2)],
        w_trans=bf(u((J, H), scale / np.sqrt(H + P))),
        w_pred=bf(u((J, P), scale / np.sqrt(H + P))),
        b1=u(J, 1.0 / np.sqrt(H + P)),
    )
    w2 = np.zeros((32, J), np.float32)
    w2[:29] = u((29, J), scale / np.sqrt(J))
    b2 = np.zeros(32, np.float32)
    b2[:29] = u(29, 1.0 / np.sqrt(J))
    b2[28] += blank_bias
    prm["w2"], prm["b2"] = bf(w2), b2
    return prm


def fresh_decode_state(n, L):
    return dict(res=np.full((n, L), -1, np.int32), res_idx=np.full(n, -1, np.int32), symbols_added=np.zeros(n, np.int32),
                time_idx=np.zeros(n, np.int32), pre_g=np.full(n, -1, np.int32),
                pre_hg=[np.zeros((n, P), np.uint16) for _ in range(2)], pre_cg=[np.zeros((n, P), np.float32) for _ in range(2)])
