"""CPU oracle of the whole int8/bf16 path: checkpoint -> quantised parameters -> encoder -> greedy decode.
TEST INFRASTRUCTURE (only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it).

numpy + the C++ operator restatements in rnnt_oracle.cpp.  Independent of the product package: the scale algebra
below restates the reference's Python (file:line cited per function), it does not import rnnt_inference_b200.
"""
import numpy as np

from . import oracle as O

H, P, J, V = 1024, 320, 512, 29
PRE_LAYERS, POST_LAYERS, PRED_LAYERS = 2, 3, 2


def _f32(t):
    return np.ascontiguousarray(t.detach().cpu().numpy() if hasattr(t, "detach") else t, dtype=np.float32)


def _quant_weight(w, scale):
    # WeightQuantizer._quant_forward (models/quant_modules.py:214-222): clamp(round_half_even(w * scale), -128, 127)
    return np.clip(np.rint(w * scale), -128, 127).astype(np.int8)


def prepare_params(sd, amax_in=None):
    """MLPerf-format state dict -> oracle parameters.

    Encoder (models/quant_lstm.py:193-215, modeling_rnnt.py:62-77): per layer one weight amax over cat[w_ih, w_hh],
    w_scale = 127/amax, in_scale = 127/amax_in[layer]; bias = (b_ih + b_hh) * (in_scale * w_scale);
    rb_scale = 1 / (in_scale * w_scale); out_scale[l] = in_scale[l+1], pre L1 out = post L0 in.
    Prediction / joint (modeling_rnnt.py:161-181,223-257; utils.py:60-81): bf16 weights, summed biases, fc1 split at
    column 1024 with the transcription bias zero, fc2 padded 29 -> 32 with zeros."""
    f32 = np.float32
    enc = []
    names = [("encoder.pre_rnn.lstm", l) for l in range(PRE_LAYERS)] + [("encoder.post_rnn.lstm", l) for l in range(POST_LAYERS)]
    amax_keys = [f"transcription.pre_rnn.lstm{l}" for l in range(PRE_LAYERS)] + \
                [f"transcription.post_rnn.lstm{l}" for l in range(POST_LAYERS)]
    in_scales = []
    for i, key in enumerate(amax_keys):
        a = amax_in[i] if amax_in is not None else float(sd[f"{key}.input_quantizer._amax"])
        in_scales.append(f32(127.0) / f32(a))          # TensorQuantizer.scale (quant_modules.py:102-104), fp32 division
    for i, (prefix, l) in enumerate(names):
        w_ih, w_hh = _f32(sd[f"{prefix}.weight_ih_l{l}"]), _f32(sd[f"{prefix}.weight_hh_l{l}"])
        b_ih, b_hh = _f32(sd[f"{prefix}.bias_ih_l{l}"]), _f32(sd[f"{prefix}.bias_hh_l{l}"])
        w_amax = f32(max(np.abs(w_ih).max(), np.abs(w_hh).max()))
        w_scale = f32(127.0) / w_amax
        k = w_ih.shape[1]
        k_pad = (k + 63) // 64 * 64                       # first layer only needs it (240 -> 256)
        w_ih_q = np.zeros((4 * H, k_pad), np.int8)
        w_ih_q[:, :k] = _quant_weight(w_ih, w_scale)
        b_scale = f32(in_scales[i] * w_scale)
        is_last = i in (PRE_LAYERS - 1, PRE_LAYERS + POST_LAYERS - 1)
        out_scale = in_scales[i + 1] if i + 1 < len(names) else f32(np.inf)
        enc.append(dict(w_ih=w_ih_q, w_hh=_quant_weight(w_hh, w_scale), bias=((b_hh + b_ih) * b_scale).astype(f32),
                        rb_scale=float(f32(1 / float(b_scale))),   # 1 / b_scale.item() in double, stored as fp32 (quant_lstm.py:209,213)
                        in_scale=float(in_scales[i]), out_scale=float(out_scale), k_pad=k_pad, last_of_stack=is_last))
    bf = O.f32_to_bf16_bits
    w1 = _f32(sd["joint_net.0.weight"])
    w2 = np.zeros((32, J), f32)
    w2[:V] = _f32(sd["joint_net.3.weight"])
    b2 = np.zeros(32, f32)
    b2[:V] = _f32(sd["joint_net.3.bias"])
    dec = dict(
        embed=bf(_f32(sd["prediction.embed.weight"])),
        pred_w_ih=[bf(_f32(sd[f"prediction.dec_rnn.lstm.weight_ih_l{l}"])) for l in range(PRED_LAYERS)],
        pred_w_hh=[bf(_f32(sd[f"prediction.dec_rnn.lstm.weight_hh_l{l}"])) for l in range(PRED_LAYERS)],
        pred_bias=[(_f32(sd[f"prediction.dec_rnn.lstm.bias_hh_l{l}"]) + _f32(sd[f"prediction.dec_rnn.lstm.bias_ih_l{l}"]))
                   for l in range(PRED_LAYERS)],
        w_trans=bf(w1[:, :H]), w_pred=bf(w1[:, H:]),
        b1=(np.zeros(J, f32) + _f32(sd["joint_net.0.bias"])).astype(f32),     # b_trans (=0) + b_pred
        w2=bf(w2), b2=b2)
    return dict(enc=enc, dec=dec)


def encode(params, x, x_lens, state=None):
    """Transcription.forward in quant mode (modeling_rnnt.py:116-144).  x {T,N,240} fp32 -> f {ceil(T/2),N,1024} bf16 bits.
    state: optional dict(hx=[5 x int8 {N,H}], cx=[5 x uint16 {N,H}]) updated in place (split_len chunks)."""
    enc = params["enc"]
    t, n, _ = x.shape
    if state is None:
        state = new_encoder_state(n)
    cur = O.quantize_pad(x, enc[0]["in_scale"], enc[0]["k_pad"])
    for i, lp in enumerate(enc):
        if i == PRE_LAYERS:
            cur = O.stack_time(cur, np.asarray(x_lens, np.int32), 2)
        last = i == len(enc) - 1
        cur = O.lstm_layer_i8(cur, state["hx"][i], state["cx"][i], lp["w_ih"], lp["w_hh"], lp["bias"], lp["rb_scale"],
                              lp["in_scale"], lp["out_scale"], last)
    return cur


def new_encoder_state(n):
    return dict(hx=[np.zeros((n, H), np.int8) for _ in range(PRE_LAYERS + POST_LAYERS)],
                cx=[np.zeros((n, H), np.uint16) for _ in range(PRE_LAYERS + POST_LAYERS)])


def new_decode_state(n, max_len):
    return dict(res=np.full((n, 30 * max_len), -1, np.int32), res_idx=np.full(n, -1, np.int32),
                symbols_added=np.zeros(n, np.int32), time_idx=np.zeros(n, np.int32), pre_g=np.full(n, -1, np.int32),
                pre_hg=[np.zeros((n, P), np.uint16) for _ in range(PRED_LAYERS)],
                pre_cg=[np.zeros((n, P), np.float32) for _ in range(PRED_LAYERS)])


def greedy_decode_batch(params, x, x_lens, split_len=-1, trace=None):
    """GreedyDecoder.forward in quant mode (models/decoder.py:21-94,171-212).  x {T,N,240} fp32 padded, x_lens {N}.
    Returns (res {N, 30*max(x_lens)} int32, res_len {N} int32, iterations)."""
    x = np.ascontiguousarray(x, np.float32)
    x_lens = np.asarray(x_lens, np.int64)
    n = x.shape[1]
    max_len = int(x_lens.max())
    ds = new_decode_state(n, max_len)
    es = new_encoder_state(n)
    iters = 0
    chunks = [(0, x.shape[0])] if split_len == -1 else [(s, min(s + split_len, max_len)) for s in range(0, max_len, split_len)]
    for s, e in chunks:
        xi = x[s:e]
        xi_lens = x_lens if split_len == -1 else np.minimum(split_len, np.clip(x_lens - s, 0, None))
        f = encode(params, xi, xi_lens, es)
        f_lens = ((xi_lens + 1) // 2).astype(np.int32)       # ceil(f_lens / 2) (decoder.py:185)
        ds["symbols_added"][:] = 0                            # decoder.py:173-174: counters restart per chunk
        ds["time_idx"][:] = 0
        iters += O.greedy_decode(f, f_lens, params["dec"], ds["res"], ds["res_idx"], ds["symbols_added"], ds["time_idx"],
                                 ds["pre_g"], ds["pre_hg"], ds["pre_cg"], trace=trace)
    return ds["res"], ds["res_idx"] + 1, iters


def pad_batch(feats, pad_to=32):
    """PytorchSUT.inference batching (models/pytorch_sut.py:85-104): time-major zero padding, N up to a multiple of 32."""
    lens = [int(f.shape[0]) for f in feats]
    n = len(feats)
    n_pad = (n + pad_to - 1) // pad_to * pad_to if pad_to else n
    x = np.zeros((max(lens), n_pad, feats[0].shape[1]), np.float32)
    for i, f in enumerate(feats):
        x[:lens[i], i] = _f32(f)
    return x, np.array(lens + [0] * (n_pad - n), np.int64)
