"""GPU parity of the whole path (checkpoint -> RNNT modules -> GreedyDecoder / PytorchSUT) against the CPU oracle pipeline.

Two tiers, as SURVEY 7.2 rules:
  * int8 / fp16 tier (encoder) and the decode state machine: bit-exact.  With `exact_order=True` the hoisted f.W_trans runs
    on CUDA cores in the oracle's accumulation order and whole transcripts are bit-exact.
  * default path: the bf16 contractions (hoisted f.W_trans: tcgen05; decode loop: mma.sync) run on the tensor cores, whose
    fp32 accumulation order is neither the oracle's nor AMX's.  Checked as: encoder output bit-exact; y1 <= 1 bf16 ulp;
    every transcript that differs from the oracle's must have departed at a near-tie of the top-2 logits (margin report,
    rnnt_inference_b200.eval_accuracy), and >= 90 % of the transcripts are identical."""
import numpy as np
import pytest
import torch

from oracle import pipeline as OP
from rnnt_inference_b200 import synthetic as S

pytestmark = pytest.mark.gpu

LENS = [37, 30, 30, 22, 9, 4, 1]


@pytest.fixture(scope="module")
def setup():
    from rnnt_inference_b200.modeling_rnnt import RNNT
    sd = S.make_checkpoint(seed=11)
    rnnt = RNNT(sd, "quant", True).eval()
    params = OP.prepare_params(sd)
    feats = [S.make_features(n, 100 + i) for i, n in enumerate(LENS)]
    return sd, rnnt, params, feats


def _oracle(params, feats, split_len=-1, pad_to=0):
    x, xl = OP.pad_batch(feats, pad_to=pad_to)
    return OP.greedy_decode_batch(params, x, xl, split_len)


def _check(res, res_len, o_res, o_len, n):
    res, res_len = res.cpu().numpy(), res_len.cpu().numpy()
    np.testing.assert_array_equal(res_len[:n], o_len[:n])
    assert o_len[:n].sum() > 0, "degenerate case: the oracle emitted nothing"
    for i in range(n):
        np.testing.assert_array_equal(res[i, :o_len[i]], o_res[i, :o_len[i]])


def test_quantised_parameters_match_oracle(setup):
    """Scale algebra / weight quantisation of the modules (quant_lstm.py:193-215) == the oracle's restatement."""
    _sd, rnnt, params, _ = setup
    stacks = [rnnt.transcription.pre_rnn, rnnt.transcription.post_rnn]
    i = 0
    for rnn in stacks:
        for layer in range(rnn.num_layers):
            w_ih, w_hh, _b, bias = rnn.weights[layer]
            p = params["enc"][i]
            np.testing.assert_array_equal(w_ih.cpu().numpy(), p["w_ih"])
            np.testing.assert_array_equal(w_hh.cpu().numpy(), p["w_hh"])
            np.testing.assert_array_equal(bias.cpu().numpy().view(np.uint32), p["bias"].view(np.uint32))
            assert float(rnn.rb_scale[layer]) == p["rb_scale"] and float(rnn.in_scale[layer]) == p["in_scale"]
            if i != 4:
                assert float(rnn.out_scale[layer]) == p["out_scale"]
            i += 1


@pytest.mark.parametrize("fused", [True, False])
def test_decoder_forward_vs_oracle(setup, fused):
    from rnnt_inference_b200.decoder import GreedyDecoder
    _sd, rnnt, params, feats = setup
    dec = GreedyDecoder(rnnt, "quant", True, -1, len(feats), fused=fused, exact_order=True)
    x, xl = OP.pad_batch(feats, pad_to=0)
    with torch.no_grad():
        res, res_len = dec(torch.from_numpy(x), torch.from_numpy(xl))
    o_res, o_len, o_iters = _oracle(params, feats)
    _check(res, res_len, o_res, o_len, len(feats))
    if fused:
        assert int(dec.last_iters.item()) == o_iters


# A transcript difference must start at a decision whose top-2 logit margin is below this.  The bound is the sensitivity
# of a logit to ONE bf16 ulp of the prediction state, which any change of the fp32 accumulation order can flip (AMX's
# TDPBF16PS differs from the oracle's order just the same): h ~ 1 has ulp 2^-8; the synthetic joint reads the label trace
# with weight 8 and pushes the label's logit with weight -8 (synthetic.py), so one ulp moves a logit by up to
# 8 * 8 * 2^-8 = 0.25.  (The encoder-side product alone moves logits by < 1e-3.)
MARGIN_TOL = DECODE_MARGIN_TOL = 0.3


def test_decoder_tensor_core_hoist_vs_oracle(setup):
    """The default path (tcgen05 hoist) on a batch of 40 utterances."""
    from oracle import oracle as O
    from rnnt_inference_b200 import eval_accuracy as E
    from rnnt_inference_b200.decoder import GreedyDecoder
    _sd, rnnt, params, _ = setup
    lens = [int(v) for v in np.linspace(120, 10, 40)]
    feats = [S.make_features(n, 700 + i) for i, n in enumerate(lens)]
    dec = GreedyDecoder(rnnt, "quant", True, -1, len(feats), keep_intermediates=True)
    x, xl = OP.pad_batch(feats, pad_to=0)
    with torch.no_grad():
        res, res_len = dec(torch.from_numpy(x), torch.from_numpy(xl))
    from rnnt_inference_b200 import _lib
    _lib.check_device_status()
    res, res_len = res.cpu().numpy(), res_len.cpu().numpy()
    bits = lambda t: t.contiguous().view(torch.int16).cpu().numpy().view(np.uint16)
    f_gpu, y1_gpu = bits(dec.last_f), bits(dec.last_y1)
    f_o = OP.encode(params, x, xl)
    np.testing.assert_array_equal(f_gpu, f_o)                                   # int8 / fp16 tier: bit-exact
    y1_o = O.linear_bf16_bf16(f_o, params["dec"]["w_trans"])
    ulp = np.abs(y1_gpu.astype(np.int32) - y1_o.astype(np.int32))
    absd = np.abs(O.bf16_bits_to_f32(y1_gpu) - O.bf16_bits_to_f32(y1_o))
    assert not ((ulp > 1) & (absd > 2e-5)).any() and (ulp != 0).mean() <= 1e-3
    f_lens = ((xl + 1) // 2).astype(np.int32)

    def oracle_decode(y1, trace=None):
        ds = OP.new_decode_state(len(feats), int(xl.max()))
        O.greedy_decode(f_o, f_lens, params["dec"], ds["res"], ds["res_idx"], ds["symbols_added"], ds["time_idx"], ds["pre_g"],
                        ds["pre_hg"], ds["pre_cg"], y1_pre=y1, trace=trace)
        return [list(ds["res"][i, :ds["res_idx"][i] + 1]) for i in range(len(feats))]

    got = [list(res[i, :res_len[i]]) for i in range(len(feats))]
    assert sum(len(t) for t in got) > 100
    # the decode loop's contractions run on the tensor cores as well: against the oracle decoding the SAME y1, and against
    # the pure oracle, a transcript may only depart at a near-tie of the top-2 logits
    for y1 in (y1_gpu, y1_o):
        trace = O.new_decision_trace(len(feats), 2 * int(f_lens.max()) + 64)
        truth = oracle_decode(y1, trace)
        rep = E.compare_transcripts(got, truth, trace, margin_tol=MARGIN_TOL)
        assert rep["unexplained"] == [], rep["differences"]
        assert rep["identical_utterances"] >= 0.9 * len(feats)


def test_decoder_split_len_vs_oracle(setup):
    """Chunked encoder with carried state (decoder.py:80-91, split_len = 8)."""
    from rnnt_inference_b200.decoder import GreedyDecoder
    _sd, rnnt, params, feats = setup
    dec = GreedyDecoder(rnnt, "quant", True, 8, len(feats), exact_order=True)
    x, xl = OP.pad_batch(feats, pad_to=0)
    with torch.no_grad():
        res, res_len = dec(torch.from_numpy(x), torch.from_numpy(xl))
    o_res, o_len, _ = _oracle(params, feats, split_len=8)
    _check(res, res_len, o_res, o_len, len(feats))


def test_forward_ragged_equals_forward(setup):
    from rnnt_inference_b200.decoder import GreedyDecoder
    _sd, rnnt, _params, feats = setup
    dec = GreedyDecoder(rnnt, "quant", True, -1, len(feats))
    x, xl = OP.pad_batch(feats, pad_to=32)
    with torch.no_grad():
        res, res_len = dec(torch.from_numpy(x), torch.from_numpy(xl))
        res, res_len = res.clone(), res_len.clone()
        packed = torch.cat(feats, 0).cuda()
        offsets = torch.zeros(32, dtype=torch.int64)
        offsets[:len(feats)] = torch.tensor(np.concatenate([[0], np.cumsum(LENS[:-1])]))
        lens = torch.zeros(32, dtype=torch.int32)
        lens[:len(feats)] = torch.tensor(LENS, dtype=torch.int32)
        res2, len2 = dec.forward_ragged(packed, offsets.cuda(), lens.cuda(), max(LENS))
        res2, len2 = res2.clone(), len2.clone()
        res3, len3 = dec.forward_ragged(packed, offsets.cuda(), lens.cuda(), max(LENS), list(LENS))   # tile-skipping encoder
    assert torch.equal(res_len, len2) and torch.equal(res, res2)
    assert torch.equal(res_len, len3) and torch.equal(res, res3)


def test_sut_offline_query_vs_oracle(setup):
    """PytorchSUT issue/complete contract through the LoadGen shim (Offline, accuracy mode): every sample's int32 token
    bytes equal the oracle's transcript of that utterance decoded in the same sorted batches."""
    from rnnt_inference_b200 import loadgen_shim as lg
    from rnnt_inference_b200.pytorch_sut import PytorchSUT
    sd, _rnnt, params, _ = setup
    lens = [25, 7, 31, 18, 12, 31, 3, 22, 9, 16]
    feats = [S.make_features(n, 300 + i) for i, n in enumerate(lens)]
    data = {"x": feats, "x_lens": [torch.tensor(n) for n in lens]}
    sut = PytorchSUT(sd, data, 4, "quant", None, enable_bf16=True, scenario="Offline", exact_order=True)
    settings = lg.TestSettings()
    settings.scenario = lg.TestScenario.Offline
    settings.mode = lg.TestMode.AccuracyOnly
    out = lg.StartTest(sut.sut, sut.qsl.qsl, settings)
    assert out["completed"] == len(lens)
    got = {e["qsl_idx"]: np.frombuffer(bytes.fromhex(e["data"]), np.int32) for e in out["accuracy_log"]}
    order = sorted(range(len(lens)), key=lambda i: -lens[i])          # stable, like list.sort(reverse=True)
    for c in range(0, len(order), 4):
        idx = order[c:c + 4]
        o_res, o_len, _ = _oracle(params, [feats[i] for i in idx], pad_to=32)
        for j, i in enumerate(idx):
            np.testing.assert_array_equal(got[i], o_res[j, :o_len[j]])


@pytest.mark.parametrize("compute_streams", [1, 2])
def test_sut_offline_head_split_vs_oracle(setup, compute_streams):
    """The pipelined Offline path with the short end of the query cut by frames (sharding.plan_chunks, forced on for this
    tiny query): every sample answered once, bytes equal to the oracle's in the planner's batches.  compute_streams=2:
    consecutive chunks on alternating streams (their cooperative kernels are ordered by the library)."""
    from rnnt_inference_b200 import loadgen_shim as lg
    from rnnt_inference_b200 import sharding
    from rnnt_inference_b200.pytorch_sut import PytorchSUT
    sd, _rnnt, params, _ = setup
    lens = [25, 7, 31, 18, 12, 31, 3, 22, 9, 16, 28, 5, 14]
    feats = [S.make_features(n, 700 + i) for i, n in enumerate(lens)]
    data = {"x": feats, "x_lens": [torch.tensor(n) for n in lens]}
    sut = PytorchSUT(sd, data, 8, "quant", None, enable_bf16=True, scenario="Offline", exact_order=True, head_fraction=0.3,
                     head_min_frames=0, compute_streams=compute_streams)
    settings = lg.TestSettings()
    settings.scenario = lg.TestScenario.Offline
    settings.mode = lg.TestMode.AccuracyOnly
    out = lg.StartTest(sut.sut, sut.qsl.qsl, settings)
    assert out["completed"] == len(lens)
    got = {e["qsl_idx"]: np.frombuffer(bytes.fromhex(e["data"]), np.int32) for e in out["accuracy_log"]}
    order = sorted(range(len(lens)), key=lambda i: -lens[i])
    plan = sharding.plan_chunks([lens[i] for i in order], 8, 0.3, 0)
    assert len(plan) == 2 and plan != [(8, 13), (0, 8)]                  # the short end really was cut by frames
    for lo, hi in plan:
        idx = order[lo:hi]
        o_res, o_len, _ = _oracle(params, [feats[i] for i in idx], pad_to=32)
        for j, i in enumerate(idx):
            np.testing.assert_array_equal(got[i], o_res[j, :o_len[j]])


def test_sut_offline_multi_tile_query_vs_oracle(setup):
    """A query of 420 utterances at batch 160 through the pipelined path: three chunks (the short end cut by frames, one full
    batch), each with two 128-row encoder tiles of different lengths and a batch width that is not a multiple of 128 (ragged
    row skipping inside straddling GEMM tiles); bytes equal to the oracle's in the planner's batches."""
    from rnnt_inference_b200 import loadgen_shim as lg
    from rnnt_inference_b200 import sharding
    from rnnt_inference_b200.pytorch_sut import PytorchSUT
    sd, _rnnt, params, _ = setup
    rng = np.random.default_rng(21)
    lens = [int(v) for v in rng.integers(6, 64, 420)]
    feats = [S.make_features(n, 2000 + i) for i, n in enumerate(lens)]
    data = {"x": feats, "x_lens": [torch.tensor(n) for n in lens]}
    sut = PytorchSUT(sd, data, 160, "quant", None, enable_bf16=True, scenario="Offline", exact_order=True, head_fraction=0.3,
                     head_min_frames=0)
    settings = lg.TestSettings()
    settings.scenario = lg.TestScenario.Offline
    settings.mode = lg.TestMode.AccuracyOnly
    out = lg.StartTest(sut.sut, sut.qsl.qsl, settings)
    assert out["completed"] == len(lens)
    got = {e["qsl_idx"]: np.frombuffer(bytes.fromhex(e["data"]), np.int32) for e in out["accuracy_log"]}
    order = sorted(range(len(lens)), key=lambda i: (-lens[i], i))
    plan = sharding.plan_chunks([lens[i] for i in order], 160, 0.3, 0)
    assert len(plan) == 3 and plan[-1] == (0, 160) and plan != sharding.plan_chunks([lens[i] for i in order], 160, 0.0, 0)
    for lo, hi in plan:
        idx = order[lo:hi]
        o_res, o_len, _ = _oracle(params, [feats[i] for i in idx], pad_to=32)
        for j, i in enumerate(idx):
            np.testing.assert_array_equal(got[i], o_res[j, :o_len[j]])


def test_sut_default_path_and_sync_inference(setup):
    """The SUT as a user runs it (tensor-core hoist): the pipelined Offline path and the synchronous
    inference() + query_samples_complete() path of the reference (pytorch_sut.py:67-118) give the same bytes, and
    against the oracle every difference is a near-tie."""
    from oracle import oracle as O
    from rnnt_inference_b200 import eval_accuracy as E
    from rnnt_inference_b200 import loadgen_shim as lg
    from rnnt_inference_b200.pytorch_sut import PytorchSUT
    sd, _rnnt, params, _ = setup
    lens = [55, 7, 61, 38, 12, 61, 3, 42, 9, 26, 80, 33]
    feats = [S.make_features(n, 900 + i) for i, n in enumerate(lens)]
    data = {"x": feats, "x_lens": [torch.tensor(n) for n in lens]}
    sut = PytorchSUT(sd, data, 8, "quant", None, enable_bf16=True, scenario="Offline")
    settings = lg.TestSettings()
    settings.scenario = lg.TestScenario.Offline
    settings.mode = lg.TestMode.AccuracyOnly
    out = lg.StartTest(sut.sut, sut.qsl.qsl, settings)
    assert out["completed"] == len(lens)
    got = {e["qsl_idx"]: list(np.frombuffer(bytes.fromhex(e["data"]), np.int32)) for e in out["accuracy_log"]}
    order = sorted(range(len(lens)), key=lambda i: -lens[i])
    # synchronous path on the first chunk
    idx = order[:8]
    res, res_len = sut.inference(idx)
    assert res.shape[0] == 32 and res_len.shape[0] == 32                    # N padded to a multiple of 32 (pytorch_sut.py:90-104)
    res, res_len = res.cpu().numpy(), res_len.cpu().numpy()
    for j, i in enumerate(idx):
        assert list(res[j, :res_len[j]]) == got[i]
    assert (res_len[8:] == 0).all()
    # against the oracle, chunk by chunk
    for c in range(0, len(order), 8):
        idx = order[c:c + 8]
        x, xl = OP.pad_batch([feats[i] for i in idx], pad_to=32)
        f_o = OP.encode(params, x, xl)
        f_lens = ((xl + 1) // 2).astype(np.int32)
        ds = OP.new_decode_state(x.shape[1], int(xl.max()))
        trace = O.new_decision_trace(x.shape[1], 2 * int(f_lens.max()) + 64)
        O.greedy_decode(f_o, f_lens, params["dec"], ds["res"], ds["res_idx"], ds["symbols_added"], ds["time_idx"], ds["pre_g"],
                        ds["pre_hg"], ds["pre_cg"], trace=trace)
        truth = [list(ds["res"][j, :ds["res_idx"][j] + 1]) for j in range(len(idx))]
        sub = {k: v[:len(idx)] for k, v in trace.items()}
        rep = E.compare_transcripts([got[i] for i in idx], truth, sub, margin_tol=MARGIN_TOL)
        assert rep["unexplained"] == [], rep["differences"]


def test_qsl_arena_and_staging_runs(setup):
    """The QSL's pinned arena holds the samples longest first (ties by id) and hands out views of it; the SUT's staging
    copies maximal runs of arena neighbours with one DMA each — a chunk of neighbours and a chunk with gaps must both
    arrive on the device as the plain concatenation of their samples."""
    from rnnt_inference_b200 import loadgen_shim as lg
    from rnnt_inference_b200.pytorch_sut import PytorchSUT
    sd = setup[0]
    lens = [25, 7, 31, 18, 12, 31, 3, 22, 9, 16]
    feats = [S.make_features(n, 500 + i) for i, n in enumerate(lens)]
    sut = PytorchSUT(sd, {"x": feats, "x_lens": [torch.tensor(n) for n in lens]}, 4, "quant", None, enable_bf16=True,
                     scenario="Offline")
    qsl = sut.qsl
    assert qsl.arena is not None and qsl.arena.is_pinned()
    order = sorted(range(len(lens)), key=qsl.sort_key)
    assert order == sorted(range(len(lens)), key=lambda i: (-lens[i], i))
    off = 0
    for i in order:
        assert qsl.arena_off[i] == off
        x, n = qsl[i]
        assert int(n) == lens[i] and x.data_ptr() == qsl.arena[off:].data_ptr()
        assert torch.equal(x, feats[i])
        off += lens[i]
    for ids in (order[2:6], [order[0], order[3], order[4], order[9]]):      # one run / three runs
        chunk = [lg.QuerySample(100 + i, i) for i in ids]
        b = sut._stage(chunk)
        b.ready.synchronize()
        total = sum(lens[i] for i in ids)
        assert torch.equal(b.packed[:total].cpu(), torch.cat([feats[i] for i in ids], 0))
        assert b.meta[1, :len(ids)].cpu().tolist() == [lens[i] for i in ids]


def test_server_sut_slot_batching_vs_oracle():
    """Server scenario (csrc/torch_sut.cpp:354-571, metadata.cpp:111-194): samples arrive one by one, run in 8 slots with
    chunked encoding (split_len 8, state carried), early response after >= 2 finished slots, long samples parked.
    Every transcript must equal the oracle's for that utterance decoded with the same chunk length (checkpoint without
    burst frames, so that a transcript cannot depend on what shares the batch; exact_order: bit-exact)."""
    import time
    from rnnt_inference_b200 import loadgen_shim as lg
    from rnnt_inference_b200.server_sut import ServerSUT
    sd = S.make_checkpoint(seed=11, burst_threshold=1e9)
    params = OP.prepare_params(sd)
    lens = [25, 7, 31, 18, 12, 44, 3, 22, 9, 16, 37, 28, 5, 40, 14, 20, 33, 11, 26, 8]
    feats = [S.make_features(n, 1300 + i) for i, n in enumerate(lens)]
    data = {"x": feats, "x_lens": [torch.tensor(n) for n in lens]}
    sut = ServerSUT(sd, data, batch_size=8, split_len=8, response_size=2, length_threshold=38, exact_order=True)
    try:
        settings = lg.TestSettings()
        settings.scenario = lg.TestScenario.Server
        settings.mode = lg.TestMode.AccuracyOnly
        settings.server_target_qps = 400.0
        settings.server_duration_s = 1.0
        settings.min_query_count = len(lens)
        out = lg.StartTest(sut.sut, sut.qsl.qsl, settings)
    finally:
        sut.close()
    assert out["completed"] == len(lens)
    assert out["latency_p99_ms"] is not None and sut.iterations >= 3
    got = {e["qsl_idx"]: list(np.frombuffer(bytes.fromhex(e["data"]), np.int32)) for e in out["accuracy_log"]}
    total = 0
    for i, f in enumerate(feats):
        x, xl = OP.pad_batch([f], pad_to=32)
        o_res, o_len, _ = OP.greedy_decode_batch(params, x, xl, split_len=8)
        assert got[i] == list(o_res[0, :o_len[0]]), f"utterance {i}"
        total += int(o_len[0])
    assert total > 20
