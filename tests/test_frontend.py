"""Audio front-end: the oracle against torch formulas (the reference's own operator KATs,
mlperf_plugins/test/pyfiles/test_{preemphasis,power_spectrum,frame_splicing}.py, restated), and the CUDA operators /
FilterbankFeatures / the SUT's --enable_process path against the oracle.

Floating point (fp32 like the reference).  Tolerances: preemphasis rtol 1e-6 (the KAT's), power spectrum and frame splicing
exact, log-mel 2e-3 absolute (fp32 FFT vs the oracle's float64 FFT, log of sums spanning ~1e8), layer norm 1e-3 (the
reference's fp32 E[x^2] - mean^2 against float64), normalised features end to end 5e-3."""
import numpy as np
import pytest
import torch

from oracle import frontend as F


def _waves(seed, lens):
    rng = np.random.default_rng(seed)
    x = np.zeros((len(lens), max(lens)), np.float32)
    for i, n in enumerate(lens):
        t = np.arange(n) / 16000.0
        x[i, :n] = (0.1 * rng.normal(size=n) + 0.3 * np.sin(2 * np.pi * (200 + 150 * i) * t) * np.sin(2 * np.pi * 3 * t)).astype(np.float32)
    return x


# ------------------------------------------------------------------------------------------------ oracle (CPU)
@pytest.mark.parametrize("seq_len", [17, 16, 18])
def test_oracle_preemphasis_kat(seq_len):
    """test_preemphasis.py:14-43 restated."""
    bs, preemph, n_fft, hop = 3, 0.97, 6, 4
    x = torch.tensor(np.arange(bs * seq_len).reshape(bs, seq_len), dtype=torch.float32)
    length = torch.tensor([seq_len, seq_len - hop, seq_len - hop - 1], dtype=torch.int32)
    for i in range(bs):
        x[i, length[i]:] = 0
    y_exp = torch.cat((x[:, 0].unsqueeze(1), x[:, 1:] - preemph * x[:, :-1]), dim=1)
    padded = torch.zeros((bs, seq_len + n_fft))
    for i in range(bs):
        tmp = torch.nn.functional.pad(y_exp[i:i + 1, :length[i]], (n_fft // 2, n_fft // 2), "reflect")
        padded[i:i + 1, :tmp.shape[1]] = tmp
    np.testing.assert_allclose(F.preemphasis(x.numpy(), length.numpy(), preemph, n_fft // 2), padded.numpy(), rtol=1e-6)


def test_oracle_stft_power_splice_norm_vs_torch():
    lens = [4000, 3700, 1601]
    x = _waves(0, lens)
    y = F.preemphasis(x, lens, 0.97, 256)
    spec = F.stft(y, 512, 160, F.hann_window(320))
    ref = torch.view_as_real(torch.stft(torch.tensor(y), 512, 160, 320, window=torch.hann_window(320, periodic=False), center=False,
                                        return_complex=True)).permute(0, 2, 1, 3).numpy()
    np.testing.assert_allclose(spec, ref, atol=2e-5)
    flens = [n // 160 + 1 for n in lens]
    p = F.power_spectrum(spec, flens)
    e = (ref * ref).sum(-1)
    for i, n in enumerate(flens):
        e[i, n:] = 0
    np.testing.assert_allclose(p, e, rtol=1e-4, atol=1e-6)
    # frame splicing: test_frame_splicing.py:14-37 restated
    for seq_len in (48, 49, 50, 51, 52, 97):
        xs = torch.tensor(np.arange(seq_len * 3 * 2).reshape(3, 2, seq_len), dtype=torch.float32)
        length = torch.tensor([seq_len, seq_len - 1, seq_len - 2], dtype=torch.int32)
        if seq_len > 96:
            length[0] -= 48
        got = F.frame_splicing(xs.numpy(), length.numpy(), 3)
        for i in range(3):
            xs[i, :, length[i]:] = 0
        seq = [xs]
        for n in range(1, 3):
            tmp = torch.zeros_like(xs)
            tmp[:, :, :-n] = xs[:, :, n:]
            seq.append(tmp)
        np.testing.assert_equal(got, torch.cat(seq, dim=1)[:, :, ::3].numpy())
    # per-feature normalisation with unbiased std
    v = np.random.default_rng(1).normal(2.0, 3.0, (2, 5, 40)).astype(np.float32)
    out = F.layernorm_pad(v, [40, 17], 1e-12, True, 32, 8)
    assert out.shape == (32, 8, 40)
    t = torch.tensor(v[1, :, :17])
    np.testing.assert_allclose(out[1, :5, :17], ((t - t.mean(1, keepdim=True)) / t.std(1, unbiased=True, keepdim=True)).numpy(), atol=2e-5)
    assert not out[1, :, 17:].any() and not out[2:].any() and not out[:, 5:].any()


def test_mel_filterbank_shape_and_area():
    """The restated librosa.filters.mel: 80 triangles over 257 bins, Slaney normalisation = unit area in Hz, corners ordered."""
    fb = F.mel_filterbank(16000, 512, 80)
    assert fb.shape == (80, 257) and fb.dtype == np.float32 and (fb >= 0).all()
    peaks = fb.argmax(1)
    assert (np.diff(peaks) >= 0).all() and peaks[0] >= 1 and peaks[-1] <= 255
    hz_per_bin = 8000.0 / 256
    area = fb.sum(1) * hz_per_bin
    assert np.all(np.abs(area[10:] - 1.0) < 0.15)            # discretised triangles of unit area
    from rnnt_inference_b200.features import mel_filterbank
    np.testing.assert_array_equal(mel_filterbank(16000, 512, 80).numpy(), fb)


# ------------------------------------------------------------------------------------------------ CUDA operators
gpu = pytest.mark.gpu


@gpu
@pytest.mark.parametrize("seq_len", [17, 16, 18])
def test_gpu_preemphasis_kat(seq_len):
    from rnnt_inference_b200 import _C as P
    bs, preemph, n_fft, hop = 3, 0.97, 6, 4
    x = torch.tensor(np.arange(bs * seq_len).reshape(bs, seq_len), dtype=torch.float32)
    length = torch.tensor([seq_len, seq_len - hop, seq_len - hop - 1], dtype=torch.int32)
    for i in range(bs):
        x[i, length[i]:] = 0
    y = P.preemphasis(x.cuda(), length.cuda(), coeff=preemph, pad_size=n_fft // 2).cpu().numpy()
    np.testing.assert_allclose(y, F.preemphasis(x.numpy(), length.numpy(), preemph, n_fft // 2), rtol=1e-6)


@gpu
@pytest.mark.parametrize("seq_len", [16, 17, 25])
def test_gpu_power_spectrum_kat(seq_len):
    """test_power_spectrum.py:12-26 through the GPU operator: exact."""
    from rnnt_inference_b200 import _C as P
    bs, freq = 3, 7
    x = torch.tensor(np.arange(bs * seq_len * freq * 2).reshape(bs, seq_len, freq, 2), dtype=torch.float32)
    length = torch.tensor([seq_len, seq_len - 1, seq_len - 2], dtype=torch.int32)
    y = P.power_spectrum(x.cuda(), length.cuda()).cpu()
    for i in range(bs):
        x[i, length[i]:] = 0
    np.testing.assert_equal(y.numpy(), (x * x).sum(-1).numpy())


@gpu
@pytest.mark.parametrize("seq_len", [48, 49, 50, 51, 52, 97])
def test_gpu_frame_splicing_kat(seq_len):
    from rnnt_inference_b200 import _C as P
    x = torch.tensor(np.arange(seq_len * 3 * 2).reshape(3, 2, seq_len), dtype=torch.float32)
    length = torch.tensor([seq_len, seq_len - 1, seq_len - 2], dtype=torch.int32)
    if seq_len > 96:
        length[0] -= 48
    y = P.frame_splicing(x.cuda(), length.cuda(), 3).cpu().numpy()
    np.testing.assert_equal(y, F.frame_splicing(x.numpy(), length.numpy(), 3))
    with pytest.raises(RuntimeError):
        P.frame_splicing(x.cuda(), length.cuda(), 2)


@gpu
def test_gpu_layernorm_pad_vs_oracle():
    from rnnt_inference_b200 import _C as P
    v = np.random.default_rng(3).normal(-1.0, 2.5, (5, 240, 77)).astype(np.float32)
    lens = np.array([77, 60, 33, 2, 1], np.int32)
    out, out_len = P.i_layernorm_pad(torch.tensor(v).cuda(), None, None, torch.tensor(lens).cuda(), 1e-12, unbiased=1,
                                     output_shape=torch.tensor([32, 256, 1680], dtype=torch.int32))
    assert tuple(out.shape) == (32, 256, 77) and out_len.tolist() == list(lens) + [0] * 27
    ref = F.layernorm_pad(v, lens, 1e-12, True, 32, 256)
    # var = E[x^2] - mean^2 in fp32 like the reference (i_layernorm_tpp.cpp:39-41): cancellation on short rows, oracle in float64
    np.testing.assert_allclose(out.cpu().numpy()[:4], ref[:4], atol=1e-3)
    assert not out[4:].cpu().numpy()[:, :, 1:].any()           # a single frame: (x - mean) = 0; padding samples: zero


@gpu
def test_gpu_logmel_and_filterbank_features_vs_oracle():
    from rnnt_inference_b200 import ops
    from rnnt_inference_b200.features import AudioProcessing, MLPERF_INPUT_EVAL
    lens = [16000 * 3 + 123, 16000 * 2, 9000, 5000, 3000]
    x = _waves(7, lens)
    # the fused STFT / power / mel / log kernel on its own
    y = F.preemphasis(x, lens, 0.97, 256)
    flens = np.array([n // 160 + 1 for n in lens], np.int32)
    fb = F.mel_filterbank()
    spec = F.power_spectrum(F.stft(y, 512, 160, F.hann_window(320)), flens).transpose(0, 2, 1).astype(np.float64) + np.float32(1e-5) ** 2
    ref = np.log(np.einsum("mf,nft->nmt", fb.astype(np.float64), spec) + 1e-20)
    got = ops.logmel(torch.tensor(y).cuda(), torch.tensor(flens).cuda(), torch.tensor(F.hann_window(320)).cuda(),
                     torch.tensor(fb).cuda(), 512, 160, float(np.float32(1e-5) ** 2), 1e-20).cpu().numpy()
    assert got.shape == ref.shape
    np.testing.assert_allclose(got, ref, atol=2e-3)
    # the whole featurizer, as the SUT builds it
    proc = AudioProcessing("quant", pad_batch_size=True, **MLPERF_INPUT_EVAL).cuda().eval()
    feats, flen = proc(torch.tensor(x), torch.tensor(lens, dtype=torch.int32), True)
    o_feats, o_len = F.filterbank_features(x, lens, True)
    assert tuple(feats.shape) == o_feats.shape == (32, 256, int(o_len.max())) and flen.cpu().tolist() == o_len.tolist()
    # E[x^2] - mean^2 in fp32 (the reference's formula) cancels badly on rows of a few frames whose log-mel values are close:
    # the tight bound on utterances of >= 0.5 s, a loose one on the two very short ones
    got = feats.cpu().numpy()
    np.testing.assert_allclose(got[:3], o_feats[:3], atol=5e-3)
    np.testing.assert_allclose(got[3:], o_feats[3:], atol=1e-1)


@gpu
def test_sut_enable_process_runs_waveforms():
    """PytorchSUT(enable_process=True) (pytorch_sut.py:24-35, 70-82): waveform samples -> GPU featurizer -> decoder; the
    transcripts equal those of the decoder fed with the featurizer's output directly."""
    from rnnt_inference_b200 import loadgen_shim as lg
    from rnnt_inference_b200 import synthetic as S
    from rnnt_inference_b200.pytorch_sut import PytorchSUT
    lens = [16000 + 57, 12000, 9000, 20000, 6000]
    x = _waves(11, lens)
    data = {"x": [torch.tensor(x[i, :n]) for i, n in enumerate(lens)], "x_lens": [torch.tensor(n) for n in lens]}
    sd = S.make_checkpoint()
    sut = PytorchSUT(sd, data, 4, "quant", None, enable_bf16=True, scenario="Offline", enable_process=True)
    settings = lg.TestSettings()
    settings.scenario = lg.TestScenario.Offline
    settings.mode = lg.TestMode.AccuracyOnly
    out = lg.StartTest(sut.sut, sut.qsl.qsl, settings)
    assert out["completed"] == len(lens)
    got = {e["qsl_idx"]: list(np.frombuffer(bytes.fromhex(e["data"]), np.int32)) for e in out["accuracy_log"]}
    order = sorted(range(len(lens)), key=lambda i: -lens[i])
    for c in range(0, len(order), 4):
        idx = order[c:c + 4]
        wavs = torch.nn.utils.rnn.pad_sequence([data["x"][i] for i in idx], batch_first=True)
        feats, flen = sut.processor(wavs, torch.tensor([lens[i] for i in idx], dtype=torch.int32), True)
        res, res_len = sut.model(feats.permute(2, 0, 1).contiguous(), flen)
        res, res_len = res.cpu().numpy(), res_len.cpu().numpy()
        for j, i in enumerate(idx):
            assert got[i] == list(res[j, :res_len[j]])
        assert sum(len(got[i]) for i in idx) > 0
