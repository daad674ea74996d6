"""CPU restatement (numpy) of the reference's audio front-end: waveform -> normalised, spliced log-mel features.

TEST INFRASTRUCTURE ONLY: tests/ compare the CUDA front-end (rnnt_inference_b200/features.py) against this file; nothing
in the product imports it.

Follows datasets/parts/features.py:186-252 (FilterbankFeatures.forward) step by step, and the plugin operators it calls:
  preemphasis + reflect padding   mlperf_plugins/csrc/preemphasis.cpp:10-46, tpps/preemphasis_tpp.cpp:8-52,
                                  tpps/reflection_copy_tpp.cpp:26-44
  torch.stft(512, hop 160, win 320 hann, center=False)     features.py:201-210
  power_spectrum                  mlperf_plugins/csrc/power_spectrum.cpp:10-33
  dither ** 2, fb (x) power + 1e-20, log                   features.py:217-227
  frame_splicing (x 3)            mlperf_plugins/csrc/frame_splicing.cpp:12-45, tpps/frame_splicing_tpp.cpp:12-28
  i_layernorm_pad                 mlperf_plugins/csrc/normalization.cpp:10-64, tpps/i_layernorm_tpp.cpp:10-76

PARITY: the three operator KATs of the reference (mlperf_plugins/test/pyfiles/test_{preemphasis,power_spectrum,
frame_splicing}.py) are replayed against this file and against the GPU operators in tests/test_frontend.py.  The mel
matrix comes from librosa.filters.mel in the reference (conda package, unpinned: docker/Dockerfile:44); librosa is absent
here, so `mel_filterbank` restates its documented algorithm (Slaney scale, htk=False, norm="slaney") — PARITY UNPINNED for
that matrix and therefore for the end-to-end feature values; every other step is pinned by the KATs or is plain arithmetic.
"""
import numpy as np

F32 = np.float32


def preemphasis(x, lens, coeff=0.97, pad_size=0):
    """x [N, L] fp32, lens [N].  out [N, L + 2 pad]: y[i] = x[i] - coeff * x[i-1] (x[-1] = 0) for i < len, reflect-padded
    by pad on both sides (edge not repeated), zeros behind."""
    x = np.asarray(x, F32)
    n, l = x.shape
    out = np.zeros((n, l + 2 * pad_size), F32)
    for i in range(n):
        ln = int(lens[i])
        y = x[i, :ln].copy()
        y[1:] = x[i, 1:ln] - F32(coeff) * x[i, :ln - 1]
        if pad_size > 0:
            y = np.pad(y, (pad_size, pad_size), mode="reflect")
        out[i, :y.shape[0]] = y
    return out


def hann_window(win_length):
    """torch.hann_window(win_length, periodic=False).float()"""
    k = np.arange(win_length, dtype=np.float64)
    return (0.5 - 0.5 * np.cos(2.0 * np.pi * k / (win_length - 1))).astype(F32)


def stft(x, n_fft=512, hop=160, window=None):
    """torch.stft(center=False, return_complex=False) of x [N, Lp]: frames t = 0 .. (Lp - n_fft) // hop, the window
    zero-padded to n_fft on both sides.  Returns [N, T, F = n_fft/2 + 1, 2] fp32 (features.py:211 permutes to this)."""
    x = np.asarray(x, F32)
    n, lp = x.shape
    t = 1 + (lp - n_fft) // hop
    win = np.zeros(n_fft, np.float64)
    if window is None:
        win[:] = 1.0
    else:
        left = (n_fft - len(window)) // 2
        win[left:left + len(window)] = window
    idx = np.arange(t)[:, None] * hop + np.arange(n_fft)[None, :]
    frames = x[:, idx].astype(np.float64) * win
    spec = np.fft.rfft(frames, axis=-1)
    return np.stack([spec.real, spec.imag], -1).astype(F32)


def power_spectrum(x, lens):
    """x [N, T, F, 2] -> re^2 + im^2 [N, T, F], frames t >= lens[n] zeroed."""
    x = np.asarray(x, F32)
    out = (x[..., 0] * x[..., 0] + x[..., 1] * x[..., 1]).astype(F32)
    for i in range(x.shape[0]):
        out[i, int(lens[i]):] = 0
    return out


def _hz_to_mel(f):
    f = np.asarray(f, np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz, logstep = 1000.0, np.log(6.4) / 27.0
    min_log_mel = min_log_hz / f_sp
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-30) / min_log_hz) / logstep, mels)


def _mel_to_hz(m):
    m = np.asarray(m, np.float64)
    f_sp = 200.0 / 3
    min_log_hz, logstep = 1000.0, np.log(6.4) / 27.0
    min_log_mel = min_log_hz / f_sp
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


def mel_filterbank(sr=16000, n_fft=512, n_mels=80, fmin=0.0, fmax=None):
    """librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax) with its defaults (htk=False, norm='slaney', float32): triangular
    filters on the Slaney mel scale, each scaled to unit area (2 / (f_hi - f_lo)).  [n_mels, n_fft/2 + 1]."""
    fmax = sr / 2.0 if fmax is None else fmax
    fftfreqs = np.linspace(0.0, sr / 2.0, 1 + n_fft // 2)
    mel_f = _mel_to_hz(np.linspace(_hz_to_mel(fmin), _hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = mel_f[:, None] - fftfreqs[None, :]
    w = np.zeros((n_mels, 1 + n_fft // 2), np.float64)
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        w[i] = np.maximum(0.0, np.minimum(lower, upper))
    w *= (2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels]))[:, None]
    return w.astype(F32)


def frame_splicing(x, lens, factor=3):
    """x [N, C, T] -> [N, C * factor, ceil(T / factor)]: out[n, k*C + c, j] = x[n, c, factor*j + k] if that frame is
    < lens[n], else 0."""
    x = np.asarray(x, F32)
    n, c, t = x.shape
    t2 = (t + factor - 1) // factor
    out = np.zeros((n, c * factor, t2), F32)
    for i in range(n):
        ln = int(lens[i])
        for k in range(factor):
            src = np.arange(t2) * factor + k
            ok = src < min(ln, t)
            out[i, k * c:(k + 1) * c, ok] = x[i][:, src[ok]].T
    return out


def layernorm_pad(x, lens, eps=1e-12, unbiased=True, padded_batch=None, out_feat=None):
    """Per (sample, feature) normalisation over the valid frames: var = E[x^2] - mean^2 + eps, times len / (len - 1) when
    unbiased; y = (x - mean) / sqrt(var).  Output [padded_batch, out_feat, max(lens)]; here everything the reference
    leaves unwritten (features >= C, frames >= len, padding samples) is zero."""
    x = np.asarray(x, F32)
    n, c, _t = x.shape
    padded_batch = n if padded_batch in (None, 1) else padded_batch
    out_feat = c if out_feat is None else out_feat
    max_len = int(np.max(lens))
    out = np.zeros((padded_batch, out_feat, max_len), F32)
    for i in range(n):
        ln = int(lens[i])
        if ln == 0:
            continue
        v = x[i, :, :ln].astype(np.float64)
        mean = v.mean(axis=1, keepdims=True)
        var = (v * v).mean(axis=1, keepdims=True) - mean * mean + eps
        if unbiased and ln > 1:
            var = var * (ln / (ln - 1.0))
        out[i, :c, :ln] = ((v - mean) / np.sqrt(var)).astype(F32)
    return out


def filterbank_features(x, x_lens, pad_batch_size=False, sample_rate=16000, window_size=0.02, window_stride=0.01, n_fft=512,
                        nfilt=80, preemph=0.97, dither=1e-5, frame_splice=3, pad_out_feat=True, fb=None):
    """FilterbankFeatures.forward (features.py:186-252) with the MLPerf RNN-T configuration (configs/rnnt.toml input_eval).
    x [N, L] fp32 waveforms, x_lens [N] samples.  Returns (features [N_pad, 256 | 240, T'], lens [N_pad])."""
    x_lens = np.asarray(x_lens, np.int64)
    win_length, hop = int(sample_rate * window_size), int(sample_rate * window_stride)
    n = x.shape[0]
    y = preemphasis(x, x_lens, preemph, n_fft // 2)
    spec = stft(y, n_fft, hop, hann_window(win_length))
    lens = np.floor(x_lens / hop + 1).astype(np.int32)
    p = power_spectrum(spec, lens).transpose(0, 2, 1)                      # [N, F, T]
    p = p + F32(dither) ** 2
    fb = mel_filterbank(sample_rate, n_fft, nfilt) if fb is None else fb
    mel = np.einsum("mf,nft->nmt", fb.astype(np.float64), p.astype(np.float64)) + 1e-20
    logmel = np.log(mel).astype(F32)
    sp = frame_splicing(logmel, lens, frame_splice)
    lens2 = np.ceil(lens / frame_splice).astype(np.int32)
    out_feat = sp.shape[1]
    if pad_out_feat:
        out_feat = (out_feat + 31) // 32 * 32
    n_pad = (n + 31) // 32 * 32 if pad_batch_size else n
    feats = layernorm_pad(sp, lens2, 1e-12, True, n_pad, out_feat)
    out_lens = np.zeros(n_pad, np.int32)
    out_lens[:n] = lens2
    return feats, out_lens
