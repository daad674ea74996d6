"""Audio front-end on the GPU: waveform -> normalised, spliced log-mel features.

Host-side mirror of the reference's featurizer, same names and arguments:
  FilterbankFeatures   datasets/parts/features.py:94-252 (forward: :186-252)
  AudioProcessing      datasets/process_librispeech.py:100-111
The arithmetic runs in CUDA kernels behind the C ABI (csrc/frontend.cu): `preemphasis` -> ONE kernel for
torch.stft + `power_spectrum` + dither^2 + mel filterbank + log (the spectrum never reaches HBM) -> `frame_splicing` ->
`i_layernorm_pad`; the reference's operators of those names are also exposed one by one in `_C.py`.

The mel matrix is librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax) in the reference; librosa is not available offline, so
`mel_filterbank` restates its documented algorithm (Slaney scale, unit-area triangles).  Pass `fb=` to use another matrix.
"""
import math
from typing import Tuple

import numpy as np
import torch
from torch import nn

from . import _C as P
from . import ops

CONSTANT = 1e-5


def _hz_to_mel(f):
    f = np.asarray(f, np.float64)
    f_sp, min_log_hz, logstep = 200.0 / 3, 1000.0, math.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_hz / f_sp + np.log(np.maximum(f, 1e-30) / min_log_hz) / logstep, f / f_sp)


def _mel_to_hz(m):
    m = np.asarray(m, np.float64)
    f_sp, min_log_hz, logstep = 200.0 / 3, 1000.0, math.log(6.4) / 27.0
    return np.where(m >= min_log_hz / f_sp, min_log_hz * np.exp(logstep * (m - min_log_hz / f_sp)), f_sp * m)


def mel_filterbank(sr, n_fft, n_mels, fmin=0.0, fmax=None):
    """[n_mels, n_fft / 2 + 1] fp32: triangular filters with corners equally spaced on the Slaney mel scale, each scaled to
    unit area — librosa.filters.mel with its defaults (htk=False, norm="slaney")."""
    fmax = sr / 2.0 if fmax is None else fmax
    freqs = np.linspace(0.0, sr / 2.0, 1 + n_fft // 2)
    corners = _mel_to_hz(np.linspace(_hz_to_mel(fmin), _hz_to_mel(fmax), n_mels + 2))
    width = np.diff(corners)
    ramps = corners[:, None] - freqs[None, :]
    rising, falling = -ramps[:-2] / width[:-1, None], ramps[2:] / width[1:, None]
    w = np.maximum(0.0, np.minimum(rising, falling))
    w *= (2.0 / (corners[2:] - corners[:-2]))[:, None]
    return torch.from_numpy(w.astype(np.float32))


class FilterbankFeatures(nn.Module):
    def __init__(self, sample_rate=8000, window_size=0.02, window_stride=0.01, window="hamming", normalize="per_feature",
                 n_fft=None, preemph=0.97, nfilt=64, lowfreq=0, highfreq=None, log=True, dither=CONSTANT, pad_to=8,
                 max_duration=16.7, frame_splicing=1, pad_out_feat=False, fb=None):
        super().__init__()
        windows = {"hann": torch.hann_window, "hamming": torch.hamming_window, "blackman": torch.blackman_window,
                   "bartlett": torch.bartlett_window}
        if window not in windows:
            raise RuntimeError(f"FilterbankFeatures: window must be one of {sorted(windows)} (got {window!r})")
        if not log or normalize != "per_feature":
            raise RuntimeError("FilterbankFeatures: the GPU front-end implements the MLPerf configuration (log=True, "
                               "normalize='per_feature')")
        self.frame_splicing = frame_splicing
        self.win_length = int(sample_rate * window_size)
        self.hop_length = int(sample_rate * window_stride)
        self.n_fft = n_fft or 2 ** math.ceil(math.log2(self.win_length))
        self.normalize, self.log, self.dither, self.preemph = normalize, log, dither, preemph
        self.use_deterministic_dithering = True          # features.py:137: always on
        highfreq = highfreq or sample_rate / 2
        self.register_buffer("window", windows[window](self.win_length, periodic=False).float())
        fb = mel_filterbank(sample_rate, self.n_fft, nfilt, lowfreq, highfreq) if fb is None else torch.as_tensor(fb).float()
        self.register_buffer("fb", fb.unsqueeze(0).contiguous())
        self.register_buffer("fb_bias", torch.full((1, fb.shape[0], 1), 1e-20))
        max_length = 1 + math.ceil((max_duration * sample_rate - self.win_length) / self.hop_length)
        self.max_length = max_length + (16 - (max_length % 16))
        self.in_feat = fb.shape[0]
        self.out_feat = self.in_feat * self.frame_splicing
        if pad_out_feat:
            self.out_feat = math.ceil(self.out_feat / 32) * 32
        self.output_shape = torch.tensor((1, self.out_feat, self.max_length), dtype=torch.int32)

    def get_seq_len(self, seq_len):
        seq_len = torch.floor(seq_len / self.hop_length + 1).to(dtype=torch.int32)
        if self.frame_splicing > 1:
            seq_len = torch.ceil(seq_len / self.frame_splicing).to(dtype=torch.int32)
        return seq_len

    @torch.no_grad()
    def forward(self, x: torch.Tensor, x_lens: torch.Tensor, pad_batch_size: bool) -> Tuple[torch.Tensor, torch.Tensor]:
        """x fp32 [N, L] waveforms (zero-padded), x_lens [N] samples.  Returns (features [N_pad, out_feat, max T'], lens)."""
        dev = self.window.device
        if dev.type != "cuda":
            raise RuntimeError("FilterbankFeatures runs on the GPU only: move the module to a CUDA device (there is no CPU path)")
        actual_bs = x.shape[0]
        x = x.to(dev, torch.float32)
        x_lens = x_lens.to(dev)
        if self.preemph is not None:
            x = P.preemphasis(x, x_lens.to(torch.int32), coeff=self.preemph, pad_size=self.n_fft // 2)
        else:
            x = torch.nn.functional.pad(x, (self.n_fft // 2, self.n_fft // 2))
        x_lens = torch.floor(x_lens / self.hop_length + 1).to(dtype=torch.int32)
        dither2 = float(self.dither) ** 2 if self.dither > 0 else 0.0
        x = ops.logmel(x, x_lens, self.window, self.fb[0], self.n_fft, self.hop_length, dither2, 1e-20)      # [N, nfilt, T]
        if self.frame_splicing > 1:
            x = P.frame_splicing(x, x_lens, self.frame_splicing)
        x_lens = torch.ceil(x_lens / self.frame_splicing).to(dtype=torch.int32)
        shape = self.output_shape.clone()
        if pad_batch_size:
            shape[0] = (actual_bs + 31) // 32 * 32
        return P.i_layernorm_pad(x, None, None, x_lens, 1e-12, unbiased=1, output_shape=shape)

    @classmethod
    def from_config(cls, cfg, log=False):
        return cls(sample_rate=cfg["sample_rate"], window_size=cfg["window_size"], window_stride=cfg["window_stride"],
                   n_fft=cfg["n_fft"], nfilt=cfg["features"], window=cfg["window"], normalize=cfg["normalize"],
                   max_duration=cfg.get("max_duration", 16.7), dither=cfg["dither"], pad_to=cfg.get("pad_to", 0),
                   frame_splicing=cfg.get("frame_splicing", 1), log=log, pad_out_feat=cfg.get("pad_out_feat"))


# configs/rnnt.toml [input_eval]: the MLPerf RNN-T evaluation front-end
MLPERF_INPUT_EVAL = dict(sample_rate=16000, window_size=0.02, window_stride=0.01, window="hann", features=80, n_fft=512,
                         frame_splicing=3, dither=0.00001, normalize="per_feature", pad_to=0, max_duration=16.7)


class AudioProcessing(nn.Module):
    """datasets/process_librispeech.py:100-111: AudioProcessing(run_mode, **featurizer_config)(wavs, wav_lens, pad_batch_size)."""

    def __init__(self, run_mode, **kwargs):
        super().__init__()
        cfg = dict(kwargs)
        cfg["pad_out_feat"] = run_mode == "quant"
        cfg.pop("pad_batch_size", None)
        feat_type = cfg.get("feat_type", "logfbank")                          # FeatureFactory.from_config (features.py:281-285)
        self.featurizer = FilterbankFeatures.from_config(cfg, log="log" in feat_type)

    def forward(self, wavs: torch.Tensor, wav_lens: torch.Tensor, pad_batch_size: bool) -> Tuple[torch.Tensor, torch.Tensor]:
        return self.featurizer(wavs, wav_lens, pad_batch_size)
