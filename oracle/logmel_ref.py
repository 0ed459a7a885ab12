"""CPU oracle for the log-mel frontend -- TEST INFRASTRUCTURE ONLY.

Restates, in numpy float32, the feature extraction that the reference reaches through
`WhisperModel.transcribe(audio_16k, ...)` (/root/reference/backend/services/transcriber.py:53-57).
The arithmetic lives in the un-vendored third-party package faster-whisper (pinned only as
`faster-whisper>=1.0.0`, /root/reference/backend/requirements.txt:2), class
`faster_whisper.feature_extractor.FeatureExtractor` (1.1.x): append 160 zero samples, centred STFT
(reflect pad n_fft/2 = 200), n_fft 400, hop 160, periodic Hann window, power spectrum of the first
201 bins, drop the last frame, Slaney mel filterbank (80 or 128 bands, norm="slaney"),
log10(clip(.,1e-10)), clamp to (global max over the utterance) - 8, (x + 4) / 4; then, per 30 s
window, `pad_or_trim` of the FEATURES to 3000 frames with 0.0.

PARITY UNPINNED: faster-whisper is not installable here and the reference's tests mock the model
(test_input_processing.py:372-429). The restatement is cross-checked against the independent
HF `transformers.WhisperFeatureExtractor` in tests/test_oracle_stt.py.
"""
from __future__ import annotations

import numpy as np

N_FFT = 400
HOP = 160
SR = 16000
N_FRAMES = 3000
TAIL_PAD = 160  # FeatureExtractor.__call__(padding=160)


def mel_filters(n_mels: int = 80, sr: int = SR, n_fft: int = N_FFT) -> np.ndarray:
    """Slaney-scale, slaney-normalised triangular filterbank, float32 [n_mels, n_fft//2+1]."""
    fftfreqs = np.fft.rfftfreq(n=n_fft, d=1.0 / sr)
    min_mel, max_mel = 0.0, 45.245640471924965  # hz_to_mel(8000) on the Slaney scale
    mels = np.linspace(min_mel, max_mel, n_mels + 2)
    f_min, f_sp = 0.0, 200.0 / 3
    freqs = f_min + f_sp * mels
    min_log_hz = 1000.0
    min_log_mel = (min_log_hz - f_min) / f_sp
    logstep = np.log(6.4) / 27.0
    log_t = mels >= min_log_mel
    freqs[log_t] = min_log_hz * np.exp(logstep * (mels[log_t] - min_log_mel))
    fdiff = np.diff(freqs)
    ramps = freqs.reshape(-1, 1) - fftfreqs.reshape(1, -1)
    lower = -ramps[:-2] / np.expand_dims(fdiff[:-1], axis=1)
    upper = ramps[2:] / np.expand_dims(fdiff[1:], axis=1)
    weights = np.maximum(np.zeros_like(lower), np.minimum(lower, upper))
    enorm = 2.0 / (freqs[2:n_mels + 2] - freqs[:n_mels])
    weights *= np.expand_dims(enorm, axis=1)
    return weights.astype(np.float32)


def hann_window() -> np.ndarray:
    return np.hanning(N_FFT + 1)[:-1].astype(np.float32)


def log_mel_utterance(audio: np.ndarray, n_mels: int = 80, padding: int = TAIL_PAD) -> np.ndarray:
    """Whole-utterance features, float32 [n_mels, n_frames] (before 30 s windowing)."""
    x = np.asarray(audio, dtype=np.float32)
    if padding:
        x = np.pad(x, (0, padding))
    x = np.pad(x, (N_FFT // 2, N_FFT // 2), mode="reflect")
    n_frames = 1 + (len(x) - N_FFT) // HOP
    idx = np.arange(N_FFT)[None, :] + HOP * np.arange(n_frames)[:, None]
    frames = x[idx] * hann_window()[None, :]
    spec = np.fft.rfft(frames, axis=1).astype(np.complex64)      # [n_frames, 201]
    magnitudes = (np.abs(spec[:-1]) ** 2).T.astype(np.float32)   # drop last frame -> [201, n_frames-1]
    mel = mel_filters(n_mels) @ magnitudes
    log_spec = np.log10(np.clip(mel, a_min=1e-10, a_max=None))
    log_spec = np.maximum(log_spec, log_spec.max() - 8.0)
    return ((log_spec + 4.0) / 4.0).astype(np.float32)


def log_mel_window(audio: np.ndarray, n_mels: int = 80, seek: int = 0) -> np.ndarray:
    """The encoder input for the 30 s window starting at frame `seek`: float32 [n_mels, 3000],
    zero-padded in feature space (faster-whisper pad_or_trim)."""
    feats = log_mel_utterance(audio, n_mels)
    content_frames = feats.shape[-1] - 1          # transcribe(): content_frames = features.shape[-1] - 1
    segment_size = min(N_FRAMES, content_frames - seek)
    seg = feats[:, seek:seek + max(segment_size, 0)]
    if seg.shape[1] < N_FRAMES:                   # pad_or_trim(segment) with 0.0 in FEATURE space
        seg = np.pad(seg, ((0, 0), (0, N_FRAMES - seg.shape[1])))
    return np.ascontiguousarray(seg, dtype=np.float32)
