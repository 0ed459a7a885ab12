"""Host side of the log-mel frontend (K1): Slaney mel filterbank construction + launch wrapper.

The filterbank matches what faster-whisper's FeatureExtractor builds (Slaney scale, slaney norm,
n_fft 400 at 16 kHz); it is handed to the kernel in sparse row form because each triangular filter
touches at most 14 of the 201 bins.
"""
from __future__ import annotations

import numpy as np
import torch

from .. import _abi
from .._device import on_device

N_FFT, HOP, SR, N_BINS, N_FRAMES = 400, 160, 16000, 201, 3000


def _hz_of_mel(m: np.ndarray) -> np.ndarray:
    """Slaney mel scale: linear below 1 kHz (200/3 Hz per mel), logarithmic above (6.4x per 27 mel)."""
    lin = m * (200.0 / 3.0)
    knee_mel = 1000.0 / (200.0 / 3.0)
    log = 1000.0 * np.exp((np.log(6.4) / 27.0) * (m - knee_mel))
    return np.where(m >= knee_mel, log, lin)


def slaney_mel_filterbank(n_mels: int) -> np.ndarray:
    """float32 [n_mels, 201] triangular filters, area-normalised (slaney)."""
    bin_hz = np.arange(N_BINS, dtype=np.float64) * (SR / N_FFT)
    top_mel = 45.245640471924965  # mel(8000 Hz)
    edges = _hz_of_mel(np.linspace(0.0, top_mel, n_mels + 2))
    fb = np.zeros((n_mels, N_BINS), dtype=np.float64)
    for m in range(n_mels):
        lo, ctr, hi = edges[m], edges[m + 1], edges[m + 2]
        rise = (bin_hz - lo) / (ctr - lo)
        fall = (hi - bin_hz) / (hi - ctr)
        fb[m] = np.clip(np.minimum(rise, fall), 0.0, None) * (2.0 / (hi - lo))
    return fb.astype(np.float32)


class LogMelFrontend:
    def __init__(self, n_mels: int, device: torch.device) -> None:
        self.n_mels, self.device = n_mels, device
        fb = slaney_mel_filterbank(n_mels)
        self.filters = fb
        starts, lens = [], []
        for row in fb:
            nz = np.nonzero(row)[0]
            starts.append(int(nz[0]) if len(nz) else 0)
            lens.append(int(nz[-1] - nz[0] + 1) if len(nz) else 0)
        self.max_len = max(lens)
        packed = np.zeros((n_mels, self.max_len), np.float32)
        for m, (s, l) in enumerate(zip(starts, lens)):
            packed[m, :l] = fb[m, s:s + l]
        self.mel_start = torch.tensor(starts, dtype=torch.int32, device=device)
        self.mel_len = torch.tensor(lens, dtype=torch.int32, device=device)
        self.mel_w = torch.from_numpy(packed).to(device)
        self.frames_alloc = int(_abi.lib().janus_logmel_frames_alloc())

    @on_device
    def __call__(self, pcm: torch.Tensor, offsets: torch.Tensor, lengths: torch.Tensor, decim: int = 1,
                 want_bf16: bool = True, want_f32: bool = False, scratch=None, out_bf16=None):
        """pcm float32 device (streams concatenated), offsets/lengths int64 device [n] (source-rate
        samples). Returns (mel_pad bf16 [n, 3002, n_mels] | None, mel fp32 [n, n_mels, 3000] | None)."""
        n = int(offsets.numel())
        dev = self.device
        if scratch is None:
            scratch = torch.empty((n, self.frames_alloc, self.n_mels), dtype=torch.float32, device=dev)
        maxbuf = torch.empty(n, dtype=torch.int32, device=dev)
        if want_bf16 and out_bf16 is None:
            out_bf16 = torch.empty((n, N_FRAMES + 2, self.n_mels), dtype=torch.bfloat16, device=dev)
        out_f32 = torch.empty((n, self.n_mels, N_FRAMES), dtype=torch.float32, device=dev) if want_f32 else None
        _abi.call("janus_logmel_f32", pcm.data_ptr(), offsets.data_ptr(), lengths.data_ptr(), n, decim, self.n_mels,
                  self.mel_start.data_ptr(), self.mel_len.data_ptr(), self.mel_w.data_ptr(), self.max_len,
                  scratch.data_ptr(), maxbuf.data_ptr(), out_bf16.data_ptr() if want_bf16 else None,
                  out_f32.data_ptr() if want_f32 else None, torch.cuda.current_stream(dev).cuda_stream)
        return (out_bf16 if want_bf16 else None), out_f32

    # ---- utterances longer than one window ------------------------------------------------------
    @staticmethod
    def n_frames(n_samples: int, decim: int = 1) -> int:
        """Feature frames faster-whisper keeps for an utterance of n_samples source-rate samples
        (+160 zero samples, centred STFT at hop 160, last frame dropped); content_frames = n_frames - 1."""
        return (-(-n_samples // decim) + 160) // HOP

    @on_device
    def utterance(self, pcm: torch.Tensor, offsets: torch.Tensor, lengths: torch.Tensor, nf_alloc: int, decim: int = 1):
        """Pass 1 over whole utterances: (scratch fp32 [n, nf_alloc, n_mels] log10-mel, maxbuf int32 [n])."""
        n, dev = int(offsets.numel()), self.device
        scratch = torch.empty((n, nf_alloc, self.n_mels), dtype=torch.float32, device=dev)
        maxbuf = torch.empty(n, dtype=torch.int32, device=dev)
        _abi.call("janus_logmel_utterance_f32", pcm.data_ptr(), offsets.data_ptr(), lengths.data_ptr(), n, decim,
                  self.n_mels, self.mel_start.data_ptr(), self.mel_len.data_ptr(), self.mel_w.data_ptr(), self.max_len,
                  scratch.data_ptr(), nf_alloc, maxbuf.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
        return scratch, maxbuf

    @on_device
    def windows(self, scratch: torch.Tensor, maxbuf: torch.Tensor, lengths: torch.Tensor, win_utt, win_seek,
                decim: int = 1, want_f32: bool = False):
        """Pass 2: window w = frames [win_seek[w], +3000) of utterance win_utt[w], clamped with that utterance's
        global max - 8 and zero-padded in feature space. Returns (mel_pad bf16 [n_w, 3002, n_mels], fp32 | None)."""
        dev = self.device
        wu = torch.as_tensor(win_utt, dtype=torch.int32).to(dev)
        ws = torch.as_tensor(win_seek, dtype=torch.int32).to(dev)
        nw = int(wu.numel())
        out_bf16 = torch.empty((nw, N_FRAMES + 2, self.n_mels), dtype=torch.bfloat16, device=dev)
        out_f32 = torch.empty((nw, self.n_mels, N_FRAMES), dtype=torch.float32, device=dev) if want_f32 else None
        _abi.call("janus_logmel_windows_f32", scratch.data_ptr(), maxbuf.data_ptr(), lengths.data_ptr(), decim,
                  self.n_mels, scratch.shape[1], wu.data_ptr(), ws.data_ptr(), nw, out_bf16.data_ptr(),
                  out_f32.data_ptr() if want_f32 else None, torch.cuda.current_stream(dev).cuda_stream)
        return out_bf16, out_f32
