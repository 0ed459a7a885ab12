"""Synthetic audio for parity tests and bench.py.

Each clip is a sequence of "notes" of 0.5-2 s: a harmonic tone (6 harmonics, amplitudes 1/k) with a
piecewise-constant f0 drawn log-uniformly from [80, 400] Hz, an overall amplitude ~U[0.02, 0.5]
(so all three energy tags occur), additive white noise at an SNR ~U[10, 40] dB, and 10-30 % of
the notes replaced by silence (digital zero, exercising aubio's -50 dB gate) or by noise only
(unvoiced, exercising YIN's no-early-exit path).

Written in torch so it runs on the CPU (tests, fixtures) or directly on the GPU (bench: 4096 x 30 s
would take minutes in numpy). Generation is per-device deterministic given `seed`; parity tests
always feed the SAME generated tensor to both the oracle and the kernels.
"""
from __future__ import annotations

import math

import torch


def make_clips(n: int, seconds: float, sample_rate: int = 16000, seed: int = 1234,
               device: str | torch.device = "cpu", chunk: int = 128) -> torch.Tensor:
    """Returns float32 [n, int(seconds*sample_rate)]."""
    dev = torch.device(device)
    N = int(round(seconds * sample_rate))
    out = torch.empty((n, N), dtype=torch.float32, device=dev)
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    max_notes = int(seconds / 0.5) + 2
    t_idx = torch.arange(N, device=dev, dtype=torch.float32) / sample_rate
    for c0 in range(0, n, chunk):
        c = min(chunk, n - c0)
        u = lambda *s: torch.rand(*s, generator=g, device=dev)  # noqa: E731
        dur = 0.5 + 1.5 * u(c, max_notes)
        ends = torch.cumsum(dur, dim=1)
        note = torch.searchsorted(ends, t_idx.expand(c, N).contiguous(), right=True).clamp_(max=max_notes - 1)
        f0_note = 80.0 * torch.exp(u(c, max_notes) * math.log(400.0 / 80.0))
        kind_u = u(c, max_notes)
        frac_off = 0.1 + 0.2 * u(c, 1)               # 10-30 % of notes are not voiced
        silent = kind_u < frac_off * 0.5              # half of them digital silence
        noise_only = (kind_u >= frac_off * 0.5) & (kind_u < frac_off)
        amp = 0.02 + 0.48 * u(c, 1)
        snr_db = 10.0 + 30.0 * u(c, 1)
        f0 = torch.gather(f0_note, 1, note)
        phase = torch.cumsum(f0.double() * (2.0 * math.pi / sample_rate), dim=1)
        phase = torch.remainder(phase, 2.0 * math.pi).float()
        tone = torch.zeros((c, N), dtype=torch.float32, device=dev)
        for k in range(1, 7):
            tone += torch.sin(phase * k) / k
        tone *= amp / 1.6  # sum of 1/k harmonics peaks near 1.6
        voiced = ~(torch.gather(silent, 1, note) | torch.gather(noise_only, 1, note))
        tone *= voiced
        sig_rms = amp / 1.6 * math.sqrt(0.5 * sum(1.0 / k ** 2 for k in range(1, 7)))
        noise_rms = sig_rms * torch.pow(10.0, -snr_db / 20.0)
        noise = torch.randn((c, N), generator=g, device=dev) * noise_rms
        noise *= ~torch.gather(silent, 1, note)
        out[c0:c0 + c] = tone + noise
    return out


def sine_wave(frequency: float = 440.0, duration: float = 1.0, sample_rate: int = 48000,
              amplitude: float = 0.5):
    """The reference tests' generator (test_input_processing.py:30-45): endpoint-inclusive linspace."""
    import numpy as np
    t = np.linspace(0, duration, int(sample_rate * duration), dtype=np.float32)
    return (amplitude * np.sin(2 * np.pi * frequency * t, dtype=np.float32)).astype(np.float32)
