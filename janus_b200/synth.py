"""Synthetic audio for parity tests and bench.py.

Each clip is a sequence of "notes" of 0.5-2 s: a harmonic tone (6 harmonics, amplitudes 1/k) with a
piecewise-constant f0 drawn log-uniformly from [80, 400] Hz, an overall amplitude ~U[0.02, 0.5]
(so all three energy tags occur), additive white noise at an SNR ~U[10, 40] dB, and 10-30 % of
the notes replaced by silence (digital zero, exercising aubio's -50 dB gate) or by noise only
(unvoiced, exercising YIN's no-early-exit path).

Written in torch so it runs on the CPU (tests, fixtures) or directly on the GPU (bench: 4096 x 30 s
would take minutes in numpy). Every clip has its own seed (`seed + i`): its
ground truth (notes, f0, amplitude, SNR) is identical on every device, its samples are deterministic
per device type (the noise comes from the device's generator; sin / cumsum differ in the last bit
between CPU and CUDA). Parity tests always feed the SAME generated tensor to both the oracle and
the kernels; bench.py copies the windows it checks back from the device.
"""
from __future__ import annotations

import math

import torch


def clip_params(seed: int, seconds: float) -> dict:
    """Ground truth of ONE clip, drawn from `np.random.default_rng(seed)` on the host (SURVEY.md section 8d:
    clip i of a batch uses seed 1234 + i), so any window of any batch can be regenerated alone, on any device:
    note durations / f0 / kinds, overall amplitude, SNR."""
    import numpy as np
    rng = np.random.default_rng(seed)
    max_notes = int(seconds / 0.5) + 2
    dur = 0.5 + 1.5 * rng.random(max_notes)
    f0 = 80.0 * np.exp(rng.random(max_notes) * math.log(400.0 / 80.0))
    kind_u = rng.random(max_notes)
    frac_off = 0.1 + 0.2 * rng.random()            # 10-30 % of notes are not voiced
    return {"ends": np.cumsum(dur), "f0": f0, "silent": kind_u < frac_off * 0.5,   # half of them digital silence
            "noise_only": (kind_u >= frac_off * 0.5) & (kind_u < frac_off),
            "amp": 0.02 + 0.48 * rng.random(), "snr_db": 10.0 + 30.0 * rng.random()}


def make_clips(n: int, seconds: float, sample_rate: int = 16000, seed: int = 1234,
               device: str | torch.device = "cpu", chunk: int = 128, first: int = 0) -> torch.Tensor:
    """Returns float32 [n, int(seconds*sample_rate)]: clips first .. first+n-1 of the stream of clips named by
    `seed`. Clip i depends on nothing but (seed + i, seconds, sample_rate, device type): its notes come from
    `clip_params(seed + i)`, its noise from a torch generator seeded seed + i on `device`."""
    import numpy as np
    dev = torch.device(device)
    N = int(round(seconds * sample_rate))
    out = torch.empty((n, N), dtype=torch.float32, device=dev)
    t_idx = torch.arange(N, device=dev, dtype=torch.float32) / sample_rate
    g = torch.Generator(device=dev)
    sig_unit = math.sqrt(0.5 * sum(1.0 / k ** 2 for k in range(1, 7))) / 1.6
    for c0 in range(0, n, chunk):
        c = min(chunk, n - c0)
        P = [clip_params(seed + first + c0 + i, seconds) for i in range(c)]
        f64 = lambda key: torch.from_numpy(np.stack([np.asarray(p[key], dtype=np.float64) for p in P]))  # noqa: E731
        ends = f64("ends").float().to(dev)
        max_notes = ends.shape[1]
        note = torch.searchsorted(ends, t_idx.expand(c, N).contiguous(), right=True).clamp_(max=max_notes - 1)
        f0 = torch.gather(f64("f0").float().to(dev), 1, note)
        silent = torch.gather(torch.from_numpy(np.stack([p["silent"] for p in P])).to(dev), 1, note)
        noise_only = torch.gather(torch.from_numpy(np.stack([p["noise_only"] for p in P])).to(dev), 1, note)
        amp = f64("amp").float().to(dev)[:, None]
        snr_db = f64("snr_db").float().to(dev)[:, None]
        phase = torch.cumsum(f0.double() * (2.0 * math.pi / sample_rate), dim=1)
        phase = torch.remainder(phase, 2.0 * math.pi).float()
        tone = torch.zeros((c, N), dtype=torch.float32, device=dev)
        for k in range(1, 7):
            tone += torch.sin(phase * k) / k
        tone *= amp / 1.6  # sum of 1/k harmonics peaks near 1.6
        tone *= ~(silent | noise_only)
        noise_rms = amp * sig_unit * torch.pow(10.0, -snr_db / 20.0)
        noise = torch.empty((c, N), dtype=torch.float32, device=dev)
        for i in range(c):
            g.manual_seed(seed + first + c0 + i)
            noise[i].normal_(generator=g)
        noise *= noise_rms
        noise *= ~silent
        out[c0:c0 + c] = tone + noise
    return out


def sine_wave(frequency: float = 440.0, duration: float = 1.0, sample_rate: int = 48000,
              amplitude: float = 0.5):
    """The reference tests' generator (test_input_processing.py:30-45): endpoint-inclusive linspace."""
    import numpy as np
    t = np.linspace(0, duration, int(sample_rate * duration), dtype=np.float32)
    return (amplitude * np.sin(2 * np.pi * frequency * t, dtype=np.float32)).astype(np.float32)
