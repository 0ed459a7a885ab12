"""Per-chunk speech gating for MANY sessions on the GPU.

The reference gates one microphone per 1536-sample chunk (32 ms at 48 kHz) with Silero VAD on the CPU
(/root/reference/backend/services/vad.py:40-77, called from engine.py:471-475). Silero's model is a download, so the
gate here is an energy decision built from what the prosody kernel already computes per hop: every
session keeps its own YIN sliding buffer on the device (as the reference's aubio object does), one launch of
`yin_frames_kernel` serves the current chunk of ALL sessions, and `vad_gate_kernel` turns the three hops of a chunk
into one flag. `MultiSessionSegmenter.push_tick` then runs the reference's utterance state machine per session.
It is NOT claimed to reproduce Silero's decisions (the speech decision is an input of the state machine).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _abi
from .services.prosody import ProsodyEngine

CHUNK_SIZE = 1536   # audio_io.py:29


class GpuVadGate:
    def __init__(self, n_sessions: int, device=None, sample_rate: int = 48000, chunk_size: int = CHUNK_SIZE,
                 hop_size: int = 512, level_db: float = -40.0, min_voiced: int = 1) -> None:
        if chunk_size % hop_size:
            raise ValueError("chunk_size must be a multiple of hop_size")
        self.n, self.chunk, self.hpc = n_sessions, chunk_size, chunk_size // hop_size
        self.level_db, self.min_voiced = float(level_db), int(min_voiced)
        self.engine = ProsodyEngine(device, sample_rate=sample_rate, hop_size=hop_size)
        dev = self.device = self.engine.device
        with torch.cuda.device(dev):
            self.plan = self.engine.plan([chunk_size] * n_sessions)
            self.out = self.engine.alloc_outputs(self.plan)
            self.state = torch.zeros((n_sessions, self.engine.buf_size), dtype=torch.float32, device=dev)
            self.out["state"] = torch.empty_like(self.state)
            self.pcm = torch.empty(n_sessions * chunk_size, dtype=torch.float32, device=dev)
            self.flags = torch.empty(n_sessions, dtype=torch.uint8, device=dev)
            self.level = torch.empty(n_sessions, dtype=torch.float32, device=dev)
            self.host_pcm = torch.empty(n_sessions * chunk_size, dtype=torch.float32, pin_memory=True)
            self.host_flags = torch.empty(n_sessions, dtype=torch.uint8, pin_memory=True)

    def step_device(self, pcm: torch.Tensor) -> torch.Tensor:
        """pcm: float32 device [n_sessions * chunk] (the current chunk of every session). Returns the device flags
        (uint8 [n_sessions]); advances every session's sliding buffer. Does not synchronise."""
        eng = self.engine
        with torch.cuda.device(self.device):
            out = eng.run(pcm, self.plan, state_in=self.state, want_state=True, out=self.out)
            self.state, self.out["state"] = out["state"], self.state      # ping-pong: no copy
            _abi.call("janus_vad_gate_f32", out["f0"].data_ptr(), out["hop_db"].data_ptr(), self.n, self.hpc,
                      self.level_db, self.min_voiced, self.flags.data_ptr(), self.level.data_ptr(),
                      torch.cuda.current_stream(self.device).cuda_stream)
        return self.flags

    def step(self, chunks: np.ndarray) -> np.ndarray:
        """chunks: float32 [n_sessions, chunk] host array -> bool [n_sessions] (H2D + kernels + D2H, synchronises)."""
        assert chunks.shape == (self.n, self.chunk)
        self.host_pcm.numpy()[:] = np.ascontiguousarray(chunks, dtype=np.float32).reshape(-1)
        with torch.cuda.device(self.device):
            self.pcm.copy_(self.host_pcm, non_blocking=True)
            self.step_device(self.pcm)
            self.host_flags.copy_(self.flags, non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
        return self.host_flags.numpy().astype(bool)
