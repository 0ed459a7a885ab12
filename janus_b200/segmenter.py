"""Utterance segmentation in front of the path, for many sessions.

The reference gates ONE microphone inside `smart_ear_loop`
(/root/reference/backend/services/engine.py:438-506): 1536-sample chunks at 48 kHz, a pre-roll of the last
10 non-speech chunks, an utterance that ends after MORE than 15 consecutive non-speech chunks (~0.5 s),
and a minimum length of 1536*6 samples. `UtteranceSegmenter` is that state machine for one session, chunk
for chunk; `MultiSessionSegmenter` runs one per session and hands finished utterances to the batched
engine (`pipeline.SemanticEncoder`), which is what BASELINE configs[4] (many concurrent sessions) needs.

The speech/non-speech decision is an input: the reference asks Silero VAD (services/vad.py:40-77), whose
model must be downloaded and is therefore not available here. `energy_gate` is a stand-in with the same
call shape (chunk -> bool); it is NOT claimed to reproduce Silero's decisions.
"""
from __future__ import annotations

from collections import deque
from typing import Callable

import numpy as np

CHUNK_SIZE = 1536                 # audio_io.py:29
PRE_ROLL_CHUNKS = 10              # engine.py:439
SILENCE_THRESHOLD_CHUNKS = 15     # engine.py:441 (trigger when the counter EXCEEDS it, engine.py:491)
MIN_UTTERANCE_SAMPLES = 1536 * 6  # engine.py:504


def energy_gate(threshold_db: float = -40.0) -> Callable[[np.ndarray], bool]:
    """chunk -> bool by mean-square level in dB (10 log10 mean x^2), the same level definition the prosody
    kernel's silence gate uses. A stand-in for services/vad.py, see module docstring."""
    def is_speech(chunk: np.ndarray) -> bool:
        x = np.asarray(chunk, dtype=np.float32)
        if x.size == 0:
            return False
        ms = float(np.mean(x * x))
        return ms > 0.0 and 10.0 * np.log10(ms) > threshold_db
    return is_speech


class UtteranceSegmenter:
    """The streaming-mode branch of engine.py:471-506 for one session."""

    def __init__(self) -> None:
        self.audio_buffer: list[np.ndarray] = []
        self.pre_roll: deque = deque(maxlen=PRE_ROLL_CHUNKS)
        self.silence_counter = 0
        self.is_talking = False
        self.skipped_short = 0

    def push(self, chunk: np.ndarray, is_speech: bool) -> np.ndarray | None:
        """Feeds one chunk; returns the finished utterance (float32) when this chunk ends one, else None."""
        trigger = False
        if is_speech:                                        # engine.py:476-483
            if len(self.audio_buffer) == 0:
                self.audio_buffer.extend(list(self.pre_roll))
            self.is_talking = True
            self.audio_buffer.append(chunk)
            self.silence_counter = 0
        else:                                                # engine.py:484-493
            self.silence_counter += 1
            if len(self.audio_buffer) > 0:
                self.audio_buffer.append(chunk)
            else:
                self.pre_roll.append(chunk)
            if self.silence_counter > SILENCE_THRESHOLD_CHUNKS:
                trigger = True
                self.is_talking = False
        if trigger and len(self.audio_buffer) > 0:           # engine.py:499-506
            combined = np.concatenate(self.audio_buffer)
            self.audio_buffer = []
            self.silence_counter = 0
            if len(combined) < MIN_UTTERANCE_SAMPLES:
                self.skipped_short += 1
                return None
            return combined
        return None

    def flush(self) -> np.ndarray | None:
        """PTT-release style end (engine.py:465-469): whatever is buffered becomes an utterance."""
        if not self.audio_buffer:
            return None
        combined = np.concatenate(self.audio_buffer)
        self.audio_buffer = []
        self.silence_counter = 0
        if len(combined) < MIN_UTTERANCE_SAMPLES:
            self.skipped_short += 1
            return None
        return combined


class MultiSessionSegmenter:
    """One `UtteranceSegmenter` per session id; collects finished utterances for the batched engine."""

    def __init__(self, is_speech: Callable[[np.ndarray], bool] | None = None) -> None:
        self.is_speech = is_speech or energy_gate()
        self.sessions: dict = {}
        self.ready: list[tuple[object, np.ndarray]] = []

    def push(self, session_id, chunk: np.ndarray, speech: bool | None = None) -> None:
        seg = self.sessions.get(session_id)
        if seg is None:
            seg = self.sessions[session_id] = UtteranceSegmenter()
        utt = seg.push(chunk, self.is_speech(chunk) if speech is None else speech)
        if utt is not None:
            self.ready.append((session_id, utt))

    def push_tick(self, session_ids, chunks: np.ndarray, speech) -> None:
        """One chunk of EVERY session at once: chunks [n, CHUNK_SIZE], speech [n] bool (e.g. the flags of
        janus_b200.vad.GpuVadGate.step for this tick)."""
        for sid, chunk, sp in zip(session_ids, chunks, speech):
            self.push(sid, chunk, bool(sp))

    def take(self, max_items: int | None = None) -> list[tuple[object, np.ndarray]]:
        """Removes and returns up to `max_items` finished (session_id, utterance) pairs, oldest first."""
        n = len(self.ready) if max_items is None else min(max_items, len(self.ready))
        out, self.ready = self.ready[:n], self.ready[n:]
        return out
