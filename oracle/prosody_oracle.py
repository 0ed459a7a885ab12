"""CPU oracle for the prosody service -- TEST INFRASTRUCTURE ONLY.

`AubioYinOracle` mimics the `aubio.pitch('yin', buf, hop, sr)` object the reference
constructs (/root/reference/backend/services/prosody.py:32-34) on top of the C
restatement in oracle/yin_aubio.c. `ProsodyOracle.analyze_buffer` then follows
/root/reference/backend/services/prosody.py:58-104 statement by statement.

PARITY UNPINNED for per-frame f0 (aubio itself is not importable here); the tag-level
facts the reference's tests pin (test_input_processing.py:447-505) are checked in
tests/test_oracle_prosody.py.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import build as _build

_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        L = ctypes.CDLL(_build.build())
        L.yin_oracle_new.restype = ctypes.c_void_p
        L.yin_oracle_new.argtypes = [ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32,
                                     ctypes.c_float, ctypes.c_float]
        L.yin_oracle_del.argtypes = [ctypes.c_void_p]
        L.yin_oracle_reset.argtypes = [ctypes.c_void_p]
        L.yin_oracle_get_state.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.yin_oracle_set_state.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.yin_oracle_lags.restype = ctypes.c_uint64
        L.yin_oracle_lags.argtypes = [ctypes.c_void_p]
        L.yin_oracle_do.restype = ctypes.c_float
        L.yin_oracle_do.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.yin_oracle_run.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64,
                                     ctypes.c_void_p, ctypes.c_void_p]
        L.np_mean_f32.restype = ctypes.c_float
        L.np_mean_f32.argtypes = [ctypes.c_void_p, ctypes.c_int64]
        _lib = L
    return _lib


class AubioYinOracle:
    """Stand-in for aubio.pitch('yin', buf_size, hop_size, samplerate)."""

    def __init__(self, buf_size: int = 4096, hop_size: int = 512, samplerate: int = 48000,
                 tolerance: float = 0.15, silence_db: float = -50.0) -> None:
        self.buf_size, self.hop_size, self.samplerate = buf_size, hop_size, samplerate
        self._tol, self._sil = tolerance, silence_db
        self._h = lib().yin_oracle_new(buf_size, hop_size, samplerate, tolerance, silence_db)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().yin_oracle_del(self._h)
            self._h = None

    def _recreate(self):
        state = self.state
        lib().yin_oracle_del(self._h)
        self._h = lib().yin_oracle_new(self.buf_size, self.hop_size, self.samplerate, self._tol, self._sil)
        self.state = state

    def set_unit(self, unit: str) -> None:
        if unit not in ("Hz", "default"):
            raise ValueError("oracle only restates unit 'Hz'")

    def set_tolerance(self, tol: float) -> None:
        self._tol = float(tol)
        self._recreate()

    def set_silence(self, db: float) -> None:
        self._sil = float(db)
        self._recreate()

    def reset(self) -> None:
        lib().yin_oracle_reset(self._h)

    @property
    def state(self) -> np.ndarray:
        out = np.empty(self.buf_size, np.float32)
        lib().yin_oracle_get_state(self._h, out.ctypes.data)
        return out

    @state.setter
    def state(self, v) -> None:
        v = np.ascontiguousarray(v, np.float32)
        assert v.shape == (self.buf_size,)
        lib().yin_oracle_set_state(self._h, v.ctypes.data)

    @property
    def lags_evaluated(self) -> int:
        return int(lib().yin_oracle_lags(self._h))

    def __call__(self, chunk: np.ndarray) -> np.ndarray:
        chunk = np.ascontiguousarray(chunk, np.float32)
        if chunk.shape != (self.hop_size,):
            raise ValueError(f"input size of {chunk.shape} does not match hop_size {self.hop_size}")
        return np.array([lib().yin_oracle_do(self._h, chunk.ctypes.data, None)], np.float32)

    def run(self, pcm: np.ndarray):
        """Whole-buffer hop loop in C: returns (f0[n_hops], level_db[n_hops])."""
        pcm = np.ascontiguousarray(pcm, np.float32)
        n_hops = -(-len(pcm) // self.hop_size)
        f0 = np.empty(n_hops, np.float32)
        db = np.empty(n_hops, np.float32)
        lib().yin_oracle_run(self._h, pcm.ctypes.data, len(pcm), f0.ctypes.data, db.ctypes.data)
        return f0, db


def np_mean_f32(a: np.ndarray) -> np.float32:
    a = np.ascontiguousarray(a, np.float32)
    return np.float32(lib().np_mean_f32(a.ctypes.data, a.size))


class ProsodyOracle:
    """Follows /root/reference/backend/services/prosody.py:10-104."""

    def __init__(self, sample_rate: int = 48000, hop_size: int = 512) -> None:
        self.sample_rate = sample_rate
        self.hop_size = hop_size
        self.pitch_detector = AubioYinOracle(4096, hop_size, sample_rate)   # prosody.py:32
        self.pitch_detector.set_unit('Hz')                                  # :33
        self.pitch_detector.set_tolerance(0.8)                              # :34

    def analyze_buffer(self, audio_buffer, python_hop_loop: bool = True, details: dict | None = None):
        if isinstance(audio_buffer, list):                                   # :58-59
            audio_buffer = np.concatenate(audio_buffer)
        if not isinstance(audio_buffer, np.ndarray):                         # :61-62
            audio_buffer = np.array(audio_buffer, dtype=np.float32)
        if audio_buffer.dtype != np.float32:                                 # :64-65
            audio_buffer = audio_buffer.astype(np.float32)
        rms = np.sqrt(np.mean(audio_buffer ** 2))                            # :67
        if rms < 0.05:                                                       # :69-74
            energy_tag = 'Quiet'
        elif rms < 0.15:
            energy_tag = 'Normal'
        else:
            energy_tag = 'Loud'
        if python_hop_loop:                                                  # :76-87 (the reference's loop shape)
            pitch_values = []
            all_f0 = []
            total = len(audio_buffer)
            for i in range(0, total, self.hop_size):
                chunk = audio_buffer[i:i + self.hop_size]
                if len(chunk) < self.hop_size:
                    chunk = np.pad(chunk, (0, self.hop_size - len(chunk)), mode='constant')
                pitch = self.pitch_detector(chunk)[0]
                all_f0.append(pitch)
                if pitch > 0.0:
                    pitch_values.append(pitch)
            f0 = np.array(all_f0, np.float32)
        else:                                                                # same arithmetic, loop in C
            f0, _ = self.pitch_detector.run(audio_buffer)
            pitch_values = list(f0[f0 > 0.0])
        if len(pitch_values) > 0:                                            # :89-99
            avg_pitch = np.mean(pitch_values)
            if avg_pitch < 120:
                pitch_tag = 'Deep'
            elif avg_pitch < 200:
                pitch_tag = 'Normal'
            else:
                pitch_tag = 'High'
        else:
            avg_pitch = None
            pitch_tag = 'Normal'
        if details is not None:
            details.update(rms=rms, f0=f0, avg_pitch=avg_pitch)
        return {'energy': energy_tag, 'pitch': pitch_tag}                    # :101-104


def vad_gate_ref(f0: np.ndarray, hop_db: np.ndarray, hops_per_chunk: int, level_db: float = -40.0,
                 min_voiced: int = 1) -> np.ndarray:
    """CPU restatement of janus_vad_gate_f32 (janus_b200/csrc/yin.cu): chunk = hops_per_chunk consecutive hops; speech
    iff its loudest hop exceeds level_db and at least min_voiced hops are voiced. (An energy gate with two thresholds in the
    place where the reference asks Silero VAD, backend/services/vad.py:40-77 -- not a restatement of Silero.)"""
    f0 = np.asarray(f0, np.float32).reshape(-1, hops_per_chunk)
    db = np.asarray(hop_db, np.float32).reshape(-1, hops_per_chunk)
    return (db.max(axis=1) > np.float32(level_db)) & ((f0 > 0).sum(axis=1) >= min_voiced)
