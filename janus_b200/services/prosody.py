"""Prosody service on B200: YIN pitch + RMS energy -> categorical tags.

Drop-in for /root/reference/backend/services/prosody.py (class ProsodyExtractor, :10-104):
same constructor (`sample_rate=48000, hop_size=512`), same attributes (`sample_rate`,
`hop_size`, `pitch_detector`), same `analyze_buffer(audio_buffer) -> {'energy', 'pitch'}` with the
same coercions, thresholds and stateful sliding buffer -- but the per-hop aubio call
(prosody.py:76-87) is replaced by one launch of the bit-exact sm_100a YIN kernel
(csrc/yin.cu) over every hop of every stream, and RMS/np.mean/tagging by a second small kernel.

`ProsodyEngine` is the batched device-side engine (many independent streams per launch);
`ProsodyExtractor` is the single-stream service the reference's callers construct with no
arguments (/root/reference/backend/services/engine.py:413).
"""
from __future__ import annotations

import threading
from dataclasses import dataclass

import numpy as np
import torch

from .. import _abi

ENERGY_TAGS = ('Quiet', 'Normal', 'Loud')   # prosody.py:69-74
PITCH_TAGS = ('Deep', 'Normal', 'High')     # prosody.py:92-97

YIN_BUF_SIZE = 4096       # prosody.py:32
YIN_TOLERANCE = 0.8       # prosody.py:34
AUBIO_DEFAULT_TOLERANCE = 0.15
AUBIO_SILENCE_DB = -50.0  # aubio DEFAULT_PITCH_SILENCE (never changed by the reference)


def _coerce(audio_buffer) -> np.ndarray:
    """The reference's input coercion, prosody.py:58-65."""
    if isinstance(audio_buffer, list):
        audio_buffer = np.concatenate(audio_buffer)
    if not isinstance(audio_buffer, np.ndarray):
        audio_buffer = np.array(audio_buffer, dtype=np.float32)
    if audio_buffer.dtype != np.float32:
        audio_buffer = audio_buffer.astype(np.float32)
    return audio_buffer


@dataclass
class ProsodyPlan:
    """Host-computed index arrays for one batch of streams (tiny; one H2D copy)."""
    n_streams: int
    total_samples: int
    total_hops: int
    total_groups: int
    lengths: np.ndarray        # int64 [n]
    offsets: np.ndarray        # int64 [n]
    hop_offsets: np.ndarray    # int64 [n+1]
    group_offsets: np.ndarray  # int32 [n+1]
    dev: torch.Tensor          # int64 device tensor holding all four arrays back to back

    def ptrs(self):
        n = self.n_streams
        base = self.dev.data_ptr()
        return base, base + 8 * n, base + 16 * n, base + 16 * n + 8 * (n + 1)


class ProsodyEngine:
    """Batched YIN + RMS over many independent streams on one GPU."""

    def __init__(self, device: torch.device | str | int | None = None, sample_rate: int = 48000,
                 hop_size: int = 512, buf_size: int = YIN_BUF_SIZE, tolerance: float = YIN_TOLERANCE,
                 silence_db: float = AUBIO_SILENCE_DB) -> None:
        self._lib = _abi.lib()
        if not torch.cuda.is_available():
            raise _abi.JanusNativeError("janus_b200 needs a CUDA device (sm_100); there is no CPU fallback")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        if self.device.type != "cuda":
            raise _abi.JanusNativeError("janus_b200 runs on CUDA devices only")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        _abi.call("janus_check_device", self.device.index)
        self.sample_rate, self.hop_size, self.buf_size = int(sample_rate), int(hop_size), int(buf_size)
        self.tolerance, self.silence_db = float(tolerance), float(silence_db)
        self.frames_per_group = int(self._lib.janus_yin_frames_per_group())

    # ---- planning -------------------------------------------------------------------------
    def plan(self, lengths) -> ProsodyPlan:
        return make_plan(lengths, self.hop_size, self.frames_per_group, self.device)

    # ---- device-side run ------------------------------------------------------------------
    def run(self, pcm: torch.Tensor, plan: ProsodyPlan, state_in: torch.Tensor | None = None,
            want_state: bool = False, out: dict | None = None) -> dict:
        """pcm: float32 device tensor, streams concatenated per `plan`. Returns device tensors:
        f0[total_hops] Hz, hop_db[total_hops], rms[n], mean_f0[n], n_voiced[n], tags[n,2] uint8
        (+ 'state' [n, buf] when want_state). Nothing here synchronises the host."""
        assert pcm.dtype == torch.float32 and pcm.is_cuda and pcm.is_contiguous()
        assert pcm.numel() >= plan.total_samples
        n, dev = plan.n_streams, self.device
        with torch.cuda.device(dev):
            stream = torch.cuda.current_stream(dev).cuda_stream
            if out is None:
                out = self.alloc_outputs(plan)
            p_off, p_len, p_hop, p_grp = plan.ptrs()
            sp = 0
            if state_in is not None:
                assert state_in.shape == (n, self.buf_size) and state_in.dtype == torch.float32 \
                    and state_in.is_cuda and state_in.is_contiguous()
                sp = state_in.data_ptr()
            _abi.call("janus_yin_frames_f32", pcm.data_ptr(), p_off, p_len, p_hop, p_grp, n,
                      plan.total_groups, self.sample_rate, self.buf_size, self.hop_size,
                      self.tolerance, self.silence_db, sp, out["f0"].data_ptr(),
                      out["hop_db"].data_ptr(), out["hop_sumsq"].data_ptr(), stream)
            _abi.call("janus_prosody_reduce_f32", out["f0"].data_ptr(), out["hop_sumsq"].data_ptr(),
                      p_hop, p_len, n, out["scratch"].data_ptr(), out["rms"].data_ptr(),
                      out["mean_f0"].data_ptr(), out["n_voiced"].data_ptr(), out["tags"].data_ptr(),
                      stream)
            if want_state:
                if "state" not in out:
                    out["state"] = torch.empty((n, self.buf_size), dtype=torch.float32, device=dev)
                _abi.call("janus_yin_state_f32", pcm.data_ptr(), p_off, p_len, n, self.buf_size,
                          self.hop_size, sp, out["state"].data_ptr(), stream)
        return out

    def alloc_outputs(self, plan: ProsodyPlan) -> dict:
        dev, n, th = self.device, plan.n_streams, max(plan.total_hops, 1)
        f32 = dict(dtype=torch.float32, device=dev)
        return {
            "f0": torch.empty(th, **f32), "hop_db": torch.empty(th, **f32),
            "hop_sumsq": torch.empty(th, **f32), "scratch": torch.empty(th, **f32),
            "rms": torch.empty(n, **f32), "mean_f0": torch.empty(n, **f32),
            "n_voiced": torch.empty(n, dtype=torch.int32, device=dev),
            "tags": torch.empty((n, 2), dtype=torch.uint8, device=dev),
        }

    # ---- host-buffer convenience (H2D + kernels + D2H) --------------------------------------
    def analyze_host(self, buffers: list[np.ndarray], state_in: torch.Tensor | None = None,
                     want_state: bool = False, details: bool = False):
        """buffers: list of float32 numpy arrays (one per stream). Returns list of tag dicts
        (and, with details=True, a dict of host arrays: f0, hop_db, rms, mean_f0, n_voiced,
        hop_offsets). Synchronises."""
        lengths = [len(b) for b in buffers]
        plan = self.plan(lengths)
        host = torch.empty(max(plan.total_samples, 1), dtype=torch.float32, pin_memory=True)
        hv = host.numpy()
        for b, o, l in zip(buffers, plan.offsets, plan.lengths):
            hv[o:o + l] = b
        with torch.cuda.device(self.device):
            pcm = host.to(self.device, non_blocking=True)
            out = self.run(pcm, plan, state_in=state_in, want_state=want_state)
            tags = out["tags"].cpu().numpy()
            res = [{'energy': ENERGY_TAGS[e], 'pitch': PITCH_TAGS[p]} for e, p in tags]
            if not details:
                return (res, out.get("state")) if want_state else res
            det = {k: out[k].cpu().numpy() for k in ("f0", "hop_db", "rms", "mean_f0", "n_voiced")}
            det["f0"] = det["f0"][:plan.total_hops]
            det["hop_db"] = det["hop_db"][:plan.total_hops]
            det["hop_offsets"] = plan.hop_offsets
            det["tags"] = tags
            return (res, det, out.get("state")) if want_state else (res, det)


def make_plan(lengths, hop_size: int, frames_per_group: int, device) -> ProsodyPlan:
    lengths = np.asarray(lengths, dtype=np.int64).reshape(-1)
    if (lengths < 0).any():
        raise ValueError("negative stream length")
    n = len(lengths)
    offsets = np.zeros(n, np.int64)
    if n:
        np.cumsum(lengths[:-1], out=offsets[1:])
    n_hops = -(-lengths // hop_size)
    hop_offsets = np.zeros(n + 1, np.int64)
    np.cumsum(n_hops, out=hop_offsets[1:])
    n_groups = -(-n_hops // frames_per_group)
    grp = np.zeros(n + 1, np.int64)
    np.cumsum(n_groups, out=grp[1:])
    if grp[-1] > 0x7fffffff:
        raise ValueError("batch too large for one launch")
    packed = np.concatenate([offsets, lengths, hop_offsets,
                             np.frombuffer(_pad8(grp.astype(np.int32)).tobytes(), dtype=np.int64)])
    dev = torch.from_numpy(packed).to(device, non_blocking=False)
    return ProsodyPlan(n, int(lengths.sum()), int(hop_offsets[-1]), int(grp[-1]), lengths, offsets,
                       hop_offsets, grp.astype(np.int32), dev)


def _pad8(a32: np.ndarray) -> np.ndarray:
    if a32.size % 2:
        a32 = np.concatenate([a32, np.zeros(1, np.int32)])
    return a32


class YinDetector:
    """GPU stand-in for the `aubio.pitch('yin', buf, hop, sr)` object held in
    `ProsodyExtractor.pitch_detector` (prosody.py:32-34): stateful sliding buffer, `set_unit`,
    `set_tolerance`, and `detector(chunk)[0]` -> f0 in Hz for one hop."""

    def __init__(self, buf_size: int, hop_size: int, samplerate: int, device=None) -> None:
        self.buf_size, self.hop_size, self.samplerate = buf_size, hop_size, samplerate
        self.engine = ProsodyEngine(device, samplerate, hop_size, buf_size, AUBIO_DEFAULT_TOLERANCE)
        self.state: torch.Tensor | None = None  # None = fresh (zeros)
        self._lock = threading.Lock()

    def set_unit(self, unit: str) -> None:
        if unit not in ("Hz", "default"):
            raise ValueError(f"unit {unit!r} not supported (the reference only uses 'Hz')")

    def set_tolerance(self, tol: float) -> None:
        self.engine.tolerance = float(tol)

    def get_tolerance(self) -> float:
        return self.engine.tolerance

    def set_silence(self, db: float) -> None:
        self.engine.silence_db = float(db)

    def reset(self) -> None:
        with self._lock:
            self.state = None

    def __call__(self, chunk) -> np.ndarray:
        chunk = np.ascontiguousarray(chunk, dtype=np.float32)
        if chunk.shape != (self.hop_size,):
            raise ValueError(f"input size of {chunk.shape} does not match hop_size {self.hop_size}")
        with self._lock:
            _, det, state = self.engine.analyze_host([chunk], state_in=self.state, want_state=True,
                                                     details=True)
            self.state = state
        return det["f0"][:1].copy()


class ProsodyExtractor:
    """Same public surface as the reference's ProsodyExtractor (prosody.py:10-104)."""

    def __init__(self, sample_rate: int = 48000, hop_size: int = 512) -> None:
        self.sample_rate = sample_rate
        self.hop_size = hop_size
        self.pitch_detector = YinDetector(YIN_BUF_SIZE, hop_size, sample_rate)  # prosody.py:32
        self.pitch_detector.set_unit('Hz')                                      # :33
        self.pitch_detector.set_tolerance(YIN_TOLERANCE)                        # :34

    def analyze_buffer(self, audio_buffer) -> dict[str, str]:
        audio_buffer = _coerce(audio_buffer)
        if audio_buffer.ndim != 1:
            audio_buffer = audio_buffer.reshape(-1)
        det = self.pitch_detector
        with det._lock:  # the aubio object is stateful; the CLI sender may call us from two threads
            res, state = det.engine.analyze_host([audio_buffer], state_in=det.state, want_state=True)
            det.state = state
        return res[0]

    def analyze_buffer_details(self, audio_buffer) -> tuple[dict[str, str], dict]:
        """Extension: also returns per-hop f0 / level and the stream RMS / mean f0 (host arrays)."""
        audio_buffer = _coerce(audio_buffer).reshape(-1)
        det = self.pitch_detector
        with det._lock:
            res, d, state = det.engine.analyze_host([audio_buffer], state_in=det.state,
                                                    want_state=True, details=True)
            det.state = state
        return res[0], d

    def analyze_batch(self, buffers) -> list[dict[str, str]]:
        """Extension: many independent utterances (fresh state each) in one launch."""
        return self.pitch_detector.engine.analyze_host([_coerce(b).reshape(-1) for b in buffers])
