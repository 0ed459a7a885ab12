"""`faster_whisper` stand-in (only what /root/reference/backend/services/transcriber.py:8,23-27,53-57,80-84 uses):
`WhisperModel(model_size, device=, compute_type=)` and `.transcribe(audio_or_path, beam_size=, language=)` ->
(iterable of Segment, TranscriptionInfo). The work runs on the B200 (log-mel, encoder, decoder kernels of
libjanus_b200.so) whatever `device` / `compute_type` say: there is no CPU path."""
from __future__ import annotations

import collections

import numpy as np

__version__ = "1.1.0+janus_b200"

Segment = collections.namedtuple("Segment", ["id", "seek", "start", "end", "text", "tokens", "avg_logprob",
                                             "compression_ratio", "no_speech_prob", "words", "temperature"])
TranscriptionInfo = collections.namedtuple("TranscriptionInfo", ["language", "language_probability", "duration",
                                                                 "duration_after_vad", "all_language_probs",
                                                                 "transcription_options", "vad_options"])


class WhisperModel:
    def __init__(self, model_size_or_path: str = "base.en", device: str = "auto", device_index=0,
                 compute_type: str = "default", **_ignored) -> None:
        import os
        from janus_b200.services.transcriber import Transcriber
        self.requested = {"device": device, "compute_type": compute_type}   # honoured as "B200, bf16"
        if os.path.exists(str(model_size_or_path)):
            self._tr = Transcriber(os.path.basename(os.path.normpath(model_size_or_path)), model_path=model_size_or_path)
        else:
            self._tr = Transcriber(model_size_or_path)

    def transcribe(self, audio, language=None, task="transcribe", beam_size=5, condition_on_previous_text=True,
                   no_speech_threshold=0.6, log_prob_threshold=-1.0, **_ignored):
        tr = self._tr
        if task != "transcribe":
            raise ValueError("faster_whisper shim: only task='transcribe' is on the Janus path")
        opts = dict(beam_size=beam_size, condition_on_previous_text=condition_on_previous_text,
                    no_speech_threshold=no_speech_threshold, log_prob_threshold=log_prob_threshold)
        if isinstance(audio, str):
            from janus_b200.services.transcriber import read_wav
            x, decim = read_wav(audio)                                   # PCM WAV only (no PyAV / FFmpeg here)
            duration = len(x) / (16000.0 * decim)
            _, segments, _ = tr._run(x, decim=decim, **opts)
        else:
            audio = np.asarray(audio, dtype=np.float32).reshape(-1)       # 16 kHz mono, as faster-whisper expects
            duration = len(audio) / 16000.0
            _, segments, _ = tr._run(audio, decim=1, **opts)
        segs = [Segment(s["id"], s["seek"], s["start"], s["end"], s["text"], s["tokens"], s["avg_logprob"], None,
                        s["no_speech_prob"], None, s["temperature"]) for s in segments]
        info = TranscriptionInfo(language or "en", 1.0, duration, duration, None, None, None)
        return iter(segs), info
