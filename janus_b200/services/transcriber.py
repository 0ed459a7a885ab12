"""STT service on B200: drop-in for /root/reference/backend/services/transcriber.py (class Transcriber,
:10-91): same constructor (`model_size='base.en'`), `transcribe_buffer(audio_buffer) -> str` with the
same coercions, the same `[::3]` 48 kHz -> 16 kHz decimation (fused into the log-mel kernel's loads),
greedy decode (beam_size=1, language 'en'), and the same `' '.join(seg.text.strip()).strip()` assembly.
`beam_size` is an optional extra constructor argument (default 1 = what the reference passes at :55); a value > 1
decodes with that many beams per window (BASELINE configs[3]) through janus_whisper_decode_beam_step.

The reference loads pretrained weights from the HF hub through faster-whisper; nothing is downloadable
here, so by default the named architecture is randomly initialised from a fixed seed (set
JANUS_B200_WHISPER_SEED / pass `weights=` to change that). Token ids are turned into text by the
shared deterministic table in whisper/config.py because no Whisper vocabulary file exists offline.
"""
from __future__ import annotations

import os
import threading
import wave

import numpy as np
import torch

from ..pipeline import join_segments, split_text_segments
from ..whisper.config import PRESETS, tokenizer_spec
from ..whisper.engine import DecodeSession, WhisperEngine
from ..whisper.weights import init_random_weights

WINDOW_SAMPLES_16K = 480000   # 30 s
MAX_NEW_TOKENS = 448          # faster-whisper: max_length=448 (DecodeSession caps at 448 - len(prompt))


class Transcriber:
    def __init__(self, model_size: str = 'base.en', weights: dict | None = None, device=None,
                 beam_size: int = 1) -> None:
        if not 1 <= beam_size <= 8:
            raise ValueError("beam_size must be in 1 .. 8")
        self.beam_size = beam_size
        if model_size not in PRESETS:
            raise ValueError(f"unknown Whisper model size {model_size!r}; known: {sorted(PRESETS)}")
        self.dims = PRESETS[model_size]
        self.spec = tokenizer_spec(self.dims)
        if weights is None:
            seed = int(os.environ.get("JANUS_B200_WHISPER_SEED", "0"))
            weights = init_random_weights(self.dims, seed=seed, profile="confident")
        self.model = WhisperEngine(self.dims, weights, device, self.spec)
        self._sessions: dict[int, DecodeSession] = {}
        self._lock = threading.Lock()  # called from a 2-worker ThreadPoolExecutor in the reference

    def _session(self, batch: int) -> DecodeSession:
        s = self._sessions.get(batch)
        if s is None:
            s = DecodeSession(self.model, batch, MAX_NEW_TOKENS, beam=self.beam_size if self.beam_size > 1 else 0)
            self._sessions[batch] = s
        return s

    def _transcribe(self, audio: np.ndarray, decim: int) -> str:
        """audio: float32 at 16 kHz * decim. Windows of 30 s are decoded as one batch."""
        if audio.size == 0:
            return ''
        win = WINDOW_SAMPLES_16K * decim
        windows = [audio[i:i + win] for i in range(0, len(audio), win)]
        dev = self.model.device
        with self._lock, torch.cuda.device(dev):
            lens = torch.tensor([len(w) for w in windows], dtype=torch.int64)
            offs = torch.cumsum(lens, 0) - lens
            pcm = torch.from_numpy(np.ascontiguousarray(audio)).to(dev)
            mel_pad, _ = self.model.frontend(pcm, offs.to(dev), lens.to(dev), decim=decim)
            enc = self.model.encode(mel_pad)
            ses = self._session(len(windows))
            ses._ensure_graph()
            ses.cross_kv(enc)
            ses.reset()
            done = 0
            for step in range(ses.n_steps):   # stop as soon as every window has emitted <|endoftext|>
                if ses.graph is not None:
                    ses.graph.replay()
                else:
                    ses._step()
                if step % 8 == 7:
                    done = int(ses.n_done.item())
                    if done >= len(windows):
                        break
            tokens = (ses.finalize()[0] if ses.beam else ses.tokens[:, ses.n_prompt:]).cpu().numpy()
        text_parts = []
        for row in tokens:
            text_parts.extend(split_text_segments(row, self.spec))
        return join_segments(text_parts)

    def transcribe_buffer(self, audio_buffer) -> str:
        if isinstance(audio_buffer, list):                      # transcriber.py:42-43
            audio_buffer = np.concatenate(audio_buffer)
        if not isinstance(audio_buffer, np.ndarray):            # :45-46
            audio_buffer = np.array(audio_buffer, dtype=np.float32)
        if audio_buffer.dtype != np.float32:                    # :48-49
            audio_buffer = audio_buffer.astype(np.float32)
        # :51 `audio_16k = audio_buffer[::3]` -- done inside the log-mel kernel (decim = 3)
        return self._transcribe(audio_buffer.reshape(-1), decim=3)

    def transcribe_file(self, file_path: str) -> str:
        """The reference lets faster-whisper decode any container through PyAV/FFmpeg
        (transcriber.py:66-91); neither exists here, so PCM WAV (8/16/32-bit, any rate that is a
        multiple of 16 kHz... in practice 16 or 48 kHz) is read directly and anything else raises."""
        with wave.open(file_path, 'rb') as wf:
            sr, ch, sw, n = wf.getframerate(), wf.getnchannels(), wf.getsampwidth(), wf.getnframes()
            raw = wf.readframes(n)
        if sw == 2:
            x = np.frombuffer(raw, dtype='<i2').astype(np.float32) / 32768.0
        elif sw == 4:
            x = np.frombuffer(raw, dtype='<i4').astype(np.float32) / 2147483648.0
        elif sw == 1:
            x = (np.frombuffer(raw, dtype=np.uint8).astype(np.float32) - 128.0) / 128.0
        else:
            raise ValueError(f"unsupported WAV sample width {sw}")
        if ch > 1:
            x = x.reshape(-1, ch).mean(axis=1).astype(np.float32)
        if sr % 16000 != 0:
            raise ValueError(f"unsupported WAV sample rate {sr} (need a multiple of 16 kHz)")
        return self._transcribe(x, decim=sr // 16000)
