"""STT service on B200: drop-in for /root/reference/backend/services/transcriber.py (class Transcriber,
:10-91): same constructor (`model_size='base.en'`), `transcribe_buffer(audio_buffer) -> str` with the
same coercions, the same `[::3]` 48 kHz -> 16 kHz decimation (fused into the log-mel kernel's loads),
greedy decode (beam_size=1, language 'en'), and the same `' '.join(seg.text.strip()).strip()` assembly.

Around the model, the utterance-level loop of `WhisperModel.transcribe` (faster-whisper, RECALLED: SURVEY.md 3.5;
restated for the tests in oracle/transcribe_ref.py) is reproduced with its defaults, which are what the reference
gets because it passes none of these arguments (transcriber.py:53-57):
  * features are computed ONCE per utterance (one global max-8 clamp, continuous STFT) and 3000-frame windows are
    sliced at `seek`, zero-padded in feature space;
  * `seek` advances by the last timestamp the model emitted (a full window if the tokens end in a single timestamp or
    hold no timestamp pair);
  * `condition_on_previous_text=True`: windows after the first are prompted with <|startofprev|> + the last 223
    tokens of the text so far;
  * the no-speech gate: a window with no_speech_prob > 0.6 whose avg_logprob is not > -1.0 is skipped, so silence
    returns '' and the caller sends no packet (/root/reference/backend/services/engine.py:536).
Not reproduced: the temperature-fallback loop (re-decoding with SAMPLING at t = 0.2 ... 1.0 when avg_logprob < -1 or
the compression ratio > 2.4): the t = 0 result is always accepted.

`beam_size` is an optional extra constructor argument (default 1 = what the reference passes at :55); a value > 1
decodes with that many beams per window (BASELINE configs[3]) through janus_whisper_decode_beam_step.

Weights: the reference's `WhisperModel(model_size)` resolves the name to a pretrained checkpoint in the Hugging Face
cache. Here `model_path=` (or $JANUS_B200_WHISPER_PATH, or a checkpoint found in the local caches for that name) is
loaded through `whisper.weights.load_weights`, with the tokenizer.json next to it (`tokenizer=`) turning ids into
text. When nothing is found -- there is no network here -- the named architecture is randomly initialised from a
fixed seed and a warning is logged; ids then map to the shared synthetic table of whisper/config.py.
"""
from __future__ import annotations

import glob
import logging
import os
import threading
import wave

import numpy as np
import torch

from ..pipeline import join_segments
from ..whisper.config import PRESETS, detokenize, tokenizer_spec
from ..whisper.engine import DecodeSession, WhisperEngine
from ..whisper.frontend import N_FRAMES
from ..whisper.weights import init_random_weights, load_weights

logger = logging.getLogger(__name__)

STEP_BLOCK = 8                # decode steps per CUDA-graph replay / between two end-of-text checks
MAX_NEW_TOKENS = 448          # faster-whisper: max_length=448 (DecodeSession caps at 448 - len(prompt))
_CAPTURE_LOCK = threading.Lock()   # CUDA-graph capture of any Transcriber in this process is serialised


def find_pretrained(model_size: str) -> str | None:
    """A local checkpoint for `model_size`, in the places the reference's stack would have put one:
    $JANUS_B200_WHISPER_PATH, the HF cache (faster-whisper's CTranslate2 conversions `Systran/faster-whisper-<size>`,
    then `openai/whisper-<size>`), openai-whisper's ~/.cache/whisper/<size>.pt. None if there is none."""
    env = os.environ.get("JANUS_B200_WHISPER_PATH")
    if env:
        return env
    hf = os.environ.get("HF_HOME", os.path.join(os.path.expanduser("~"), ".cache", "huggingface"))
    for repo in (f"models--Systran--faster-whisper-{model_size}", f"models--openai--whisper-{model_size}"):
        for snap in sorted(glob.glob(os.path.join(hf, "hub", repo, "snapshots", "*"))):
            if any(os.path.exists(os.path.join(snap, f)) for f in ("model.bin", "model.safetensors", "pytorch_model.bin")):
                return snap
    pt = os.path.join(os.path.expanduser("~"), ".cache", "whisper", f"{model_size}.pt")
    return pt if os.path.exists(pt) else None


class Transcriber:
    def __init__(self, model_size: str = 'base.en', weights: dict | None = None, device=None,
                 beam_size: int = 1, model_path: str | None = None, tokenizer: str | None = None) -> None:
        if not 1 <= beam_size <= 8:
            raise ValueError("beam_size must be in 1 .. 8")
        self.beam_size = beam_size
        # WhisperModel.transcribe() defaults (the reference overrides none of them)
        self.condition_on_previous_text = True
        self.no_speech_threshold = 0.6
        self.log_prob_threshold = -1.0
        self.max_windows: int | None = None   # safety valve for tests with random weights (faster-whisper has none)
        self.pretrained = False
        path = model_path if model_path is not None else (find_pretrained(model_size) if weights is None else None)
        if weights is None and path is not None:
            self.dims, weights = load_weights(path, name=model_size)
            self.pretrained = True
            if tokenizer is None:
                cand = os.path.join(path if os.path.isdir(path) else os.path.dirname(path), "tokenizer.json")
                tokenizer = cand if os.path.exists(cand) else None
        else:
            if model_size not in PRESETS:
                raise ValueError(f"unknown Whisper model size {model_size!r}; known: {sorted(PRESETS)}")
            self.dims = PRESETS[model_size]
        if tokenizer is not None:
            from ..whisper.tokenizer import WhisperTokenizer
            self.spec = WhisperTokenizer.from_file(tokenizer, self.dims).spec()
        else:
            self.spec = tokenizer_spec(self.dims)
        if weights is None:
            seed = int(os.environ.get("JANUS_B200_WHISPER_SEED", "0"))
            logger.warning("no pretrained Whisper checkpoint for %r found locally (set JANUS_B200_WHISPER_PATH or pass "
                           "model_path=): using seeded random-init weights, text is synthetic", model_size)
            weights = init_random_weights(self.dims, seed=seed, profile="confident")
        self.model = WhisperEngine(self.dims, weights, device, self.spec)
        self._sessions: dict[tuple, DecodeSession] = {}
        self._lock = threading.Lock()  # called from a 2-worker ThreadPoolExecutor in the reference
        self.last_trace: list = []     # per window of the last call: seek, tokens, no_speech_prob, avg_logprob, skipped
        self.last_segments: list = []  # per emitted segment of the last call: faster-whisper's Segment fields
        # capture the single-window step graph now, not inside the first transcribe_buffer call on a worker thread
        with self._lock, torch.cuda.device(self.model.device):
            self._session(len(self.spec.sot_sequence), self.beam_size)

    def _session(self, n_prompt: int, beam_size: int) -> DecodeSession:
        """One single-window session per prompt LENGTH (the length is baked into the step; the tokens are data).
        The bare sot prompt replays a CUDA graph; conditioned prompts (windows after the first of a > 30 s utterance)
        run the same kernels eagerly."""
        key = (n_prompt, beam_size)
        s = self._sessions.get(key)
        if s is None:
            n_sot = len(self.spec.sot_sequence)
            base = n_prompt == n_sot
            if len(self._sessions) >= 8 and not base:   # bound the cache: conditioned prompt lengths vary
                for k in [k for k in self._sessions if k[0] != n_sot][:1]:
                    del self._sessions[k]
            prompt = [self.spec.sot_prev] * (n_prompt - n_sot) + list(self.spec.sot_sequence)
            s = DecodeSession(self.model, 1, MAX_NEW_TOKENS, beam=beam_size if beam_size > 1 else 0,
                              use_graph=base, prompt=prompt)
            if base:
                with _CAPTURE_LOCK:
                    s._ensure_graph()
                    # single steps until <|startoftranscript|> has been consumed (no_speech_prob is read there), then
                    # STEP_BLOCK steps per replay: an utterance is launch-latency bound, fewer host launches help
                    s.block_graph = s.capture(STEP_BLOCK) if not s.beam else None
            self._sessions[key] = s
        return s

    # ---- one window ---------------------------------------------------------------------------
    def _decode_window(self, enc: torch.Tensor, prompt: list[int], beam_size: int):
        """-> (tokens before <|endoftext|>, sum_logprob, no_speech_prob)."""
        ses = self._session(len(prompt), beam_size)
        ses.set_prompt(prompt)
        ses.cross_kv(enc)
        ses.reset()
        nsp = None
        block = getattr(ses, "block_graph", None)
        step = 0
        while step < ses.n_steps:   # stop as soon as the window has emitted <|endoftext|>
            if block is not None and step > ses.sot_index and ses.n_steps - step >= STEP_BLOCK:
                block.replay()
                step += STEP_BLOCK
                check = True
            else:
                if ses.graph is not None:
                    ses.graph.replay()
                else:
                    ses._step()
                if step == ses.sot_index:
                    nsp = ses.no_speech_prob()
                step += 1
                check = step % STEP_BLOCK == 0
            if check and step > ses.n_prompt and int(ses.n_done.item()) >= 1:
                break
        if ses.beam:
            tok, lp = ses.finalize()
            row, sum_lp = tok[0].cpu().tolist(), float(lp[0])
        else:
            row, sum_lp = ses.tokens[0, ses.n_prompt:].cpu().tolist(), float(ses.sum_logprob[0])
        eot = self.spec.eot
        tokens = row[:row.index(eot)] if eot in row else row
        return tokens, sum_lp, float(nsp[0]) if nsp is not None else 0.0

    # ---- the utterance loop -------------------------------------------------------------------
    def _transcribe(self, audio: np.ndarray, decim: int, **overrides) -> str:
        return self._run(audio, decim, **overrides)[0]

    def _run(self, audio: np.ndarray, decim: int, beam_size: int | None = None,
             condition_on_previous_text: bool | None = None, no_speech_threshold: float | None = None,
             log_prob_threshold: float | None = None):
        """audio: float32 at 16 kHz * decim (the kernel reads audio[::decim]) -> (text, segments, per-window trace).
        The keyword arguments override the object's settings for this call only (the faster_whisper shim passes
        transcribe()'s arguments through them); the returned lists belong to this call (race-free under the
        reference's 2-worker pool), `last_trace` / `last_segments` are diagnostics of the most recent call."""
        beam_size = self.beam_size if beam_size is None else beam_size
        nst = self.no_speech_threshold if no_speech_threshold is None else no_speech_threshold
        lpt = self.log_prob_threshold if log_prob_threshold is None else log_prob_threshold
        if not 1 <= beam_size <= 8:
            raise ValueError("beam_size must be in 1 .. 8")
        cond = self.condition_on_previous_text if condition_on_previous_text is None else condition_on_previous_text
        if audio.size == 0:
            self.last_trace, self.last_segments = [], []
            return '', [], []
        spec, fe, dev = self.spec, self.model.frontend, self.model.device
        tb, eot = spec.timestamp_begin, spec.eot
        n_frames = fe.n_frames(len(audio), decim)
        content_frames = n_frames - 1
        texts: list[str] = []
        trace: list = []
        segments: list = []
        with self._lock, torch.cuda.device(dev):
            pcm = torch.from_numpy(np.ascontiguousarray(audio)).to(dev)
            offs = torch.zeros(1, dtype=torch.int64, device=dev)
            lens = torch.tensor([len(audio)], dtype=torch.int64, device=dev)
            scratch, maxbuf = fe.utterance(pcm, offs, lens, max(n_frames, 1), decim)
            seek, all_tokens = 0, []
            while seek < content_frames and (self.max_windows is None or len(trace) < self.max_windows):
                segment_size = min(N_FRAMES, content_frames - seek)
                mel_pad, _ = fe.windows(scratch, maxbuf, lens, [0], [seek], decim)
                enc = self.model.encode(mel_pad)
                prompt = list(spec.sot_sequence)
                if cond and all_tokens:
                    prompt = [spec.sot_prev] + all_tokens[-(spec.max_length // 2 - 1):] + prompt
                tokens, sum_lp, nsp = self._decode_window(enc, prompt, beam_size)
                avg_logprob = sum_lp / (len(tokens) + 1)
                skip = nsp > nst and not avg_logprob > lpt
                trace.append({"seek": seek, "tokens": tokens, "no_speech_prob": nsp, "avg_logprob": avg_logprob,
                              "skipped": skip, "n_prompt": len(prompt)})
                if skip:
                    seek += segment_size
                    continue
                segs, advance = split_window_segments(tokens, tb, segment_size)
                t0 = seek * 0.01                      # time_offset: 10 ms per feature frame
                for start, end, sl in segs:
                    text = detokenize([t for t in sl if t < eot], spec)
                    if start == end or not text.strip():
                        continue
                    texts.append(text)
                    segments.append({"id": len(segments) + 1, "seek": seek, "start": t0 + 0.02 * start,
                                     "end": t0 + 0.02 * end, "text": text, "tokens": sl, "avg_logprob": avg_logprob,
                                     "no_speech_prob": nsp, "temperature": 0.0})
                seek += advance
                all_tokens.extend(tokens)
        self.last_trace, self.last_segments = trace, segments
        return join_segments(texts), segments, trace

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
        x, decim = read_wav(file_path)
        return self._transcribe(x, decim=decim)


def read_wav(file_path: str):
    """PCM WAV -> (mono float32 samples, decimation factor to 16 kHz)."""
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
    return x, sr // 16000


def split_window_segments(tokens: list[int], timestamp_begin: int, segment_size: int):
    """One window's tokens (before <|endoftext|>) -> ([(start_pos, end_pos, tokens)], frames to advance `seek` by),
    the segment / seek rule of faster-whisper's generate_segments (timestamp positions are 0.02 s = 2 feature
    frames): split after every CONSECUTIVE timestamp pair; with at least one pair the window ends at the last pair
    unless the tokens end in a single timestamp; with none the whole list is one segment and the window is consumed."""
    tb = timestamp_begin
    is_ts = [t >= tb for t in tokens]
    single_ending = len(tokens) >= 2 and not is_ts[-2] and is_ts[-1]
    cuts = [i for i in range(1, len(tokens)) if is_ts[i] and is_ts[i - 1]]
    segs = []
    if cuts:
        if single_ending:
            cuts = cuts + [len(tokens)]
        last = 0
        for cur in cuts:
            sl = tokens[last:cur]
            segs.append((sl[0] - tb, sl[-1] - tb, sl))
            last = cur
        return segs, (segment_size if single_ending else (tokens[last - 1] - tb) * 2)
    stamps = [t for t in tokens if t >= tb]
    end = stamps[-1] - tb if stamps and stamps[-1] != tb else segment_size // 2
    return [(0, end, tokens)], segment_size
