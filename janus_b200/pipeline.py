"""Batched sender-side semantic encode on one GPU: PCM windows -> (text tokens, prosody tags) -> packets.

This is the multi-stream form of what the reference does per utterance in
/root/reference/backend/services/engine.py:510-550 (process_audio_blocking + transmit_packet_blocking):
STT (`Transcriber.transcribe_buffer`), prosody (`ProsodyExtractor.analyze_buffer`), then
`JanusPacket(...).serialize()`. Streams are independent; one instance serves one GPU and N GPUs run
N instances on disjoint stream shards (no collective on the data path).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from .protocol import JanusMode, pack_packets
from .services.prosody import ENERGY_TAGS, PITCH_TAGS, ProsodyEngine
from .whisper.config import PRESETS, WhisperDims, detokenize, tokenizer_spec
from .whisper.engine import DecodeSession, WhisperEngine
from .whisper.weights import init_random_weights


def shard_streams(n_streams: int, world_size: int, rank: int) -> range:
    """Contiguous, balanced shard of stream ids for `rank` (pure data parallelism)."""
    base, rem = divmod(n_streams, world_size)
    lo = rank * base + min(rank, rem)
    return range(lo, lo + base + (1 if rank < rem else 0))


def split_text_segments(tokens, spec) -> list[str]:
    """Token ids of one window -> segment texts, splitting at timestamp tokens the way faster-whisper
    forms `Segment`s (the reference then strips and joins them: transcriber.py:59-64)."""
    segs, cur = [], []
    for t in tokens:
        t = int(t)
        if t == spec.eot:
            break
        if t >= spec.timestamp_begin:
            if cur:
                segs.append(detokenize(cur, spec))
                cur = []
        else:
            cur.append(t)
    if cur:
        segs.append(detokenize(cur, spec))
    return segs


def join_segments(segs: list[str]) -> str:
    return ' '.join(s.strip() for s in segs).strip()   # transcriber.py:59-64


@dataclass
class EncodeResult:
    tokens: torch.Tensor        # int32 [B, max_new] (device)
    sum_logprob: torch.Tensor   # fp32 [B]
    tags: torch.Tensor          # uint8 [B, 2] (device)


class SemanticEncoder:
    """Fixed-batch engine: B windows of up to 30 s each, one launch sequence per step."""

    def __init__(self, model: str | WhisperDims = "small", batch: int = 64, max_new_tokens: int = 128,
                 device=None, seed: int = 0, weights: dict | None = None, weight_profile: str = "confident",
                 prosody_sample_rate: int = 16000, input_sample_rate: int = 16000, enc_chunk: int = 64) -> None:
        self.dims = PRESETS[model] if isinstance(model, str) else model
        self.spec = tokenizer_spec(self.dims)
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        if weights is None:  # no checkpoints exist offline: random init of the named architecture
            weights = init_random_weights(self.dims, seed=seed, profile=weight_profile)
        self.whisper = WhisperEngine(self.dims, weights, self.device, self.spec)
        self.session = DecodeSession(self.whisper, batch, max_new_tokens)
        self.prosody = ProsodyEngine(self.device, sample_rate=prosody_sample_rate)
        self.batch = batch
        self.decim = input_sample_rate // 16000
        assert input_sample_rate in (16000, 48000)
        self._plans: dict = {}
        d = self.dims
        # the encoder is compute-bound and runs in micro-batches of `enc_chunk` windows (bounded workspace);
        # the decoder is latency/memory-bound and steps all `batch` windows at once so its per-step fixed
        # costs are amortised over as many streams as fit
        self.enc_chunk = min(enc_chunk, batch)
        self.mel_pad = torch.empty((self.enc_chunk, 2 * d.n_audio_ctx + 2, d.n_mels), dtype=torch.bfloat16,
                                   device=self.device)
        self.mel_scratch = torch.empty((self.enc_chunk, self.whisper.frontend.frames_alloc, d.n_mels),
                                       dtype=torch.float32, device=self.device)
        self.enc_out = torch.empty((batch, d.n_audio_ctx, d.n_audio_state), dtype=torch.bfloat16, device=self.device)
        self._pros_out = None

    def _plan(self, lengths: tuple):
        p = self._plans.get(lengths)
        if p is None:
            plan = self.prosody.plan(list(lengths))
            offs = torch.from_numpy(plan.offsets).to(self.device)
            lens = torch.from_numpy(plan.lengths).to(self.device)
            p = (plan, offs, lens, self.prosody.alloc_outputs(plan))
            self._plans[lengths] = p
        return p

    def kernels_per_step(self) -> int:
        n_chunks = -(-self.batch // self.enc_chunk)
        enc = n_chunks * (2 + 7 * self.dims.n_audio_layer + 1 + 2)  # stem, layers, ln_post, 2 logmel kernels
        dec = 1 + self.session.n_steps * self.session.kernels_per_step
        return enc + dec + 2

    def encode_device(self, pcm: torch.Tensor, lengths: tuple) -> EncodeResult:
        """pcm: float32 device tensor, the B windows concatenated (source-rate samples); lengths: per
        window sample counts. Asynchronous on the current stream."""
        assert len(lengths) == self.batch
        plan, offs, lens, pros_out = self._plan(tuple(lengths))
        self.prosody.run(pcm, plan, out=pros_out)
        self.encode_windows(pcm, offs, lens)
        tokens, sum_lp = self.session.decode(self.enc_out)
        return EncodeResult(tokens, sum_lp, pros_out["tags"])

    def encode_windows(self, pcm: torch.Tensor, offs: torch.Tensor, lens: torch.Tensor) -> torch.Tensor:
        """K1 + K2 over all windows in micro-batches; fills self.enc_out [batch, n_ctx, d]."""
        for c0 in range(0, self.batch, self.enc_chunk):
            c1 = min(c0 + self.enc_chunk, self.batch)
            n = c1 - c0
            self.whisper.frontend(pcm, offs[c0:c1], lens[c0:c1], decim=self.decim, scratch=self.mel_scratch[:n],
                                  out_bf16=self.mel_pad[:n])
            self.whisper.encode(self.mel_pad[:n], out=self.enc_out[c0:c1])
        return self.enc_out

    def to_packets(self, tokens_host: np.ndarray, tags_host: np.ndarray, timestamps=None,
                   mode: JanusMode = JanusMode.SEMANTIC_VOICE, override_emotion=None):
        texts = [join_segments(split_text_segments(row, self.spec)) for row in tokens_host]
        pros = [{'energy': ENERGY_TAGS[e], 'pitch': PITCH_TAGS[p]} for e, p in tags_host]
        return texts, pack_packets(texts, pros, mode, override_emotion, timestamps)

    def encode_host(self, buffers: list[np.ndarray], timestamps=None):
        """Host buffers in, packets out (H2D + kernels + D2H + MessagePack). Synchronises."""
        lengths = tuple(len(b) for b in buffers)
        host = torch.empty(sum(lengths), dtype=torch.float32, pin_memory=True)
        hv, o = host.numpy(), 0
        for b in buffers:
            hv[o:o + len(b)] = b
            o += len(b)
        with torch.cuda.device(self.device):
            pcm = host.to(self.device, non_blocking=True)
            res = self.encode_device(pcm, lengths)
            tok = res.tokens.cpu().numpy()
            tags = res.tags.cpu().numpy()
        return self.to_packets(tok, tags, timestamps)


def gather_packets(local_packets: list, local_stream_ids: list[int], dst: int = 0):
    """Host-side gather of the per-rank packet bytes (55-500 B each) onto rank `dst`, ordered by global
    stream id. This is the only cross-rank step of the path and it is not on the GPU data path: an
    NCCL collective over NVLink for a few KB would be pure latency."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return dict(zip(local_stream_ids, local_packets))
    payload = list(zip(local_stream_ids, local_packets))
    gathered = [None] * dist.get_world_size() if dist.get_rank() == dst else None
    dist.gather_object(payload, gathered, dst=dst)
    if dist.get_rank() != dst:
        return None
    out = {}
    for part in gathered:
        out.update(dict(part))
    return dict(sorted(out.items()))
