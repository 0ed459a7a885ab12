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

from ._device import on_device
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
    """Device-side results of one `encode_device` call. The work is still in flight on the lane
    streams when this is returned: call `wait()` (stream-side join) or `synchronize()` before reading.
    The tensors are slots of a small ring and stay valid for `SemanticEncoder.result_slots - 1`
    further calls."""
    tokens: torch.Tensor        # int32 [B, max_new] (device)
    sum_logprob: torch.Tensor   # fp32 [B]
    tags: torch.Tensor          # uint8 [B, 2] (device)
    events: list                # one per lane: recorded after the lane wrote its rows
    mean_f0: torch.Tensor | None = None   # fp32 [B]: mean f0 of the voiced hops in Hz (0 = none voiced)
    rms: torch.Tensor | None = None       # fp32 [B]: stream RMS (prosody.py:67)

    def wait(self, stream: torch.cuda.Stream | None = None) -> "EncodeResult":
        st = stream if stream is not None else torch.cuda.current_stream(self.tokens.device)
        for ev in self.events:
            st.wait_event(ev)
        return self

    def synchronize(self) -> "EncodeResult":
        for ev in self.events:
            ev.synchronize()
        return self


class _Lane:
    """One decode lane: `lane_batch` windows with their own KV caches, CUDA graph and stream."""

    def __init__(self, enc: "SemanticEncoder", index: int) -> None:
        dev = enc.device
        self.index = index
        self.stream = torch.cuda.Stream(dev)
        with torch.cuda.stream(self.stream):
            self.session = DecodeSession(enc.whisper, enc.lane_batch, enc.max_new_tokens, light=enc.light,
                                         steps_per_graph=enc.steps_per_graph, latent=enc.latent,
                                         beam=enc.beam)
            self.session._ensure_graph()
        self.stream.synchronize()
        self.ready = torch.cuda.Event()    # encoder stream -> lane: cross-KV of this lane is written
        self.free = torch.cuda.Event()     # lane -> encoder stream: the lane finished decoding
        self.free.record(self.stream)
        self.pros_out: dict = {}


class SemanticEncoder:
    """Batched engine for one GPU. A call encodes `batch` windows (each up to 30 s) as
    `batch / lane_batch` *lanes*:

      encoder stream   for each lane in turn: YIN/RMS prosody, log-mel, Whisper encoder (tensor-pipe
                       bound, tcgen05) and the cross-K/V projection of that lane's windows
      lane streams     the lane's greedy decode (HBM-bound: 27 MB of encoder states per window and step with
                       latent cross-attention -- lanes of >= 128 windows -- else 55 MB of cross-K/V),
                       replayed from a CUDA graph, then a copy of its tokens/tags into the result slot

    The decode kernels of a lane use no TMEM and < 32 KB of shared memory (JANUS_DECODE_LIGHT), so they
    share the SMs with the persistent tcgen05 GEMM / attention CTAs of the encoder stream: the HBM-bound
    decode of earlier lanes overlaps the tensor-bound encoder of later lanes -- also across calls, since
    there are `n_lanes` >= 2 * batch / lane_batch lanes used round-robin and a call only enqueues work.
    """

    def __init__(self, model: str | WhisperDims = "small", batch: int = 64, max_new_tokens: int = 128,
                 device=None, seed: int = 0, weights: dict | None = None, weight_profile: str = "confident",
                 prosody_sample_rate: int = 16000, input_sample_rate: int = 16000, lane_batch: int = 64,
                 n_lanes: int | None = None, steps_per_graph: int = 10, light: bool | None = None,
                 enc_chunk: int = 64, latent: bool | None = None, beam: int = 0) -> None:
        self.dims = PRESETS[model] if isinstance(model, str) else model
        self.spec = tokenizer_spec(self.dims)
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        if weights is None:  # no checkpoints exist offline: random init of the named architecture
            weights = init_random_weights(self.dims, seed=seed, profile=weight_profile)
        self.batch = batch
        self.lane_batch = min(lane_batch, batch)
        if batch % self.lane_batch:
            raise ValueError(f"batch {batch} must be a multiple of lane_batch {self.lane_batch}")
        self.lanes_per_call = batch // self.lane_batch
        self.n_lanes = n_lanes if n_lanes is not None else 2 * self.lanes_per_call
        if self.n_lanes < self.lanes_per_call:
            raise ValueError("n_lanes must be >= batch / lane_batch")
        # 0: greedy (what the reference calls: beam_size=1, transcriber.py:55); >= 1: beam search with that many beams
        # per window (BASELINE configs[3] uses 5), decode rows = lane_batch * beam
        self.beam = beam
        # None: decode sessions pick latent cross-attention (over the encoder states, half the HBM bytes per
        # step) whenever the model fits and a lane holds >= DecodeSession.LATENT_MIN_BATCH windows
        self.latent = latent
        self.light = (self.lane_batch <= 64 and not latent and not beam) if light is None else light
        self.max_new_tokens, self.steps_per_graph = max_new_tokens, steps_per_graph
        self.decim = input_sample_rate // 16000
        assert input_sample_rate in (16000, 48000)
        d = self.dims
        with torch.cuda.device(self.device):
            self.whisper = WhisperEngine(self.dims, weights, self.device, self.spec)
            self.prosody = ProsodyEngine(self.device, sample_rate=prosody_sample_rate)
            # high priority: when an SM frees up, the (few, large, persistent) tcgen05 CTAs of the encoder are
            # placed before the many small decode CTAs, which then fill the remaining threads / shared memory
            self.enc_stream = torch.cuda.Stream(self.device, priority=-1)
            self.lanes = [_Lane(self, i) for i in range(self.n_lanes)]
            self._next_lane = 0
            self._last_decode = None
            lb = self.lane_batch
            # the encoder is tensor-bound and runs in micro-batches of `enc_chunk` windows (bounded workspace);
            # encoder-stream buffers are shared by all lanes (the encoder stream is in-order)
            self.enc_chunk = ec = min(enc_chunk, lb)
            self.mel_pad = torch.empty((ec, 2 * d.n_audio_ctx + 2, d.n_mels), dtype=torch.bfloat16, device=self.device)
            self.mel_scratch = torch.empty((ec, self.whisper.frontend.frames_alloc, d.n_mels), dtype=torch.float32,
                                           device=self.device)
            # latent decode sessions own their encoder-state buffer (they read it during the whole decode)
            self.enc_out = None if self.lanes[0].session.latent else \
                torch.empty((lb, d.n_audio_ctx, d.n_audio_state), dtype=torch.bfloat16, device=self.device)
            self.max_new = self.lanes[0].session.max_new
            self.result_slots = -(-self.n_lanes // self.lanes_per_call) + 1
            self._results = [(torch.zeros((batch, self.max_new), dtype=torch.int32, device=self.device),
                              torch.zeros(batch, dtype=torch.float32, device=self.device),
                              torch.zeros((batch, 2), dtype=torch.uint8, device=self.device),
                              torch.zeros(batch, dtype=torch.float32, device=self.device),
                              torch.zeros(batch, dtype=torch.float32, device=self.device))
                             for _ in range(self.result_slots)]
            self._slot = 0
        self._plans: dict = {}
        self.session = self.lanes[0].session   # (single-lane callers / diagnostics)

    # ---- planning -------------------------------------------------------------------------
    def _lane_plan(self, lengths: tuple):
        """Index arrays for one lane's windows (offsets relative to the lane's first sample)."""
        p = self._plans.get(lengths)
        if p is None:
            plan = self.prosody.plan(list(lengths))
            offs = torch.from_numpy(plan.offsets).to(self.device)
            lens = torch.from_numpy(plan.lengths).to(self.device)
            p = (plan, offs, lens)
            self._plans[lengths] = p
        return p

    def kernels_per_step(self) -> int:
        chunks = -(-self.lane_batch // self.enc_chunk)
        # prosody (2), cross-KV (1) per lane; log-mel (2), stem (2), layers, ln_post per encoder micro-batch
        enc = self.lanes_per_call * (3 + chunks * (2 + 2 + 7 * self.dims.n_audio_layer + 1))
        ses = self.lanes[0].session
        return enc + self.lanes_per_call * ses.n_steps * ses.kernels_per_step

    # ---- the path -------------------------------------------------------------------------
    @on_device
    def encode_lane_front(self, lane: _Lane, pcm: torch.Tensor, lengths: tuple) -> None:
        """Prosody + K1 + K2 + cross-K/V of one lane's windows; call under the encoder stream."""
        plan, offs, lens = self._lane_plan(lengths)
        out = lane.pros_out.get(lengths)
        if out is None:
            out = lane.pros_out[lengths] = self.prosody.alloc_outputs(plan)
        lane.cur_pros = out
        self.prosody.run(pcm, plan, out=out)
        enc = self.encode_lane_windows(pcm, offs, lens, out=lane.session.enc_target())
        lane.session.cross_kv(enc)

    @on_device
    def encode_lane_windows(self, pcm: torch.Tensor, offs: torch.Tensor, lens: torch.Tensor,
                            out: torch.Tensor | None = None) -> torch.Tensor:
        """K1 + K2 over one lane's windows in encoder micro-batches; fills `out` (a latent decode session's own
        encoder-state buffer) or the shared self.enc_out [lane_batch, n_ctx, d]."""
        enc_out = self.enc_out if out is None else out
        for c0 in range(0, self.lane_batch, self.enc_chunk):
            n = min(self.enc_chunk, self.lane_batch - c0)
            self.whisper.frontend(pcm, offs[c0:c0 + n], lens[c0:c0 + n], decim=self.decim, scratch=self.mel_scratch[:n],
                                  out_bf16=self.mel_pad[:n])
            self.whisper.encode(self.mel_pad[:n], out=enc_out[c0:c0 + n])
        return enc_out

    @on_device
    def encode_device(self, pcm: torch.Tensor, lengths, host_pcm: torch.Tensor | None = None) -> EncodeResult:
        """pcm: float32 device tensor holding the `batch` windows concatenated (source-rate samples);
        lengths: per-window sample counts. If `host_pcm` (pinned) is given, each lane's slice is first
        copied host->device into `pcm` on the encoder stream. Only enqueues work; see EncodeResult."""
        lengths = tuple(int(v) for v in lengths)
        assert len(lengths) == self.batch
        lb = self.lane_batch
        dev = self.device
        tokens, sum_lp, tags, mean_f0, rms = self._results[self._slot]
        self._slot = (self._slot + 1) % self.result_slots
        cur = torch.cuda.current_stream(dev)
        ev_in = torch.cuda.Event()
        ev_in.record(cur)
        self.enc_stream.wait_event(ev_in)
        events = []
        start = 0
        for j in range(self.lanes_per_call):
            lane = self.lanes[self._next_lane]
            self._next_lane = (self._next_lane + 1) % self.n_lanes
            ll = lengths[j * lb:(j + 1) * lb]
            n = sum(ll)
            sl = pcm[start:start + n]
            with torch.cuda.stream(self.enc_stream):
                self.enc_stream.wait_event(lane.free)
                if host_pcm is not None:
                    sl.copy_(host_pcm[start:start + n], non_blocking=True)
                self.encode_lane_front(lane, sl, ll)
                lane.ready.record(self.enc_stream)
            with torch.cuda.stream(lane.stream):
                lane.stream.wait_event(lane.ready)
                ses = lane.session
                ses.reset()
                ses.run_steps()
                r = slice(j * lb, (j + 1) * lb)
                if ses.beam:
                    best_tok, best_lp = ses.finalize()
                    tokens[r].copy_(best_tok, non_blocking=True)
                    sum_lp[r].copy_(best_lp, non_blocking=True)
                else:
                    tokens[r].copy_(ses.tokens[:, ses.n_prompt:], non_blocking=True)
                    sum_lp[r].copy_(ses.sum_logprob, non_blocking=True)
                tags[r].copy_(lane.cur_pros["tags"], non_blocking=True)
                mean_f0[r].copy_(lane.cur_pros["mean_f0"], non_blocking=True)
                rms[r].copy_(lane.cur_pros["rms"], non_blocking=True)
                lane.free = torch.cuda.Event()
                lane.free.record(lane.stream)
                self._last_decode = lane.free
                events.append(lane.free)
            start += n
        return EncodeResult(tokens, sum_lp, tags, events, mean_f0, rms)

    @on_device
    def synchronize(self) -> None:
        self.enc_stream.synchronize()
        for lane in self.lanes:
            lane.stream.synchronize()

    def to_packets(self, tokens_host: np.ndarray, tags_host: np.ndarray, timestamps=None,
                   mode: JanusMode = JanusMode.SEMANTIC_VOICE, override_emotion=None,
                   avg_pitch_hz: np.ndarray | None = None, avg_energy: np.ndarray | None = None):
        """Token ids + tag ids -> texts and MessagePack packets. `avg_pitch_hz` / `avg_energy` (per stream)
        opt in to the richer prosody the reference's receiver already reads but its sender never fills
        (/root/reference/backend/services/engine.py:231-233,554-560): off by default, because
        the extra keys change the packet bytes."""
        texts = [join_segments(split_text_segments(row, self.spec)) for row in tokens_host]
        pros = [{'energy': ENERGY_TAGS[e], 'pitch': PITCH_TAGS[p]} for e, p in tags_host]
        if avg_pitch_hz is not None or avg_energy is not None:
            for i, d_ in enumerate(pros):
                if avg_pitch_hz is not None:
                    d_['avg_pitch_hz'] = float(avg_pitch_hz[i])
                if avg_energy is not None:
                    d_['avg_energy'] = float(avg_energy[i])
        return texts, pack_packets(texts, pros, mode, override_emotion, timestamps)

    def encode_host(self, buffers: list[np.ndarray], timestamps=None, rich_prosody: bool = False):
        """Host buffers in, packets out (H2D + kernels + D2H + MessagePack). Synchronises. `rich_prosody` adds the
        stream's mean voiced f0 / RMS (already on the device: prosody_reduce_kernel) as `avg_pitch_hz` / `avg_energy`
        to the packet's prosody dict -- the fields the reference's receiver reads but its sender leaves None
        (/root/reference/backend/services/engine.py:231-233,554-560); off by default: it changes the bytes."""
        assert len(buffers) == self.batch
        lengths = tuple(len(b) for b in buffers)
        host = torch.empty(max(sum(lengths), 1), dtype=torch.float32, pin_memory=True)
        hv, o = host.numpy(), 0
        for b in buffers:
            hv[o:o + len(b)] = b
            o += len(b)
        with torch.cuda.device(self.device):
            pcm = torch.empty(host.numel(), dtype=torch.float32, device=self.device)
            res = self.encode_device(pcm, lengths, host_pcm=host).wait()
            tok = res.tokens.cpu().numpy()
            tags = res.tags.cpu().numpy()
            if rich_prosody:
                return self.to_packets(tok, tags, timestamps, avg_pitch_hz=res.mean_f0.cpu().numpy(),
                                       avg_energy=res.rms.cpu().numpy())
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
