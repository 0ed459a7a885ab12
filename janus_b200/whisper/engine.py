"""Device-side Whisper engine: weight packing (fp32 dict -> bf16 device tensors + C structs) and the
launch wrappers for the encoder (K2) and decoder (K3) drivers in libjanus_b200.so."""
from __future__ import annotations

import ctypes as C

import torch

from .. import _abi
from .._device import on_device
from .config import TokenizerSpec, WhisperDims, tokenizer_spec
from .frontend import LogMelFrontend


class DimsStruct(C.Structure):
    _fields_ = [("n_mels", C.c_int32), ("n_ctx", C.c_int32), ("d", C.c_int32), ("n_heads", C.c_int32),
                ("n_layers", C.c_int32)]


class EncLayerStruct(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("ln1_w", "ln1_b", "w_qkv", "b_qkv", "w_o", "b_o", "ln2_w", "ln2_b",
                                          "w_fc1", "b_fc1", "w_fc2", "b_fc2")]


class EncWeightsStruct(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("conv1_w", "conv1_b", "conv2_w", "conv2_b", "pos", "ln_post_w",
                                          "ln_post_b")] + [("layers", C.POINTER(EncLayerStruct))]


class WhisperEngine:
    def __init__(self, dims: WhisperDims, weights: dict[str, torch.Tensor], device=None,
                 spec: TokenizerSpec | None = None) -> None:
        _abi.lib()
        if not torch.cuda.is_available():
            raise _abi.JanusNativeError("janus_b200 needs a CUDA device (sm_100); there is no CPU fallback")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        _abi.call("janus_check_device", self.device.index or 0)
        self.dims = dims
        self.spec = spec or tokenizer_spec(dims)
        self._keep: list[torch.Tensor] = []   # owns every device tensor referenced by the C structs
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        with torch.cuda.device(self.device):
            self.frontend = LogMelFrontend(dims.n_mels, self.device)
            self._pack_encoder(weights)
            self._pack_decoder(weights)
        self._enc_ws: torch.Tensor | None = None

    # ---- packing ----------------------------------------------------------------------------
    def _dev(self, t: torch.Tensor, dtype) -> torch.Tensor:
        x = t.detach().to(dtype=dtype).contiguous().to(self.device)
        self._keep.append(x)
        return x

    def _bf16(self, t):
        return self._dev(t, torch.bfloat16)

    def _f32(self, t):
        return self._dev(t, torch.float32)

    def _pack_encoder(self, W) -> None:
        d = self.dims
        self.c_dims = DimsStruct(d.n_mels, d.n_audio_ctx, d.n_audio_state, d.n_audio_head, d.n_audio_layer)
        n = d.n_audio_state
        self.enc_layers = (EncLayerStruct * d.n_audio_layer)()
        for i in range(d.n_audio_layer):
            p = f"enc.l{i}"
            L = self.enc_layers[i]
            wqkv = torch.cat([W[p + ".attn.q.w"], W[p + ".attn.k.w"], W[p + ".attn.v.w"]], 0)
            bqkv = torch.cat([W[p + ".attn.q.b"], torch.zeros(n), W[p + ".attn.v.b"]], 0)
            L.ln1_w, L.ln1_b = self._f32(W[p + ".attn_ln.w"]).data_ptr(), self._f32(W[p + ".attn_ln.b"]).data_ptr()
            L.w_qkv, L.b_qkv = self._bf16(wqkv).data_ptr(), self._f32(bqkv).data_ptr()
            L.w_o, L.b_o = self._bf16(W[p + ".attn.o.w"]).data_ptr(), self._f32(W[p + ".attn.o.b"]).data_ptr()
            L.ln2_w, L.ln2_b = self._f32(W[p + ".mlp_ln.w"]).data_ptr(), self._f32(W[p + ".mlp_ln.b"]).data_ptr()
            L.w_fc1, L.b_fc1 = self._bf16(W[p + ".mlp.fc1.w"]).data_ptr(), self._f32(W[p + ".mlp.fc1.b"]).data_ptr()
            L.w_fc2, L.b_fc2 = self._bf16(W[p + ".mlp.fc2.w"]).data_ptr(), self._f32(W[p + ".mlp.fc2.b"]).data_ptr()
        ew = EncWeightsStruct()
        # conv weights [d, c, 3] -> tap-major [d, 3, c] so a GEMM row is (x[t-1], x[t], x[t+1])
        ew.conv1_w = self._bf16(W["enc.conv1.w"].permute(0, 2, 1).reshape(n, -1)).data_ptr()
        ew.conv1_b = self._f32(W["enc.conv1.b"]).data_ptr()
        ew.conv2_w = self._bf16(W["enc.conv2.w"].permute(0, 2, 1).reshape(n, -1)).data_ptr()
        ew.conv2_b = self._f32(W["enc.conv2.b"]).data_ptr()
        ew.pos = self._f32(W["enc.pos"]).data_ptr()
        ew.ln_post_w, ew.ln_post_b = self._f32(W["enc.ln_post.w"]).data_ptr(), self._f32(W["enc.ln_post.b"]).data_ptr()
        ew.layers = C.cast(self.enc_layers, C.POINTER(EncLayerStruct))
        self.enc_weights = ew

    # ---- encoder ------------------------------------------------------------------------------
    @on_device
    def encode(self, mel_pad: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
        """mel_pad: bf16 [B, 2*n_ctx+2, n_mels] (LogMelFrontend output) -> bf16 [B, n_ctx, d]."""
        d = self.dims
        B = mel_pad.shape[0]
        assert mel_pad.dtype == torch.bfloat16 and mel_pad.is_contiguous() and mel_pad.is_cuda
        assert mel_pad.shape[1] == 2 * d.n_audio_ctx + 2 and mel_pad.shape[2] == d.n_mels
        need = int(_abi.lib().janus_whisper_encode_workspace_bytes(C.byref(self.c_dims), B))
        if self._enc_ws is None or self._enc_ws.numel() < need:
            self._enc_ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        if out is None:
            out = torch.empty((B, d.n_audio_ctx, d.n_audio_state), dtype=torch.bfloat16, device=self.device)
        _abi.call("janus_whisper_encode_bf16", C.byref(self.c_dims), C.byref(self.enc_weights), mel_pad.data_ptr(), B,
                  out.data_ptr(), self._enc_ws.data_ptr(), self._enc_ws.numel(),
                  torch.cuda.current_stream(self.device).cuda_stream)
        return out

    def mel_pad_from_f32(self, mel: torch.Tensor) -> torch.Tensor:
        """[B, n_mels, 2*n_ctx] fp32 (reference layout) -> the encoder's padded time-major bf16 input."""
        B, M, T = mel.shape
        out = torch.zeros((B, T + 2, M), dtype=torch.bfloat16, device=self.device)
        out[:, 1:T + 1] = mel.to(self.device).permute(0, 2, 1).to(torch.bfloat16)
        return out


# ==============================================================================================
# decoder (K3)
# ==============================================================================================
class DecLayerStruct(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("ln1_w", "ln1_b", "w_qkv", "b_qkv", "w_o", "b_o", "lnc_w", "lnc_b",
                                          "w_cq", "b_cq", "w_co", "b_co", "ln2_w", "ln2_b", "w_fc1", "b_fc1",
                                          "w_fc2", "b_fc2", "w_ck_lat", "w_cv", "b_cv")]


class DecWeightsStruct(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("tok_emb", "pos", "ln_w", "ln_b", "w_ckv", "b_ckv", "suppress_mask")] + \
               [("layers", C.POINTER(DecLayerStruct))] + \
               [(n, C.c_int32) for n in ("n_vocab", "n_vocab_pad", "n_text_ctx", "kv_ld", "eot", "blank",
                                         "timestamp_begin", "max_initial_timestamp_index")]


class DecodeStateStruct(C.Structure):
    _fields_ = [("batch", C.c_int32), ("max_len", C.c_int32), ("n_prompt", C.c_int32), ("flags", C.c_int32),
                ("self_ld", C.c_int32), ("reserved0", C.c_int32)] + \
               [(n, C.c_void_p) for n in ("tokens", "forced", "choice", "logprob", "sum_logprob", "pos", "done",
                                          "n_done", "self_k", "self_v", "cross_k", "cross_v", "x", "xn", "qkv",
                                          "att", "q", "ff", "logits", "reserved1", "enc_out", "qlat", "zlat")]


class BeamStateStruct(C.Structure):
    _fields_ = [("beam", C.c_int32), ("n_streams", C.c_int32), ("max_candidates", C.c_int32), ("reserved0", C.c_int32)] + \
               [(n, C.c_void_p) for n in ("cum", "indir", "cand_score", "cand_tok", "fin_tokens", "fin_score", "fin_len",
                                          "n_fin", "stream_done")]


def _pack_decoder(self, W) -> None:
    d, spec = self.dims, self.spec
    n, L = d.n_text_state, d.n_text_layer
    self.c_dec_dims = DimsStruct(d.n_mels, d.n_audio_ctx, n, d.n_text_head, L)
    self.dec_layers = (DecLayerStruct * L)()
    # latent cross-attention (dec_xlat.cu): d <= 768 holds one 128-key tile of encoder states in shared memory,
    # d = 1024 (medium) streams every tile twice (second pass from L2); 16 heads at most (MMA N = 16)
    self.latent_ok = n % 128 == 0 and n <= 1024 and d.n_text_head <= 16 and n == 64 * d.n_text_head and \
        d.n_audio_state == n
    ckv_w, ckv_b = [], []
    for i in range(L):
        p = f"dec.l{i}"
        S = self.dec_layers[i]
        wqkv = torch.cat([W[p + ".attn.q.w"], W[p + ".attn.k.w"], W[p + ".attn.v.w"]], 0)
        bqkv = torch.cat([W[p + ".attn.q.b"], torch.zeros(n), W[p + ".attn.v.b"]], 0)
        S.ln1_w, S.ln1_b = self._f32(W[p + ".attn_ln.w"]).data_ptr(), self._f32(W[p + ".attn_ln.b"]).data_ptr()
        S.w_qkv, S.b_qkv = self._bf16(wqkv).data_ptr(), self._f32(bqkv).data_ptr()
        S.w_o, S.b_o = self._bf16(W[p + ".attn.o.w"]).data_ptr(), self._f32(W[p + ".attn.o.b"]).data_ptr()
        S.lnc_w, S.lnc_b = self._f32(W[p + ".cross_ln.w"]).data_ptr(), self._f32(W[p + ".cross_ln.b"]).data_ptr()
        S.w_cq, S.b_cq = self._bf16(W[p + ".cross.q.w"]).data_ptr(), self._f32(W[p + ".cross.q.b"]).data_ptr()
        S.w_co, S.b_co = self._bf16(W[p + ".cross.o.w"]).data_ptr(), self._f32(W[p + ".cross.o.b"]).data_ptr()
        S.ln2_w, S.ln2_b = self._f32(W[p + ".mlp_ln.w"]).data_ptr(), self._f32(W[p + ".mlp_ln.b"]).data_ptr()
        S.w_fc1, S.b_fc1 = self._bf16(W[p + ".mlp.fc1.w"]).data_ptr(), self._f32(W[p + ".mlp.fc1.b"]).data_ptr()
        S.w_fc2, S.b_fc2 = self._bf16(W[p + ".mlp.fc2.w"]).data_ptr(), self._f32(W[p + ".mlp.fc2.b"]).data_ptr()
        if self.latent_ok:   # JANUS_DECODE_LATENT weights: Wk regrouped per head [H*d, 64], Wv / bv per layer
            H = d.n_text_head
            S.w_ck_lat = self._bf16(W[p + ".cross.k.w"].view(H, 64, n).permute(0, 2, 1).reshape(H * n, 64)
                                    .contiguous()).data_ptr()
            S.w_cv = self._bf16(W[p + ".cross.v.w"]).data_ptr()
            S.b_cv = self._f32(W[p + ".cross.v.b"]).data_ptr()
        ckv_w += [W[p + ".cross.k.w"], W[p + ".cross.v.w"]]
        ckv_b += [torch.zeros(n), W[p + ".cross.v.b"]]
    dw = DecWeightsStruct()
    vpad = -(-d.n_vocab // 256) * 256   # multiple of the widest GEMM N tile (zero rows: logits 0, never picked)
    emb = torch.zeros(vpad, n)
    emb[:d.n_vocab] = W["dec.tok_emb"]
    dw.tok_emb = self._bf16(emb).data_ptr()
    dw.pos = self._f32(W["dec.pos"]).data_ptr()
    dw.ln_w, dw.ln_b = self._f32(W["dec.ln.w"]).data_ptr(), self._f32(W["dec.ln.b"]).data_ptr()
    dw.w_ckv = self._bf16(torch.cat(ckv_w, 0)).data_ptr()
    dw.b_ckv = self._f32(torch.cat(ckv_b, 0)).data_ptr()
    mask = torch.zeros(d.n_vocab, dtype=torch.uint8)
    mask[list(spec.suppress_ids)] = 1
    dw.suppress_mask = self._dev(mask, torch.uint8).data_ptr()
    dw.layers = C.cast(self.dec_layers, C.POINTER(DecLayerStruct))
    dw.n_vocab, dw.n_vocab_pad, dw.n_text_ctx = d.n_vocab, vpad, d.n_text_ctx
    dw.kv_ld = -(-d.n_audio_ctx // 64) * 64
    dw.eot, dw.blank, dw.timestamp_begin = spec.eot, spec.blank, spec.timestamp_begin
    dw.max_initial_timestamp_index = spec.max_initial_timestamp_index
    self.dec_weights = dw
    self.n_vocab_pad, self.kv_ld = vpad, int(dw.kv_ld)


WhisperEngine._pack_decoder = _pack_decoder


class DecodeSession:
    """Greedy (beam = 0) or beam-search (beam >= 1) decode of a fixed-size batch of windows: owns the KV caches and
    the captured CUDA graph of one step. With beam search the decode rows are the beams of the windows
    (row = window * beam + k); cross-attention inputs exist once per window."""

    DECODE_LIGHT = 1   # JANUS_DECODE_LIGHT
    DECODE_LATENT = 4  # JANUS_DECODE_LATENT
    DECODE_KV_SWAPPED = 8  # JANUS_DECODE_KV_SWAPPED
    LATENT_MIN_BATCH = 128   # the latent kernel runs one CTA per window: below ~one window per SM the K/V kernel
                             # (one CTA per window and head) fills the machine better

    def __init__(self, engine: WhisperEngine, batch: int, max_new_tokens: int, use_graph: bool = True,
                 light: bool = False, steps_per_graph: int = 1,
                 latent: bool | None = None, beam: int = 0, patience: float = 1.0,
                 prompt: tuple | list | None = None) -> None:
        if not hasattr(engine, "dec_weights"):
            raise RuntimeError("engine was built without decoder weights")
        if beam and light:
            raise ValueError("beam search excludes light decode sessions")
        if not 0 <= beam <= 8:
            raise ValueError("beam must be in 0 .. 8 (JANUS_MAX_BEAM)")
        self.eng, self.B = engine, batch
        self.beam, self.n_windows = beam, batch
        NW = batch                 # windows: cross-attention inputs, results
        batch = batch * max(beam, 1)   # decode rows
        d, spec, dev = engine.dims, engine.spec, engine.device
        # `prompt`: the whole forced prefix, e.g. [<|startofprev|>, previous text ..., <|startoftranscript|>, ...]
        # (faster-whisper's condition_on_previous_text); default: the sot sequence alone
        prompt = tuple(int(t) for t in (prompt if prompt is not None else spec.sot_sequence))
        if len(prompt) >= spec.max_length:
            raise ValueError("prompt leaves no room for new tokens")
        self.n_prompt = len(prompt)
        self.sot_index = max(i for i, t in enumerate(prompt) if t == spec.sot) if spec.sot in prompt else 0
        self.max_new = min(max_new_tokens, spec.max_length - self.n_prompt)
        self.max_len = self.n_prompt + self.max_new
        n, L, H = d.n_text_state, d.n_text_layer, d.n_text_head
        bf = dict(dtype=torch.bfloat16, device=dev)
        i32 = dict(dtype=torch.int32, device=dev)
        B = batch
        self.tokens = torch.zeros((B, self.max_len), **i32)
        self.forced = torch.full((B, self.max_len), -1, **i32)
        self.choice = torch.zeros((B, self.max_len), **i32)
        self.logprob = torch.zeros((B, self.max_len), dtype=torch.float32, device=dev)
        self.sum_logprob = torch.zeros(B, dtype=torch.float32, device=dev)
        self.pos = torch.zeros(1, **i32)
        self.done = torch.zeros(B, dtype=torch.uint8, device=dev)
        self.n_done = torch.zeros(1, **i32)
        self.self_ld = min(d.n_text_ctx, -(-self.max_len // 8) * 8)   # cache only the positions this session can reach
        self.self_k = torch.zeros((L, B, H, self.self_ld, 64), **bf)
        self.self_v = torch.zeros((L, B, H, self.self_ld, 64), **bf)
        if latent is None:
            # beam search defaults to the K/V-cache form: its kernel serves all beams of a window from ONE pass over
            # K / V (4.6 MB per window and layer), while the latent kernel walks the encoder states once per beam
            latent = engine.latent_ok and not light and NW >= self.LATENT_MIN_BATCH and beam <= 1
        if latent and not engine.latent_ok:
            raise ValueError("latent cross-attention needs d % 128 == 0, d <= 1024 and <= 16 heads")
        if latent and light:
            raise ValueError("latent cross-attention excludes light decode sessions")
        self.latent = latent
        if latent:
            # cross-attention over the encoder states themselves: no per-layer K/V (2 * L times smaller)
            self.cross_k = self.cross_v = None
            self.enc_buf = torch.zeros((NW, d.n_audio_ctx, n), **bf)
            self.qlat = torch.zeros((B, 16, n), **bf)      # rows >= H stay zero
            self.zlat = torch.zeros((B, H, n), **bf)
        elif beam:
            # beam search, K/V form: K row-major / V transposed, the layouts of the tensor-core kernel
            # (dec_cross_attn_beam_mma_kernel) that serves all beams of a window from one pass over K / V
            self.cross_k = torch.zeros((L, NW, H, d.n_audio_ctx, 64), **bf)
            self.cross_v = torch.zeros((L, NW, H, 64, engine.kv_ld), **bf)
        else:
            self.cross_k = torch.zeros((L, NW, H, 64, engine.kv_ld), **bf)
            self.cross_v = torch.zeros((L, NW, H, d.n_audio_ctx, 64), **bf)
        self.x, self.xn = torch.zeros((B, n), **bf), torch.zeros((B, n), **bf)
        self.qkv, self.att = torch.zeros((B, 3 * n), **bf), torch.zeros((B, n), **bf)
        self.q, self.ff = torch.zeros((B, n), **bf), torch.zeros((B, 4 * n), **bf)
        self.logits = torch.zeros((B, engine.n_vocab_pad), dtype=torch.float32, device=dev)
        self.prompt = torch.tensor(list(prompt), **i32)
        s = DecodeStateStruct()
        s.batch, s.max_len, s.n_prompt = B, self.max_len, self.n_prompt
        s.self_ld = self.self_ld
        if light and B > 64:
            raise ValueError("light decode sessions hold at most 64 streams")
        s.flags = (self.DECODE_LIGHT if light else 0) | \
                  (self.DECODE_LATENT if latent else 0) | (self.DECODE_KV_SWAPPED if beam and not latent else 0)
        for name in ("tokens", "forced", "choice", "logprob", "sum_logprob", "pos", "done", "n_done", "self_k",
                     "self_v", "x", "xn", "qkv", "att", "q", "ff", "logits"):
            setattr(s, name, getattr(self, name).data_ptr())
        if latent:
            s.enc_out, s.qlat, s.zlat = self.enc_buf.data_ptr(), self.qlat.data_ptr(), self.zlat.data_ptr()
        else:
            s.cross_k, s.cross_v = self.cross_k.data_ptr(), self.cross_v.data_ptr()
        self.state = s
        self.beam_state = None
        if beam:
            K, MC = beam, max(1, int(round(beam * patience)))
            if MC > 16:
                raise ValueError("beam * patience too large")
            f32 = dict(dtype=torch.float32, device=dev)
            self.max_candidates = MC
            self.cum = torch.zeros(B, **f32)
            self.indir = torch.zeros((B, self.self_ld), dtype=torch.uint8, device=dev)
            self.cand_score = torch.zeros((B, K + 1), **f32)
            self.cand_tok = torch.zeros((B, K + 1), **i32)
            self.fin_tokens = torch.zeros((NW, MC, self.max_len), **i32)
            self.fin_score = torch.zeros((NW, MC), **f32)
            self.fin_len = torch.zeros((NW, MC), **i32)
            self.n_fin = torch.zeros(NW, **i32)
            self.stream_done = torch.zeros(NW, dtype=torch.uint8, device=dev)
            self.best_tokens = torch.zeros((NW, self.max_len), **i32)
            self.best_score = torch.zeros(NW, **f32)
            self.best_len = torch.zeros(NW, **i32)
            self._cum0 = torch.full((NW, K), float("-inf"), **f32)
            self._cum0[:, 0] = 0.0
            self._cum0 = self._cum0.reshape(-1)
            self._indir0 = (torch.arange(B, device=dev) % K).to(torch.uint8)[:, None].expand(B, self.self_ld).contiguous()
            bs = BeamStateStruct()
            bs.beam, bs.n_streams, bs.max_candidates = K, NW, MC
            for name in ("cum", "indir", "cand_score", "cand_tok", "fin_tokens", "fin_score", "fin_len", "n_fin",
                         "stream_done"):
                setattr(bs, name, getattr(self, name).data_ptr())
            self.beam_state = bs
        self.use_graph = use_graph
        self.graph: torch.cuda.CUDAGraph | None = None
        self.n_steps = self.n_prompt + self.max_new - 1
        # several steps per captured graph: fewer host launches per lane when many lanes are in flight
        self.steps_per_graph = max(1, min(steps_per_graph, self.n_steps))
        self.tail_graph: torch.cuda.CUDAGraph | None = None
        # embed, final LN, logits, pick, advance + 11 (13 with latent cross-attention) per layer
        self.kernels_per_step = 3 + (13 if latent else 11) * L + 2 + 1
        if beam:
            self.kernels_per_step += 1   # candidates + update instead of pick

    @on_device
    def _step(self) -> None:
        e = self.eng
        if self.beam_state is not None:
            _abi.call("janus_whisper_decode_beam_step", C.byref(e.c_dec_dims), C.byref(e.dec_weights),
                      C.byref(self.state), C.byref(self.beam_state), torch.cuda.current_stream(e.device).cuda_stream)
            return
        _abi.call("janus_whisper_decode_step", C.byref(e.c_dec_dims), C.byref(e.dec_weights), C.byref(self.state),
                  torch.cuda.current_stream(e.device).cuda_stream)

    @on_device
    def set_prompt(self, prompt) -> None:
        """Replaces the forced prefix by another one of the SAME length (the length is baked into the captured
        step; the tokens are data)."""
        prompt = [int(t) for t in prompt]
        if len(prompt) != self.n_prompt:
            raise ValueError("set_prompt: the prompt length is fixed per session")
        spec = self.eng.spec
        self.prompt.copy_(torch.tensor(prompt, dtype=torch.int32))
        self.sot_index = max(i for i, t in enumerate(prompt) if t == spec.sot) if spec.sot in prompt else 0

    @on_device
    def no_speech_prob(self) -> torch.Tensor:
        """softmax(logits)[<|nospeech|>] per row, fp32 device tensor. Only meaningful right after the step that
        consumed <|startoftranscript|> (step index `sot_index`): that is the distribution faster-whisper /
        CTranslate2 report as `no_speech_prob` (return_no_speech_prob=True)."""
        V = self.eng.dims.n_vocab
        return torch.softmax(self.logits[:, :V], dim=-1)[:, self.eng.spec.no_speech].clone()

    @on_device
    def finalize(self):
        """Beam search: ranks every window's finished / live hypotheses (score / n_tokens) into best_tokens /
        best_score / best_len. Enqueued on the current stream."""
        e = self.eng
        _abi.call("janus_whisper_beam_finalize", C.byref(self.state), C.byref(self.beam_state), e.spec.eot,
                  self.best_tokens.data_ptr(), self.best_score.data_ptr(), self.best_len.data_ptr(),
                  torch.cuda.current_stream(e.device).cuda_stream)
        return self.best_tokens[:, self.n_prompt:], self.best_score

    @on_device
    def reset(self, forced: torch.Tensor | None = None) -> None:
        self.tokens.zero_()
        self.tokens[:, :self.n_prompt] = self.prompt
        self.forced.fill_(-1)
        if forced is not None:  # [B, max_new] tokens to teacher-force after the prompt
            k = min(forced.shape[1], self.max_new)
            self.forced[:, self.n_prompt:self.n_prompt + k] = forced[:, :k].to(self.forced)
        self.pos.zero_()
        self.done.zero_()
        self.n_done.zero_()
        self.sum_logprob.zero_()
        self.logprob.zero_()
        if self.beam_state is not None:
            if forced is not None:
                raise ValueError("teacher forcing is a greedy-session feature")
            self.cum.copy_(self._cum0)
            self.indir.copy_(self._indir0)
            self.n_fin.zero_()
            self.stream_done.zero_()

    def enc_target(self) -> torch.Tensor | None:
        """Latent sessions read the encoder states during the whole decode: the buffer the encoder should write
        this session's windows into (None for K/V sessions, which copy what they need in cross_kv)."""
        return self.enc_buf if self.latent else None

    @on_device
    def cross_kv(self, enc_out: torch.Tensor) -> None:
        e = self.eng
        assert enc_out.shape[0] == self.n_windows and enc_out.dtype == torch.bfloat16 and enc_out.is_contiguous()
        if self.latent:
            if enc_out.data_ptr() != self.enc_buf.data_ptr():
                self.enc_buf.copy_(enc_out)
            return
        _abi.call("janus_whisper_cross_kv_swapped_bf16" if self.beam else "janus_whisper_cross_kv_bf16",
                  C.byref(e.c_dec_dims), C.byref(e.dec_weights), enc_out.data_ptr(), self.n_windows, self.cross_k.data_ptr(), self.cross_v.data_ptr(),
                  torch.cuda.current_stream(e.device).cuda_stream)

    @on_device
    def _ensure_graph(self) -> None:
        if not self.use_graph or self.graph is not None:
            return
        self.reset()
        self._step()  # eager warm-up: sets kernel attributes outside capture
        torch.cuda.current_stream(self.eng.device).synchronize()
        self.graph = self.capture(self.steps_per_graph)
        rem = self.n_steps % self.steps_per_graph
        if rem:
            self.tail_graph = self.capture(rem)

    @on_device
    def capture(self, n_steps: int) -> torch.cuda.CUDAGraph:
        """A CUDA graph of `n_steps` decode steps (the step reads its position from device memory, so graphs of
        different lengths can be mixed freely). Call after one eager step (kernel attributes are set outside capture)."""
        g = torch.cuda.CUDAGraph()
        # an explicit capture stream ON THIS ENGINE'S DEVICE: torch.cuda.graph's default capture stream is a class-level
        # singleton created on whichever device was current first (a second engine on cuda:1 captured an empty graph)
        if getattr(self, "_cap_stream", None) is None:
            self._cap_stream = torch.cuda.Stream(self.eng.device)
        # thread_local: another service thread (the reference runs STT and prosody from a 2-worker pool) may call
        # cudaMalloc / pin memory while this thread captures
        with torch.cuda.graph(g, stream=self._cap_stream, capture_error_mode="thread_local"):
            for _ in range(n_steps):
                self._step()
        return g

    @on_device
    def run_steps(self, n_steps: int | None = None) -> None:
        """Enqueues n_steps decode steps (default: all) on the current stream."""
        n = self.n_steps if n_steps is None else n_steps
        if self.graph is None:
            for _ in range(n):
                self._step()
            return
        full, rem = divmod(n, self.steps_per_graph)
        for _ in range(full):
            self.graph.replay()
        if rem:
            if rem == self.n_steps % self.steps_per_graph and self.tail_graph is not None:
                self.tail_graph.replay()
            else:
                for _ in range(rem):
                    self._step()

    @on_device
    def decode(self, enc_out: torch.Tensor, forced: torch.Tensor | None = None, n_steps: int | None = None):
        """Returns (tokens [n_windows, max_new] int32 device, sum_logprob [n_windows]); with beam search those of the
        best hypothesis of every window. Does not synchronise."""
        self._ensure_graph()
        self.cross_kv(enc_out)
        self.reset(forced)
        self.run_steps(n_steps)
        if self.beam_state is not None:
            return self.finalize()
        return self.tokens[:, self.n_prompt:], self.sum_logprob
