"""Device-side Whisper engine: weight packing (fp32 dict -> bf16 device tensors + C structs) and the
launch wrappers for the encoder (K2) and decoder (K3) drivers in libjanus_b200.so."""
from __future__ import annotations

import ctypes as C

import torch

from .. import _abi
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
        self.frontend = LogMelFrontend(dims.n_mels, self.device)
        self._pack_encoder(weights)
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
