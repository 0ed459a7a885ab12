"""Random-init Whisper weights of a named architecture (there are no checkpoints offline).

One flat dict of fp32 CPU tensors, OpenAI-Whisper shaped:
  enc.conv1.{w[d,n_mels,3],b}  enc.conv2.{w[d,d,3],b}  enc.pos[n_audio_ctx,d] (fixed sinusoids)
  enc.l{i}.attn_ln.{w,b}  enc.l{i}.attn.{q,k,v,o}.w[d,d]  (+ .b for q, v, o; the key projection has no bias)
  enc.l{i}.mlp_ln.{w,b}   enc.l{i}.mlp.fc1.{w[4d,d],b}  enc.l{i}.mlp.fc2.{w[d,4d],b}   enc.ln_post.{w,b}
  dec.tok_emb[V,d]  dec.pos[n_text_ctx,d]
  dec.l{i}.{attn_ln,cross_ln,mlp_ln}.{w,b}  dec.l{i}.{attn,cross}.{q,k,v,o}  dec.l{i}.mlp.{fc1,fc2}  dec.ln.{w,b}

The same dict feeds the fp32 CPU oracle (oracle/whisper_ref.py) and, packed to bf16, the CUDA path,
so both sides see bit-identical parameters (bf16-representable values: every tensor is rounded to
bf16 once at init so the oracle's fp32 copy and the kernels' bf16 copy hold the same numbers).
"""
from __future__ import annotations

import math

import torch

from .config import WhisperDims


def sinusoids(length: int, channels: int, max_timescale: float = 10000.0) -> torch.Tensor:
    inc = math.log(max_timescale) / (channels // 2 - 1)
    inv = torch.exp(-inc * torch.arange(channels // 2, dtype=torch.float32))
    t = torch.arange(length, dtype=torch.float32)[:, None] * inv[None, :]
    return torch.cat([torch.sin(t), torch.cos(t)], dim=1)


PROFILES = {
    # near-flat logits, varied tokens: exercises every code path; greedy margins are small, so token
    # parity is asserted only where the fp32 margin exceeds the stated bf16 logit tolerance
    "diverse": dict(gain=1.0, emb_row_spread=0.3, qk_gain=3.0, pos_std=1.0),
    # heavy-tailed tied embedding -> peaked distribution (like a trained model): top-1 margins are far
    # above bf16 noise, so free-running greedy tokens must be bit-exact
    "confident": dict(gain=0.7, emb_row_spread=0.3, qk_gain=3.0, pos_std=0.3),
}


def init_random_weights(dims: WhisperDims, seed: int = 0, gain: float = 0.7, emb_row_spread: float = 0.3,
                        qk_gain: float = 1.0, pos_std: float = 0.3, bf16_exact: bool = True,
                        profile: str | None = None) -> dict[str, torch.Tensor]:
    """Seeded random weights. `emb_row_spread` is the log-normal sigma of the per-token norm of the
    tied token embedding: it makes the logit distribution heavy-tailed (like a trained model's peaked
    distribution) so greedy top-1 margins sit far above bf16 noise."""
    if profile is not None:
        return init_random_weights(dims, seed, bf16_exact=bf16_exact, **PROFILES[profile])
    g = torch.Generator().manual_seed(seed)
    W: dict[str, torch.Tensor] = {}

    def randn(*shape, std=1.0):
        return torch.randn(*shape, generator=g, dtype=torch.float32) * std

    def linear(prefix, out_f, in_f, bias=True):
        W[prefix + ".w"] = randn(out_f, in_f, std=gain / math.sqrt(in_f))
        if bias:
            W[prefix + ".b"] = randn(out_f, std=0.1)

    def ln(prefix, n):
        W[prefix + ".w"] = 1.0 + randn(n, std=0.1)
        W[prefix + ".b"] = randn(n, std=0.1)

    def attn(prefix, n):
        linear(prefix + ".q", n, n)
        linear(prefix + ".k", n, n, bias=False)
        W[prefix + ".q.w"] *= qk_gain
        W[prefix + ".k.w"] *= qk_gain
        linear(prefix + ".v", n, n)
        linear(prefix + ".o", n, n)

    d = dims.n_audio_state
    W["enc.conv1.w"] = randn(d, dims.n_mels, 3, std=gain / math.sqrt(3 * dims.n_mels))
    W["enc.conv1.b"] = randn(d, std=0.1)
    W["enc.conv2.w"] = randn(d, d, 3, std=gain / math.sqrt(3 * d))
    W["enc.conv2.b"] = randn(d, std=0.1)
    W["enc.pos"] = sinusoids(dims.n_audio_ctx, d)
    for i in range(dims.n_audio_layer):
        p = f"enc.l{i}"
        ln(p + ".attn_ln", d)
        attn(p + ".attn", d)
        ln(p + ".mlp_ln", d)
        linear(p + ".mlp.fc1", 4 * d, d)
        linear(p + ".mlp.fc2", d, 4 * d)
    ln("enc.ln_post", d)

    t = dims.n_text_state
    row = torch.exp(randn(dims.n_vocab, std=emb_row_spread))
    W["dec.tok_emb"] = randn(dims.n_vocab, t, std=1.0 / math.sqrt(t)) * row[:, None]
    W["dec.pos"] = randn(dims.n_text_ctx, t, std=pos_std)
    for i in range(dims.n_text_layer):
        p = f"dec.l{i}"
        ln(p + ".attn_ln", t)
        attn(p + ".attn", t)
        ln(p + ".cross_ln", t)
        attn(p + ".cross", t)
        ln(p + ".mlp_ln", t)
        linear(p + ".mlp.fc1", 4 * t, t)
        linear(p + ".mlp.fc2", t, 4 * t)
    ln("dec.ln", t)
    if bf16_exact:
        for k in W:
            W[k] = W[k].to(torch.bfloat16).to(torch.float32)
    return W


# ==============================================================================================
# pretrained checkpoints -> the flat dict above
# ==============================================================================================
# The reference loads a pretrained model by name (/root/reference/backend/services/transcriber.py:23-27:
# `WhisperModel(model_size, device='cpu', compute_type='int8')`, which resolves the name to a CTranslate2
# conversion of the OpenAI checkpoint in the Hugging Face cache). `load_weights(path)` accepts the three
# public serialisations of those same parameters and returns (WhisperDims, dict) in the layout documented at the
# top of this module; WhisperEngine then packs it to bf16.
_HF_ATTN = (("q_proj", "q"), ("k_proj", "k"), ("v_proj", "v"), ("out_proj", "o"))
_OA_ATTN = (("query", "q"), ("key", "k"), ("value", "v"), ("out", "o"))


def _f32(t) -> torch.Tensor:
    return torch.as_tensor(t).detach().to(torch.float32).contiguous()


def dims_from_weights(W: dict, n_heads: int | None = None, name: str = "loaded") -> WhisperDims:
    """Architecture from tensor shapes. Every public Whisper size has 64-wide heads, which fixes the head count
    unless the caller (a config file) names it."""
    d_a, n_mels, _ = W["enc.conv1.w"].shape
    d_t = W["dec.tok_emb"].shape[1]
    import re
    n_al = 1 + max(int(m.group(1)) for m in (re.match(r"enc\.l(\d+)\.", k) for k in W) if m)
    n_tl = 1 + max(int(m.group(1)) for m in (re.match(r"dec\.l(\d+)\.", k) for k in W) if m)
    return WhisperDims(name, int(n_mels), int(W["enc.pos"].shape[0]), int(d_a), n_heads or d_a // 64, n_al,
                       int(W["dec.tok_emb"].shape[0]), int(W["dec.pos"].shape[0]), int(d_t), n_heads or d_t // 64, n_tl)


def from_hf_state_dict(sd: dict) -> dict:
    """Hugging Face `WhisperForConditionalGeneration.state_dict()` (keys `model.encoder...` / `model.decoder...`,
    `proj_out.weight` tied to the token embedding) -> flat dict."""
    sd = {(k[len("model."):] if k.startswith("model.") else k): v for k, v in sd.items()}
    W: dict[str, torch.Tensor] = {}

    def ln(p, hp):
        W[p + ".w"], W[p + ".b"] = _f32(sd[hp + ".weight"]), _f32(sd[hp + ".bias"])

    def attn(p, hp):
        for a, b in _HF_ATTN:
            W[f"{p}.{b}.w"] = _f32(sd[f"{hp}.{a}.weight"])
            if b != "k":
                W[f"{p}.{b}.b"] = _f32(sd[f"{hp}.{a}.bias"])

    def mlp(p, hp):
        for fc in ("fc1", "fc2"):
            W[f"{p}.mlp.{fc}.w"], W[f"{p}.mlp.{fc}.b"] = _f32(sd[f"{hp}.{fc}.weight"]), _f32(sd[f"{hp}.{fc}.bias"])

    for c in ("conv1", "conv2"):
        W[f"enc.{c}.w"], W[f"enc.{c}.b"] = _f32(sd[f"encoder.{c}.weight"]), _f32(sd[f"encoder.{c}.bias"])
    W["enc.pos"] = _f32(sd["encoder.embed_positions.weight"])
    i = 0
    while f"encoder.layers.{i}.fc1.weight" in sd:
        hp, p = f"encoder.layers.{i}", f"enc.l{i}"
        attn(p + ".attn", hp + ".self_attn"); ln(p + ".attn_ln", hp + ".self_attn_layer_norm")
        mlp(p, hp); ln(p + ".mlp_ln", hp + ".final_layer_norm")
        i += 1
    ln("enc.ln_post", "encoder.layer_norm")
    W["dec.tok_emb"] = _f32(sd["decoder.embed_tokens.weight"])
    W["dec.pos"] = _f32(sd["decoder.embed_positions.weight"])
    i = 0
    while f"decoder.layers.{i}.fc1.weight" in sd:
        hp, p = f"decoder.layers.{i}", f"dec.l{i}"
        attn(p + ".attn", hp + ".self_attn"); ln(p + ".attn_ln", hp + ".self_attn_layer_norm")
        attn(p + ".cross", hp + ".encoder_attn"); ln(p + ".cross_ln", hp + ".encoder_attn_layer_norm")
        mlp(p, hp); ln(p + ".mlp_ln", hp + ".final_layer_norm")
        i += 1
    ln("dec.ln", "decoder.layer_norm")
    return W


def from_openai_state_dict(sd: dict) -> dict:
    """OpenAI `whisper.load_model(...).state_dict()` (the `model_state_dict` of a `<size>.pt` checkpoint) -> flat dict."""
    W: dict[str, torch.Tensor] = {}

    def ln(p, op):
        W[p + ".w"], W[p + ".b"] = _f32(sd[op + ".weight"]), _f32(sd[op + ".bias"])

    def attn(p, op):
        for a, b in _OA_ATTN:
            W[f"{p}.{b}.w"] = _f32(sd[f"{op}.{a}.weight"])
            if b != "k":
                W[f"{p}.{b}.b"] = _f32(sd[f"{op}.{a}.bias"])

    def mlp(p, op):
        W[p + ".mlp.fc1.w"], W[p + ".mlp.fc1.b"] = _f32(sd[op + ".mlp.0.weight"]), _f32(sd[op + ".mlp.0.bias"])
        W[p + ".mlp.fc2.w"], W[p + ".mlp.fc2.b"] = _f32(sd[op + ".mlp.2.weight"]), _f32(sd[op + ".mlp.2.bias"])

    for c in ("conv1", "conv2"):
        W[f"enc.{c}.w"], W[f"enc.{c}.b"] = _f32(sd[f"encoder.{c}.weight"]), _f32(sd[f"encoder.{c}.bias"])
    W["enc.pos"] = _f32(sd["encoder.positional_embedding"])
    i = 0
    while f"encoder.blocks.{i}.mlp.0.weight" in sd:
        op, p = f"encoder.blocks.{i}", f"enc.l{i}"
        attn(p + ".attn", op + ".attn"); ln(p + ".attn_ln", op + ".attn_ln")
        mlp(p, op); ln(p + ".mlp_ln", op + ".mlp_ln")
        i += 1
    ln("enc.ln_post", "encoder.ln_post")
    W["dec.tok_emb"] = _f32(sd["decoder.token_embedding.weight"])
    W["dec.pos"] = _f32(sd["decoder.positional_embedding"])
    i = 0
    while f"decoder.blocks.{i}.mlp.0.weight" in sd:
        op, p = f"decoder.blocks.{i}", f"dec.l{i}"
        attn(p + ".attn", op + ".attn"); ln(p + ".attn_ln", op + ".attn_ln")
        attn(p + ".cross", op + ".cross_attn"); ln(p + ".cross_ln", op + ".cross_attn_ln")
        mlp(p, op); ln(p + ".mlp_ln", op + ".mlp_ln")
        i += 1
    ln("dec.ln", "decoder.ln")
    return W


# ---- CTranslate2 model.bin (what faster-whisper itself loads) -------------------------------
# Binary layout and variable names as published by CTranslate2 (`Model::load` / `WhisperSpec`), RECALLED -- the
# package is not installable here, so this reader is checked only against the writer below (tests) and is the
# first thing to re-verify when a real `model.bin` is at hand: u32 version; (v >= 2) string spec name;
# (v >= 3) u32 spec revision; u32 n_variables; per variable: string name, u8 rank, u32 dims[rank],
# (v >= 4) u8 dtype + u32 n_bytes else u8 item_size + u32 n_items, raw data; (v >= 3) u32 n_aliases + string pairs.
# Strings are u16 length (including the trailing NUL) + bytes. dtype ids: 0 f32, 1 i8, 2 i16, 3 i32, 4 f16, 5 bf16.
_CT2_DTYPES = {0: torch.float32, 1: torch.int8, 2: torch.int16, 3: torch.int32, 4: torch.float16, 5: torch.bfloat16}


def read_ct2_variables(path: str) -> dict:
    import struct
    import numpy as np
    out, aliases = {}, []
    with open(path, "rb") as f:
        def u(fmt):
            return struct.unpack("<" + fmt, f.read(struct.calcsize("<" + fmt)))[0]

        def string():
            n = u("H")
            return f.read(n).rstrip(b"\0").decode()
        version = u("I")
        if version >= 2:
            string()
        if version >= 3:
            u("I")
        for _ in range(u("I")):
            name = string()
            shape = [u("I") for _ in range(u("B"))]
            if version >= 4:
                dtype, nbytes = _CT2_DTYPES[u("B")], u("I")
            else:
                item, n_items = u("B"), u("I")
                dtype, nbytes = {4: torch.float32, 2: torch.int16, 1: torch.int8}[item], item * n_items
            raw = np.frombuffer(f.read(nbytes), dtype=np.uint8).copy()
            out[name] = torch.from_numpy(raw).view(dtype).reshape(shape)
        if version >= 3:
            for _ in range(u("I")):
                aliases.append((string(), string()))
    for alias, target in aliases:
        out[alias] = out[target]
    return out


def write_ct2_variables(path: str, variables: dict, aliases=(), spec: str = "WhisperSpec", version: int = 6) -> None:
    """Inverse of read_ct2_variables (test helper; also documents the layout)."""
    import struct
    ids = {v: k for k, v in _CT2_DTYPES.items()}
    with open(path, "wb") as f:
        def string(s):
            b = s.encode() + b"\0"
            f.write(struct.pack("<H", len(b)) + b)
        f.write(struct.pack("<I", version)); string(spec); f.write(struct.pack("<I", 3))
        f.write(struct.pack("<I", len(variables)))
        for name, t in variables.items():
            t = t.contiguous()
            string(name)
            f.write(struct.pack("<B", t.dim()) + b"".join(struct.pack("<I", s) for s in t.shape))
            raw = t.view(torch.uint8).numpy().tobytes() if t.numel() else b""
            f.write(struct.pack("<BI", ids[t.dtype], len(raw)) + raw)
        f.write(struct.pack("<I", len(aliases)))
        for a, b in aliases:
            string(a); string(b)


def from_ct2_variables(V: dict) -> dict:
    """CTranslate2 WhisperSpec variables -> flat dict. Self-attention q/k/v are fused in `linear_0` ([3d, d]),
    cross-attention k/v in `attention/linear_1` ([2d, d]); int8 / int16 weights carry a per-row `weight_scale`
    (w = q / scale); the fused bias keeps a zero block for the bias-free key projection."""
    def lin(prefix):
        w = V[prefix + "/weight"]
        if w.dtype in (torch.int8, torch.int16):
            w = w.to(torch.float32) / _f32(V[prefix + "/weight_scale"]).reshape(-1, 1)
        b = V.get(prefix + "/bias")
        return _f32(w), (None if b is None else _f32(b))

    W: dict[str, torch.Tensor] = {}

    def ln(p, cp):
        W[p + ".w"], W[p + ".b"] = _f32(V[cp + "/gamma"]), _f32(V[cp + "/beta"])

    def put(p, w, b, bias=True):
        W[p + ".w"] = w.contiguous()
        if bias:
            W[p + ".b"] = (b if b is not None else torch.zeros(w.shape[0])).contiguous()

    def self_attn(p, cp):
        ln(p + "_ln", cp + "/layer_norm")
        w, b = lin(cp + "/linear_0")
        d = w.shape[1]
        b = b if b is not None else torch.zeros(3 * d)
        put(p + ".q", w[:d], b[:d]); put(p + ".k", w[d:2 * d], None, bias=False); put(p + ".v", w[2 * d:], b[2 * d:])
        put(p + ".o", *lin(cp + "/linear_1"))

    def ffn(p, cp):
        ln(p + ".mlp_ln", cp + "/layer_norm")
        put(p + ".mlp.fc1", *lin(cp + "/linear_0")); put(p + ".mlp.fc2", *lin(cp + "/linear_1"))

    for c in ("conv1", "conv2"):
        W[f"enc.{c}.w"], W[f"enc.{c}.b"] = _f32(V[f"encoder/{c}/weight"]), _f32(V[f"encoder/{c}/bias"])
    W["enc.pos"] = _f32(V["encoder/position_encodings/encodings"])
    i = 0
    while f"encoder/layer_{i}/ffn/linear_0/weight" in V:
        cp, p = f"encoder/layer_{i}", f"enc.l{i}"
        self_attn(p + ".attn", cp + "/self_attention"); ffn(p, cp + "/ffn")
        i += 1
    ln("enc.ln_post", "encoder/layer_norm")
    W["dec.tok_emb"] = _f32(V["decoder/embeddings/weight"])
    W["dec.pos"] = _f32(V["decoder/position_encodings/encodings"])
    i = 0
    while f"decoder/layer_{i}/ffn/linear_0/weight" in V:
        cp, p = f"decoder/layer_{i}", f"dec.l{i}"
        self_attn(p + ".attn", cp + "/self_attention")
        ln(p + ".cross_ln", cp + "/attention/layer_norm")
        put(p + ".cross.q", *lin(cp + "/attention/linear_0"))
        w, b = lin(cp + "/attention/linear_1")
        d = w.shape[1]
        b = b if b is not None else torch.zeros(2 * d)
        put(p + ".cross.k", w[:d], None, bias=False); put(p + ".cross.v", w[d:], b[d:])
        put(p + ".cross.o", *lin(cp + "/attention/linear_2"))
        ffn(p, cp + "/ffn")
        i += 1
    ln("dec.ln", "decoder/layer_norm")
    return W


def load_weights(path: str, name: str | None = None):
    """Pretrained Whisper parameters from `path` -> (WhisperDims, flat fp32 dict).

    Accepted: a Hugging Face model directory or file (`model.safetensors` / `pytorch_model.bin`, keys `model.encoder...`),
    an OpenAI checkpoint (`<size>.pt` with `dims` + `model_state_dict`), a CTranslate2 / faster-whisper model directory
    or its `model.bin`. A `config.json` next to the file supplies the head counts when present."""
    import json
    import os
    cfg, file = {}, path
    if os.path.isdir(path):
        for cand in ("model.safetensors", "pytorch_model.bin", "model.bin"):
            if os.path.exists(os.path.join(path, cand)):
                file = os.path.join(path, cand)
                break
        else:
            cands = [f for f in os.listdir(path) if f.endswith((".pt", ".safetensors"))]
            if len(cands) != 1:
                raise FileNotFoundError(f"no Whisper checkpoint found in {path}")
            file = os.path.join(path, cands[0])
    cfg_path = os.path.join(os.path.dirname(file), "config.json")
    if os.path.exists(cfg_path):
        with open(cfg_path) as f:
            cfg = json.load(f)
    heads = cfg.get("encoder_attention_heads")
    label = name or os.path.basename(os.path.normpath(path))
    if file.endswith(".safetensors"):
        from safetensors.torch import load_file
        W = from_hf_state_dict(load_file(file))
    elif os.path.basename(file) == "model.bin":
        W = from_ct2_variables(read_ct2_variables(file))
    else:
        ck = torch.load(file, map_location="cpu", weights_only=True)
        if "model_state_dict" in ck:
            heads = heads or ck.get("dims", {}).get("n_audio_head")
            W = from_openai_state_dict(ck["model_state_dict"])
        elif any(k.startswith(("model.encoder.", "encoder.layers.")) for k in ck):
            W = from_hf_state_dict(ck)
        else:
            W = from_openai_state_dict(ck)
    return dims_from_weights(W, heads, label), W
