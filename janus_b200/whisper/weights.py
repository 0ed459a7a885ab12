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
